"""Helpers for the GPU parity tests: run the CUDA path through the C ABI and compare key-sorted with the oracle.

Comparison is key-sorted because group order is unspecified in the reference too
(/root/reference/cpp/tests/groupby/groupby_test_util.cpp:86-94 sorts the result before comparing).
"""
import threading

import numpy as np

import cudf_b200 as cb
from cudf_b200 import aggregation as agg_mod
from oracle import oracle as orc

KIND_TO_AGG = {
    "sum": agg_mod.sum, "product": agg_mod.product, "min": agg_mod.min, "max": agg_mod.max,
    "count_valid": lambda: agg_mod.count(cb.NullPolicy.EXCLUDE), "count_all": lambda: agg_mod.count(cb.NullPolicy.INCLUDE),
    "sum_of_squares": agg_mod.sum_of_squares, "mean": agg_mod.mean, "m2": agg_mod.m2, "argmin": agg_mod.argmin,
    "argmax": agg_mod.argmax, "sum_overflow": agg_mod.sum_overflow,
}


def make_agg(kind, ddof=1):
    if kind == "variance":
        return agg_mod.variance(ddof)
    if kind == "std":
        return agg_mod.std(ddof)
    return KIND_TO_AGG[kind]()


def to_device(hc: orc.HostColumn) -> cb.Column:
    return cb.Column.from_numpy(hc.data, hc.valid, type_id=hc.type_id)


def sort_order(key_cols):
    """key_cols: list of (data, valid). nulls first, ascending, NaN last -- same order as the oracle's output."""
    ks = []
    for data, valid in reversed(key_cols):
        d = data.copy()
        if d.dtype.kind == "f":
            d[~valid] = 0
        else:
            d = d.astype(np.int64) if d.dtype.kind in "ib" else d
            d[~valid] = 0
        ks.append(d)
        ks.append(valid.astype(np.int8))
    return np.lexsort(ks) if ks else np.arange(0)


def run_gpu(keys, values, aggs, include_null_keys=False, options=None):
    """keys/values: lists of orc.HostColumn; aggs: [(value_index, kind, ddof)].
    Returns (G, [(data, valid)] keys, [(data, valid)] aggs) key-sorted, plus the path string."""
    dkeys = [to_device(k) for k in keys]
    dvals = [to_device(v) for v in values]
    # one request per value column, aggregations in first-seen order
    per_val = {}
    for pos, (vi, kind, ddof) in enumerate(aggs):
        per_val.setdefault(vi, []).append((pos, kind, ddof))
    reqs, where = [], {}
    for vi, lst in per_val.items():
        reqs.append(cb.GroupByRequest(dvals[vi], [make_agg(k, d) for _, k, d in lst]))
        for j, (pos, _, _) in enumerate(lst):
            where[pos] = (len(reqs) - 1, j)
    gb = cb.GroupBy(cb.Table(dkeys), cb.NullPolicy.INCLUDE if include_null_keys else cb.NullPolicy.EXCLUDE, options=options)
    ktab, rtabs = gb.aggregate(reqs)
    kout = [c.to_numpy() for c in ktab.columns()]
    G = ktab.columns()[0].size
    order = sort_order(kout)
    kout = [(d[order], v[order]) for d, v in kout]
    aout = []
    for pos in range(len(aggs)):
        r, j = where[pos]
        d, v = rtabs[r].columns()[j].to_numpy()
        aout.append((d[order], v[order]))
    return G, kout, aout, gb.last_path


_ORACLE_LOCK = threading.Lock()  # the C oracle keeps its last result in static storage: one call at a time


def compare_with_oracle(keys, values, aggs, include_null_keys=False, options=None, float_rtol=1e-9, name=""):
    with _ORACLE_LOCK:
        Go, ko, ao = orc.groupby(keys, values, aggs, include_null_keys=include_null_keys, threads=4)
    G, kg, ag, path = run_gpu(keys, values, aggs, include_null_keys, options)
    assert G == Go, f"{name}: {G} groups on GPU, oracle has {Go} (path {path})"
    for c in range(len(keys)):
        np.testing.assert_array_equal(kg[c][1], ko[c][1], err_msg=f"{name}: key {c} validity")
        v = ko[c][1]
        a, b = kg[c][0][v], ko[c][0][v]
        if a.dtype.kind == "f":
            np.testing.assert_array_equal(np.where(np.isnan(a), 0, a), np.where(np.isnan(b), 0, b), err_msg=f"{name}: key {c}")
            np.testing.assert_array_equal(np.isnan(a), np.isnan(b))
        else:
            np.testing.assert_array_equal(a, b, err_msg=f"{name}: key {c}")
    for i, (vi, kind, ddof) in enumerate(aggs):
        np.testing.assert_array_equal(ag[i][1], ao[i][1], err_msg=f"{name}: agg {i} ({kind}) validity")
        v = ao[i][1]
        a, b = ag[i][0][v], ao[i][0][v]
        assert a.dtype == b.dtype, f"{name}: agg {i} ({kind}) dtype {a.dtype} vs oracle {b.dtype}"
        if a.dtype.kind == "f":
            rt = float_rtol if a.dtype == np.float64 and values[vi].data.dtype != np.float32 else 2e-5
            np.testing.assert_allclose(a, b, rtol=rt, atol=0, equal_nan=True, err_msg=f"{name}: agg {i} ({kind})")
        else:
            np.testing.assert_array_equal(a, b, err_msg=f"{name}: agg {i} ({kind}) must be bit-exact")
    return G, path
