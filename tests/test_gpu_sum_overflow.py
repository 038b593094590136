"""GPU parity of SUM_OVERFLOW (struct{sum, overflow}; /root/reference/cpp/tests/groupby/sum_overflow_tests.cpp) through the C
ABI: the reference's own vectors, then seeded inputs against the oracle on every kernel family that serves the two exact-total
accumulators (generic, partitioned, bucketed). Flags and validity bit-exact; sums bit-exact wherever the flag is clear (the
sum of an overflowed group is unspecified in the reference -- here both sides hold the wrapped total, compared too)."""
import numpy as np
import pytest

from oracle import oracle as orc

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():  # pragma: no cover
    pytest.skip("needs a CUDA device", allow_module_level=True)

import cudf_b200 as cb  # noqa: E402
from tests import gpu_util  # noqa: E402

H = orc.HostColumn
TYPES = [(np.int8, None), (np.int16, None), (np.int32, None), (np.int64, None), (np.int32, 25), (np.int64, 26)]
IDS = ["int8", "int16", "int32", "int64", "decimal32", "decimal64"]


def run(keys, vals, include=False, options=None, extra=()):
    """-> G, key data (sorted), sum child, overflow child, struct validity, [extra results], path"""
    gb = cb.GroupBy(cb.Table([gpu_util.to_device(k) for k in keys]), cb.NullPolicy.INCLUDE if include else cb.NullPolicy.EXCLUDE, options=options)
    aggs = [gpu_util.make_agg("sum_overflow")] + [gpu_util.make_agg(k) for k in extra]
    ktab, rtabs = gb.aggregate([cb.GroupByRequest(gpu_util.to_device(vals), aggs)])
    kout = [c.to_numpy() for c in ktab.columns()]
    order = gpu_util.sort_order(kout)
    col = rtabs[0].columns()[0]
    assert col.type_id == cb.TypeId.STRUCT and len(col.children) == 2
    sums, valid = cb.Column(col.children[0].data, col.mask, col.null_count, col.children[0].type_id).to_numpy()
    flags, _ = col.children[1].to_numpy()
    if col.mask is None:
        valid = np.ones(len(sums), dtype=bool)
    others = [tuple(x[order] for x in c.to_numpy()) for c in rtabs[0].columns()[1:]]
    return len(sums), [k[0][order] for k in kout], sums[order], flags[order].astype(bool), valid[order], others, gb.last_path


def check_against_oracle(keys, vals, include=False, options=None, name=""):
    Go, ko, ao = orc.groupby(keys, [vals], [(0, "sum_overflow", 1), (0, "sum_overflow_flag", 1), (0, "sum", 1), (0, "count_valid", 1)],
                             include_null_keys=include, threads=4)
    G, k, s, f, v, others, path = run(keys, vals, include, options, extra=("sum", "count_valid"))
    assert G == Go, f"{name}: {G} groups, oracle {Go} ({path})"
    np.testing.assert_array_equal(v, ao[0][1] if ao[0][2] else np.ones(G, dtype=bool), err_msg=f"{name}: struct validity")
    np.testing.assert_array_equal(f[v], ao[1][0][v], err_msg=f"{name}: overflow flags ({path})")
    assert s.dtype == ao[0][0].dtype
    np.testing.assert_array_equal(s[v], ao[0][0][v], err_msg=f"{name}: sums ({path})")
    np.testing.assert_array_equal(others[0][0][v], ao[2][0][v], err_msg=f"{name}: plain SUM beside it")
    np.testing.assert_array_equal(others[1][0], ao[3][0], err_msg=f"{name}: COUNT beside it")
    return G, path


@pytest.mark.parametrize("dt,tid", TYPES, ids=IDS)
def test_reference_vectors(dt, tid):
    # Basic
    G, k, s, f, v, _, _ = run([H(np.array([1, 2, 3, 1, 2, 2, 1, 3, 3, 2], dtype=np.int32))], H(np.arange(10).astype(dt), type_id=tid))
    assert G == 3 and s.dtype == np.dtype(dt)
    np.testing.assert_array_equal(s, np.array([9, 19, 17], dtype=dt)); assert not f.any() and v.all()
    # ZeroValidValues: null struct, children 0 / false
    G, k, s, f, v, _, _ = run([H(np.array([1, 1, 1], dtype=np.int32))], H(np.array([3, 4, 5], dtype=dt), np.zeros(3, dtype=bool), type_id=tid))
    assert G == 1 and s[0] == 0 and not f[0] and not v[0]
    # NullKeysAndValues
    keys = H(np.array([1, 2, 3, 1, 2, 2, 1, 3, 3, 2, 4], dtype=np.int32), np.array([1, 1, 1, 1, 1, 1, 1, 0, 1, 1, 1], dtype=bool))
    vals = H(np.array([0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 4], dtype=dt), np.array([0, 1, 1, 1, 1, 0, 1, 1, 1, 1, 0], dtype=bool), type_id=tid)
    G, k, s, f, v, _, _ = run([keys], vals)
    np.testing.assert_array_equal(k[0], [1, 2, 3, 4])
    np.testing.assert_array_equal(s, np.array([9, 14, 10, 0], dtype=dt)); assert not f.any()
    np.testing.assert_array_equal(v, [True, True, True, False])
    # OverflowDetection: the reference checks the flags
    mx, mn = np.iinfo(dt).max, np.iinfo(dt).min
    keys = H(np.array([1, 2, 3, 4, 1, 2, 2, 1, 3, 3, 2, 4, 4], dtype=np.int32))
    vals = H(np.array([mx - 5, 10, 20, mn + 5, 10, 20, 30, mx - 5, mx - 5, 1, 40, -10, mn + 5], dtype=dt), type_id=tid)
    G, k, s, f, v, _, _ = run([keys], vals)
    np.testing.assert_array_equal(f, [True, False, True, True]); assert s[1] == 100


@pytest.mark.parametrize("dt,tid", TYPES, ids=IDS)
@pytest.mark.parametrize("include", [False, True])
def test_seeded_against_oracle(dt, tid, include):
    rng = np.random.default_rng(31)
    n = 200_000
    k = H(rng.integers(0, 3000, n).astype(np.int32), rng.random(n) > 0.03)
    info = np.iinfo(dt)
    v = rng.integers(info.min // 50, info.max // 50, n).astype(dt)     # ~67 rows per group: some totals leave the type, some do not
    small = rng.random(n) < 0.5
    v[small] = rng.integers(-3, 4, int(small.sum())).astype(dt)
    vals = H(v, rng.random(n) > 0.1, type_id=tid)
    G, path = check_against_oracle([k], vals, include=include, name=f"seeded/{np.dtype(dt).name}/{include}")
    assert path == "global", path


def test_int64_extremes_need_the_full_128_bit_total():
    # many INT64_MAX / INT64_MIN values per group: the wrapped 64-bit sum alone cannot tell whether the total fits
    rng = np.random.default_rng(32)
    n = 300_000
    k = H(rng.integers(0, 500, n).astype(np.int64))
    choice = rng.integers(0, 4, n)
    v = np.where(choice == 0, np.iinfo(np.int64).max, np.where(choice == 1, np.iinfo(np.int64).min, rng.integers(-5, 5, n))).astype(np.int64)
    half = rng.integers(0, 500, 250)
    bal = np.isin(k.data, half)
    v[bal] = np.where(np.arange(n)[bal] % 2 == 0, np.iinfo(np.int64).max, -np.iinfo(np.int64).max)   # totals near zero: no overflow
    check_against_oracle([k], H(v), name="int64 extremes")


@pytest.mark.parametrize("force,expect", [(3, "partitioned"), (7, "bucketed")])
def test_other_kernel_families(force, expect):
    rng = np.random.default_rng(33)
    n = 1_000_000
    k = H(rng.integers(0, 200_000, n).astype(np.int64))
    v = rng.integers(-(1 << 62), 1 << 62, n).astype(np.int64)
    G, path = check_against_oracle([k], H(v, rng.random(n) > 0.2), options={"force_path": force, "bucket_slice_log2": 8}, name=f"force {force}")
    assert path == expect, path


def test_unsupported_types_and_streaming():
    k = gpu_util.to_device(H(np.array([1, 1], dtype=np.int32)))
    for arr in (np.array([1, 2], dtype=np.uint32), np.array([True, False]), np.array([1.0, 2.0])):
        with pytest.raises((cb.LogicError, cb.UnsupportedError)):
            cb.GroupBy(cb.Table([k])).aggregate([cb.GroupByRequest(gpu_util.to_device(H(arr)), [gpu_util.make_agg("sum_overflow")])])
