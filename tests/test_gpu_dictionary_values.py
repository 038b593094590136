"""DICTIONARY32 value columns on the hash path: the element of row r is keys[indices[r]], read
through the indirection by the generic kernels, and every aggregation is typed by the dictionary's KEYS
(/root/reference/cpp/src/groupby/hash/groupby.cu:118-120, hash/output_utils.cu:66, device_aggregators.cuh:164-200). The expected
values come from the oracle run on the DECODED column."""
import numpy as np
import pytest

from oracle import oracle as orc

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():  # pragma: no cover
    pytest.skip("needs a CUDA device", allow_module_level=True)

import cudf_b200 as cb  # noqa: E402
from cudf_b200 import _lib  # noqa: E402
from tests import gpu_util  # noqa: E402

H = orc.HostColumn
N = 200_000


def _run(keys_np, dict_keys, indices, valid, kinds, include=False):
    """groupby(keys).agg(kinds) of the dictionary column (dict_keys, indices, valid) on the GPU vs the oracle on dict_keys[indices]"""
    dk = [cb.Column.from_numpy(k) for k in keys_np]
    dcol = cb.Column.dictionary(cb.Column.from_numpy(indices), cb.Column.from_numpy(dict_keys),
                                *( (torch.from_numpy(cb.column.pack_mask(valid)).cuda(), int((~valid).sum())) if valid is not None else (None, 0)))
    gb = cb.GroupBy(cb.Table(dk), cb.NullPolicy.INCLUDE if include else cb.NullPolicy.EXCLUDE)
    kt, rts = gb.aggregate([cb.GroupByRequest(dcol, [gpu_util.make_agg(k, d) for k, d in kinds])])
    assert gb.last_path == "global", gb.last_path
    kout = [c.to_numpy() for c in kt.columns()]
    order = gpu_util.sort_order(kout)
    got = [tuple(x[order] for x in c.to_numpy()) for c in rts[0].columns()]
    decoded = H(dict_keys[indices.astype(np.int64)], valid)
    Go, ko, ao = orc.groupby([H(k) for k in keys_np], [decoded], [(0, k, d) for k, d in kinds], include_null_keys=include, threads=4)
    assert len(order) == Go
    for c in range(len(keys_np)):
        np.testing.assert_array_equal(kout[c][0][order], ko[c][0])
    for i, (kind, _) in enumerate(kinds):
        np.testing.assert_array_equal(got[i][1], ao[i][1], err_msg=f"{kind} validity")
        a, b = got[i][0][ao[i][1]], ao[i][0][ao[i][1]]
        assert a.dtype == b.dtype, (kind, a.dtype, b.dtype)
        if a.dtype.kind == "f":
            np.testing.assert_allclose(a, b, rtol=1e-9, atol=0, err_msg=kind)
        else:
            np.testing.assert_array_equal(a, b, err_msg=kind)


@pytest.mark.parametrize("idx_dtype", [np.uint8, np.uint16, np.uint32, np.int32])
def test_int64_dictionary_every_kind(idx_dtype):
    rng = np.random.default_rng(1)
    D = 200
    dict_keys = np.sort(rng.integers(-(1 << 40), 1 << 40, D)).astype(np.int64)
    indices = rng.integers(0, D, N).astype(idx_dtype)
    valid = rng.random(N) > 0.15
    k = rng.integers(0, 3000, N).astype(np.int64)
    _run([k], dict_keys, indices, valid,
         [("sum", 1), ("min", 1), ("max", 1), ("mean", 1), ("count_valid", 1), ("count_all", 1), ("variance", 1), ("argmin", 1), ("argmax", 1)])


def test_float64_dictionary_without_nulls_and_two_keys():
    rng = np.random.default_rng(2)
    dict_keys = np.unique(rng.integers(0, 100_000, 5000)).astype(np.float64) / 8.0
    indices = rng.integers(0, len(dict_keys), N).astype(np.uint32)
    _run([rng.integers(0, 40, N).astype(np.int32), rng.integers(0, 7, N).astype(np.int8)], dict_keys, indices, None,
         [("sum", 1), ("min", 1), ("max", 1), ("mean", 1), ("std", 0)])


def test_int32_dictionary_sum_promotes_to_int64():
    rng = np.random.default_rng(3)
    dict_keys = np.array([2_000_000_000, -2_000_000_000, 7, 0], np.int32)
    indices = rng.integers(0, 4, N).astype(np.uint8)
    _run([rng.integers(0, 5, N).astype(np.int64)], dict_keys, indices, rng.random(N) > 0.5, [("sum", 1), ("max", 1), ("count_valid", 1)], include=True)


def test_capability_and_errors():
    L = _lib.lib()
    k = cb.Column.from_numpy(np.arange(10, dtype=np.int64))
    d = cb.Column.dictionary(cb.Column.from_numpy(np.zeros(10, np.uint8)), cb.Column.from_numpy(np.array([1.5])))
    gb = cb.GroupBy(cb.Table([k]), options={"force_path": 7})
    with pytest.raises(cb.UnsupportedError):
        gb.aggregate([cb.GroupByRequest(d, [gpu_util.make_agg("sum")])])
    # a dictionary of dictionaries is not a thing (device_aggregators.cuh:166-167)
    dd = cb.Column.dictionary(cb.Column.from_numpy(np.zeros(10, np.uint8)), d)
    with pytest.raises((cb.UnsupportedError, _lib.LogicError)):  # no target type for (DICTIONARY32, SUM): the reference's "Invalid type/aggregation combination."
        cb.GroupBy(cb.Table([k])).aggregate([cb.GroupByRequest(dd, [gpu_util.make_agg("sum")])])
    assert L is not None
