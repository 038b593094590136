"""The reference's own known-answer tests for DICTIONARY32 value columns with fixed-width keys, through the CUDA path:
sum_tests.cpp:146-164 (typed), max_tests.cpp:226-247, min_tests.cpp:219-240, mean_tests.cpp:130-146. A
`dictionary_column_wrapper<V>{...}` holds the sorted distinct values as its keys child and uint32 positions as its indices."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():  # pragma: no cover
    pytest.skip("needs a CUDA device", allow_module_level=True)

import cudf_b200 as cb  # noqa: E402
from tests import gpu_util  # noqa: E402

K10 = np.array([1, 2, 3, 1, 2, 2, 1, 3, 3, 2], np.int32)


def _dictionary(values, dtype):
    v = np.asarray(values).astype(dtype)
    keys, idx = np.unique(v, return_inverse=True)
    return cb.Column.dictionary(cb.Column.from_numpy(idx.astype(np.uint32)), cb.Column.from_numpy(keys))


def _single_agg(vals, kind):
    gb = cb.GroupBy(cb.Table([cb.Column.from_numpy(K10)]))
    kt, rts = gb.aggregate([cb.GroupByRequest(vals, [gpu_util.make_agg(kind)])])
    k = kt.columns()[0].to_numpy()[0]
    order = np.argsort(k)
    d, valid = rts[0].columns()[0].to_numpy()
    assert valid.all()
    np.testing.assert_array_equal(k[order], [1, 2, 3])
    return d[order]


@pytest.mark.parametrize("dtype", [np.int8, np.int16, np.int32, np.int64, np.float32, np.float64])
def test_sum_dictionary(dtype):
    got = _single_agg(_dictionary(range(10), dtype), "sum")
    assert got.dtype == (np.int64 if np.dtype(dtype).kind == "i" else np.dtype(dtype))
    np.testing.assert_array_equal(got, np.array([9, 19, 17]).astype(got.dtype))


def test_max_min_fixed_width_dictionary():
    vals = [0xABC, 0xBBB, 0xF1, 0xAAA, 0xFFF, 0xBAA, 0xAAA, 0x01, 0xF1, 0xEEE]
    got = _single_agg(_dictionary(vals, np.int64), "max")
    assert got.dtype == np.int64
    np.testing.assert_array_equal(got, [0xABC, 0xFFF, 0xF1])
    got = _single_agg(_dictionary(vals, np.int64), "min")
    np.testing.assert_array_equal(got, [0xAAA, 0xBAA, 0x01])


def test_mean_int16_dictionary():
    got = _single_agg(_dictionary(range(10), np.int16), "mean")
    assert got.dtype == np.float64
    np.testing.assert_allclose(got, [9.0 / 3, 19.0 / 4, 17.0 / 3], rtol=1e-15, atol=0)
