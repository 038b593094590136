"""Pins the oracle's SUM_OVERFLOW (struct{sum, overflow}) against the reference's own known-answer tests
(/root/reference/cpp/tests/groupby/sum_overflow_tests.cpp: Basic :41-90, ZeroValidValues :190-220, NullKeysAndValues :222-258,
OverflowDetection :260-378). The vectors below are the ones those tests spell out; decimal32/64 share the integer arithmetic."""
import numpy as np
import pytest

from oracle import oracle as orc

H = orc.HostColumn
TYPES = [(np.int8, None), (np.int16, None), (np.int32, None), (np.int64, None), (np.int32, 25), (np.int64, 26)]  # 25/26: DECIMAL32/64
IDS = ["int8", "int16", "int32", "int64", "decimal32", "decimal64"]


def _run(keys, vals, include=False, threads=1):
    G, kout, aout = orc.groupby([keys], [vals], [(0, "sum_overflow", 1), (0, "sum_overflow_flag", 1)], include_null_keys=include, threads=threads)
    return G, kout[0], aout[0], aout[1]


@pytest.mark.parametrize("dt,tid", TYPES, ids=IDS)
def test_basic(dt, tid):
    keys = H(np.array([1, 2, 3, 1, 2, 2, 1, 3, 3, 2], dtype=np.int32))
    vals = H(np.arange(10).astype(dt), type_id=tid)
    G, k, s, f = _run(keys, vals)
    assert G == 3
    np.testing.assert_array_equal(k[0], [1, 2, 3])
    assert s[0].dtype == np.dtype(dt)  # the sum child keeps the Source type
    np.testing.assert_array_equal(s[0], np.array([9, 19, 17], dtype=dt))
    np.testing.assert_array_equal(f[0], [False, False, False])
    assert s[1].all() and not s[2]  # no nulls in the input: no struct mask


@pytest.mark.parametrize("dt,tid", TYPES, ids=IDS)
def test_zero_valid_values(dt, tid):
    keys = H(np.array([1, 1, 1], dtype=np.int32))
    vals = H(np.array([3, 4, 5], dtype=dt), np.zeros(3, dtype=bool), type_id=tid)
    G, k, s, f = _run(keys, vals)
    assert G == 1 and s[0][0] == 0 and not f[0][0]
    assert not s[1][0] and not f[1][0] and s[2]  # null struct


@pytest.mark.parametrize("dt,tid", TYPES, ids=IDS)
def test_null_keys_and_values(dt, tid):
    keys = H(np.array([1, 2, 3, 1, 2, 2, 1, 3, 3, 2, 4], dtype=np.int32), np.array([1, 1, 1, 1, 1, 1, 1, 0, 1, 1, 1], dtype=bool))
    vals = H(np.array([0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 4], dtype=dt), np.array([0, 1, 1, 1, 1, 0, 1, 1, 1, 1, 0], dtype=bool), type_id=tid)
    G, k, s, f = _run(keys, vals)
    assert G == 4
    np.testing.assert_array_equal(k[0], [1, 2, 3, 4])
    np.testing.assert_array_equal(s[0], np.array([9, 14, 10, 0], dtype=dt))
    np.testing.assert_array_equal(f[0], [False] * 4)
    np.testing.assert_array_equal(s[1], [True, True, True, False])


@pytest.mark.parametrize("threads", [1, 4])
@pytest.mark.parametrize("dt,tid", TYPES, ids=IDS)
def test_overflow_detection(dt, tid, threads):
    mx, mn = np.iinfo(dt).max, np.iinfo(dt).min
    keys = H(np.array([1, 2, 3, 4, 1, 2, 2, 1, 3, 3, 2, 4, 4], dtype=np.int32))
    vals = H(np.array([mx - 5, 10, 20, mn + 5, 10, 20, 30, mx - 5, mx - 5, 1, 40, -10, mn + 5], dtype=dt), type_id=tid)
    G, k, s, f = _run(keys, vals, threads=threads)
    np.testing.assert_array_equal(k[0], [1, 2, 3, 4])
    np.testing.assert_array_equal(f[0], [True, False, True, True])  # the reference checks the flags only; group 2 also has its sum
    assert s[0][1] == 100


def test_unsigned_bool_and_float_are_rejected():
    keys = H(np.array([1, 1], dtype=np.int32))
    for v in (np.array([1, 2], dtype=np.uint32), np.array([True, False]), np.array([1.0, 2.0])):
        with pytest.raises(orc.OracleError):
            orc.groupby([keys], [H(v)], [(0, "sum_overflow", 1)])


def test_total_that_fits_is_not_flagged_whatever_the_order():
    # max, 1, -5: the reference flags this group only when its atomics happen to land in that order; the oracle reports the
    # order-independent outcome (the exact total fits)
    keys = H(np.zeros(3, dtype=np.int32))
    vals = H(np.array([np.iinfo(np.int64).max, 1, -5], dtype=np.int64))
    G, k, s, f = _run(keys, vals)
    assert not f[0][0] and s[0][0] == np.iinfo(np.int64).max - 4
