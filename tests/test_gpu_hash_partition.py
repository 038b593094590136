"""GPU parity tests of hash_partition (bgb_hash_partition through cudf_b200.partitioning) against oracle/hash_partition.py.

Modelled on the reference's own tests: cpp/tests/partitioning/hash_partition_test.cpp (offsets shape, empty inputs,
every row lands in the partition its hash names, partitions are a permutation of the input, nulls, mixed types,
identity hash) and python/cudf/cudf/tests/dataframe/methods/test_partition_by_hash.py (known-answer partitions).
The order of rows inside a partition is unspecified in the reference too, so each partition is compared as a multiset
of (row values, validity) tuples; offsets and per-row partition numbers are compared exactly.
"""
import numpy as np
import pytest

import cudf_b200 as cb
from cudf_b200 import partitioning as part
from oracle import hash_partition as hp

pytestmark = pytest.mark.gpu


def col(a, valid=None, dtype=None, type_id=None):
    return cb.Column.from_numpy(np.asarray(a, dtype=dtype), None if valid is None else np.asarray(valid, dtype=bool), type_id=type_id)


def rows_of(columns, lo, hi):
    """multiset key of rows [lo, hi): tuples of (bits | None) per column"""
    out = []
    cols = []
    for data, valid in columns:
        bits = np.ascontiguousarray(data).view({1: np.uint8, 2: np.uint16, 4: np.uint32, 8: np.uint64}[data.dtype.itemsize])
        cols.append((bits, valid))
    for r in range(lo, hi):
        out.append(tuple(int(b[r]) if (v is None or v[r]) else None for b, v in cols))
    return sorted(out, key=lambda t: tuple((-1 if x is None else x) for x in t))


def check(host_cols, key_idx, nparts, seed=0, hash_function="murmur3", keys_as_table=False):
    """host_cols: list of (ndarray, valid | None)"""
    dcols = [col(d, v) for d, v in host_cols]
    t = cb.Table(dcols)
    keys = cb.Table([dcols[i] for i in key_idx]) if keys_as_table else key_idx
    hid = part.HashId.HASH_MURMUR3 if hash_function == "murmur3" else part.HashId.HASH_IDENTITY
    out, offsets = part.hash_partition(t, keys, nparts, hid, seed)
    order, want_offsets = hp.hash_partition(host_cols, key_idx, nparts, seed, hash_function)
    assert offsets == want_offsets
    n = len(host_cols[0][0])
    assert out.num_columns() == len(host_cols) and out.num_rows() == n
    got = [c.to_numpy() for c in out.columns()]
    for (d, v), (gd, gv), dc in zip(host_cols, got, out.columns()):
        assert gd.dtype == d.dtype
        nulls = 0 if v is None else int((~np.asarray(v)).sum())
        assert dc.null_count == nulls
        assert int((~gv).sum()) == nulls
    # every output row sits in the partition its own hash names
    gpid = hp.partition_numbers([(got[i][0], got[i][1]) for i in key_idx], nparts, seed, hash_function)
    want_pid = np.repeat(np.arange(nparts), np.diff(offsets))
    assert (gpid == want_pid).all()
    # and every partition holds exactly the oracle's rows
    ordered = [(d[order], None if v is None else np.asarray(v)[order]) for d, v in host_cols]
    gv_cols = [(gd, gv) for gd, gv in got]
    for p in range(nparts):
        lo, hi = offsets[p], offsets[p + 1]
        if hi - lo > 20000:
            continue  # the permutation check below covers big partitions
        assert rows_of(gv_cols, lo, hi) == rows_of(ordered, lo, hi)
    return out, offsets


@pytest.mark.parametrize("nparts", [1, 2, 3, 5, 8, 32, 100, 257, 4096])
def test_int64_key_with_value(nparts):
    rng = np.random.default_rng(nparts)
    n = 30011
    check([(rng.integers(-10**12, 10**12, n), None), (rng.integers(0, 1000, n).astype(np.int32), None)], [0], nparts)


@pytest.mark.parametrize("dtype", [np.int8, np.int16, np.int32, np.int64, np.uint8, np.uint16, np.uint32, np.uint64, np.float32,
                                   np.float64, np.bool_])
@pytest.mark.parametrize("seed", [0, 12345])
def test_key_types(dtype, seed):
    rng = np.random.default_rng(7)
    n = 5003
    if dtype == np.bool_:
        k = rng.integers(0, 2, n).astype(bool)
    elif np.dtype(dtype).kind == "f":
        k = rng.standard_normal(n).astype(dtype)
        k[::17] = 0.0
        k[1::17] = -0.0
        k[2::17] = np.nan
        k[3::17] = -np.nan
        k[4::17] = np.inf
    else:
        info = np.iinfo(dtype)
        k = rng.integers(info.min, info.max, n, dtype=dtype, endpoint=True)
    check([(k, None), (np.arange(n, dtype=np.int64), None)], [0], 7, seed)


def test_bool_bytes_other_than_one():
    # element<bool>() reads any non-zero byte as true; the hash sees uint8 0/1 (murmurhash3_x86_32.cuh:46-51)
    raw = np.array([0, 1, 2, 255, 0, 7], dtype=np.uint8)
    dc = cb.Column.from_numpy(raw, None, type_id=cb.TypeId.BOOL8)
    out, offsets = part.hash_partition(cb.Table([dc, col(np.arange(6, dtype=np.int32))]), [0], 2)
    want = hp.partition_numbers([(raw != 0, None)], 2)
    got_rows = out.columns()[1].to_numpy()[0]
    got_pid = np.repeat(np.arange(2), np.diff(offsets))
    assert (want[got_rows] == got_pid).all()


@pytest.mark.parametrize("type_id,np_dtype", [(cb.TypeId.TIMESTAMP_DAYS, np.int32), (cb.TypeId.TIMESTAMP_NANOSECONDS, np.int64),
                                              (cb.TypeId.DURATION_SECONDS, np.int64), (cb.TypeId.DECIMAL32, np.int32),
                                              (cb.TypeId.DECIMAL64, np.int64)])
def test_rep_types_hash_as_their_integers(type_id, np_dtype):
    rng = np.random.default_rng(11)
    k = rng.integers(-10**6, 10**6, 4000).astype(np_dtype)
    dk = cb.Column.from_numpy(k, None, type_id=type_id)
    out, offsets = part.hash_partition(cb.Table([dk]), [0], 5, seed=3)
    _, want = hp.hash_partition([(k, None)], [0], 5, 3)
    assert offsets == want
    assert out.columns()[0].type_id == type_id


@pytest.mark.parametrize("keys_as_table", [False, True])
def test_multi_column_keys_and_nulls(keys_as_table):
    rng = np.random.default_rng(5)
    n = 20000
    k0 = rng.integers(0, 50, n).astype(np.int32)
    k1 = rng.integers(0, 9, n).astype(np.int64)
    k2 = rng.standard_normal(n).astype(np.float32)
    v0 = rng.integers(0, 1 << 40, n)
    v1 = rng.integers(0, 200, n).astype(np.uint8)
    check([(k0, rng.random(n) > 0.1), (v0, rng.random(n) > 0.3), (k1, None), (k2, rng.random(n) > 0.5), (v1, rng.random(n) > 0.2)],
          [0, 2, 3], 13, seed=99, keys_as_table=keys_as_table)


def test_column_order_matters():
    rng = np.random.default_rng(2)
    n = 4096
    a, b = rng.integers(0, 100, n).astype(np.int32), rng.integers(0, 100, n).astype(np.int32)
    check([(a, None), (b, None)], [0, 1], 11)
    check([(a, None), (b, None)], [1, 0], 11)
    p01 = hp.partition_numbers([(a, None), (b, None)], 11)
    p10 = hp.partition_numbers([(b, None), (a, None)], 11)
    assert (p01 != p10).any()


def test_all_null_keys_go_to_one_partition():
    n = 1000
    out, offsets = check([(np.arange(n, dtype=np.int64), np.zeros(n, dtype=bool)), (np.arange(n, dtype=np.int32), None)], [0], 3)
    assert offsets == [0, n, n, n]  # UINT32_MAX % 3 == 0


def test_mask_without_nulls():
    n = 3000
    rng = np.random.default_rng(4)
    check([(rng.integers(0, 99, n), np.ones(n, dtype=bool)), (rng.integers(0, 99, n), np.ones(n, dtype=bool))], [0], 4)


def test_reference_known_answers():
    # python/cudf/cudf/tests/dataframe/methods/test_partition_by_hash.py:41-133
    out, offsets = check([(np.arange(10, dtype=np.int64), None), (np.arange(10, dtype=np.int64) + 100, np.arange(10) % 2 == 0)], [0], 3)
    assert offsets == [0, 1, 3, 10]
    keys = out.columns()[0].to_numpy()[0]
    assert keys[:1].tolist() == [9] and sorted(keys[1:3].tolist()) == [0, 8] and sorted(keys[3:].tolist()) == list(range(1, 8))
    out, offsets = check([(np.arange(5, dtype=np.int64), np.array([1, 1, 1, 0, 0], dtype=bool)), (np.arange(5, dtype=np.int64) + 100, None)], [0], 3)
    assert offsets == [0, 2, 3, 5]
    assert sorted(out.columns()[1].to_numpy()[0][:2].tolist()) == [103, 104]
    out, offsets = check([(np.array([1, 2, 3, 4, 5], dtype=np.int64), None), (np.array([3, 2, 1, 4, 5], dtype=np.int64), None)], [1], 2)
    assert sorted(out.columns()[1].to_numpy()[0][:offsets[1]].tolist()) == [1, 2, 4]


def test_empty_inputs():
    t = cb.Table([col(np.arange(4, dtype=np.int32)), col(np.arange(4, dtype=np.float64))])
    out, offsets = part.hash_partition(t, [0], 0)
    assert offsets == [0] and out.num_rows() == 0 and out.num_columns() == 2
    out, offsets = part.hash_partition(t, [], 3)
    assert offsets == [0, 0, 0, 0] and out.num_rows() == 0
    out, offsets = part.hash_partition(t, cb.Table([]), 3)
    assert offsets == [0, 0, 0, 0] and out.num_rows() == 0
    e = cb.Table([col(np.zeros(0, dtype=np.int32))])
    out, offsets = part.hash_partition(e, [0], 2)
    assert offsets == [0, 0, 0] and out.num_rows() == 0


def test_more_partitions_than_rows():
    check([(np.array([5, 6, 7], dtype=np.int32), None)], [0], 64)


def test_errors():
    t = cb.Table([col(np.arange(4, dtype=np.int32))])
    with pytest.raises(ValueError):
        part.hash_partition(t, cb.Table([col(np.arange(3, dtype=np.int32))]), 2)
    with pytest.raises(IndexError):
        part.hash_partition(t, [1], 2)
    with pytest.raises(cb.UnsupportedError):
        part.hash_partition(t, [0], 5000)
    ts = cb.Table([cb.Column.from_numpy(np.arange(4, dtype=np.int64), None, type_id=cb.TypeId.TIMESTAMP_SECONDS)])
    with pytest.raises(cb.LogicError, match="IdentityHash does not support this data type"):
        part.hash_partition(ts, [0], 2, part.HashId.HASH_IDENTITY)


@pytest.mark.parametrize("dtype", [np.int8, np.int16, np.int32, np.int64, np.uint8, np.uint32, np.uint64, np.bool_])
def test_identity_hash(dtype):
    rng = np.random.default_rng(9)
    n = 6000
    if dtype == np.bool_:
        k = rng.integers(0, 2, n).astype(bool)
    else:
        info = np.iinfo(dtype)
        k = rng.integers(info.min, info.max, n, dtype=dtype, endpoint=True)
    check([(k, rng.random(n) > 0.1), (np.arange(n, dtype=np.int32), None)], [0], 6, hash_function="identity")
    check([(k, None), (k, None)], [0, 1], 10, hash_function="identity")


def test_key_table_not_from_input():
    rng = np.random.default_rng(12)
    n = 5000
    v = rng.integers(0, 1000, n)
    k = rng.integers(0, 77, n).astype(np.int32)
    out, offsets = part.hash_partition(cb.Table([col(v)]), cb.Table([col(k)]), 4)
    assert out.num_columns() == 1
    _, want = hp.hash_partition([(k, None)], [0], 4)
    assert offsets == want
    assert sorted(out.columns()[0].to_numpy()[0].tolist()) == sorted(v.tolist())


def test_wide_table_uses_gather_map():
    rng = np.random.default_rng(13)
    n = 9000
    cols = [(rng.integers(0, 500, n).astype(np.int32), None)]
    for i in range(20):
        dt = [np.int8, np.int16, np.int32, np.int64, np.float64][i % 5]
        cols.append((rng.integers(0, 100, n).astype(dt), (rng.random(n) > 0.25) if i % 3 == 0 else None))
    check(cols, [0], 9)


def test_sliced_input():
    rng = np.random.default_rng(14)
    n, off = 4000, 37
    k = rng.integers(0, 1000, n + off).astype(np.int64)
    v = rng.integers(0, 1000, n + off).astype(np.int32)
    valid = rng.random(n + off) > 0.3
    dk, dv = col(k).sliced(off, n), col(v, valid).sliced(off, n)
    dv.null_count = int((~valid[off:]).sum())
    out, offsets = part.hash_partition(cb.Table([dk, dv]), [0], 5)
    order, want = hp.hash_partition([(k[off:], None), (v[off:], valid[off:])], [0], 5)
    assert offsets == want
    gk = out.columns()[0].to_numpy()[0]
    gv, gvalid = out.columns()[1].to_numpy()
    assert sorted(zip(gk.tolist(), np.where(gvalid, gv, -1).tolist())) == sorted(zip(k[off:].tolist(), np.where(valid[off:], v[off:], -1).tolist()))


def test_large_is_a_permutation():
    n = 20_000_000
    rng = np.random.default_rng(15)
    k = rng.integers(0, 1 << 40, n)
    v = np.arange(n, dtype=np.int64)
    out, offsets = part.hash_partition(cb.Table([col(k), col(v)]), [0], 64)
    gk, gv = out.columns()[0].to_numpy()[0], out.columns()[1].to_numpy()[0]
    assert (k[gv] == gk).all()                      # rows stayed intact
    assert (np.sort(gv) == v).all()                 # a permutation
    pid = hp.partition_numbers([(gk, None)], 64)
    assert (pid == np.repeat(np.arange(64), np.diff(offsets))).all()
    assert offsets[-1] == n
