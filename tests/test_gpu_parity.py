"""GPU parity: the CUDA path (through the C ABI) against the reference's known-answer tests and against the oracle.

Bars (BASELINE.json north_star): keys, counts, integer sums, MIN/MAX bit-exact; float64 SUM/MEAN/VAR within 1e-9
relative (atomic order is nondeterministic in the reference too); float32 sums accumulate in float32 like the
reference (tolerance 2e-5, stated here).
"""
import numpy as np
import pytest

from oracle import oracle as orc
from tests import kat

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():  # pragma: no cover
    pytest.skip("needs a CUDA device", allow_module_level=True)

from tests import gpu_util  # noqa: E402

CASES = kat.load_cases()
H = orc.HostColumn


@pytest.mark.parametrize("name,c,kt,vt", CASES, ids=[x[0] for x in CASES])
def test_reference_known_answers(name, c, kt, vt):
    keys, vals, exp = kat.materialize(c, kt, vt)
    G, kout, aout, _ = gpu_util.run_gpu([keys], [vals], [(0, c["agg"], c["ddof"])], include_null_keys=c["include_null_keys"])
    kat.check(name, G, (kout[0][0], kout[0][1], None), (aout[0][0], aout[0][1], None), exp)


def _rand_col(rng, n, dtype, lo, hi, null_frac=0.0):
    if np.dtype(dtype).kind == "f":
        d = rng.integers(lo, hi, n).astype(dtype)
    elif np.dtype(dtype).kind == "b":
        d = rng.integers(0, 2, n).astype(bool)
    else:
        d = rng.integers(lo, hi, n).astype(dtype)
    v = None
    if null_frac > 0:
        v = rng.random(n) >= null_frac
    return H(d, v)


ALL_AGGS = [(0, "sum", 1), (0, "count_valid", 1), (0, "count_all", 1), (0, "min", 1), (0, "max", 1), (0, "mean", 1),
            (0, "variance", 1), (0, "std", 2), (0, "m2", 1), (0, "sum_of_squares", 1)]


@pytest.mark.parametrize("ktype", ["int8", "int16", "int32", "int64", "uint32", "uint64", "float32", "float64"])
@pytest.mark.parametrize("vtype", ["int8", "int32", "int64", "uint16", "float32", "float64"])
def test_random_single_key_all_aggs(ktype, vtype):
    import zlib
    rng = np.random.default_rng(zlib.crc32(f"{ktype}-{vtype}".encode()))
    n = 50_000
    keys = [_rand_col(rng, n, ktype, 0, 100, 0.05)]
    vals = [_rand_col(rng, n, vtype, 0, 100, 0.1)]
    gpu_util.compare_with_oracle(keys, vals, ALL_AGGS, name=f"{ktype}/{vtype}")
    gpu_util.compare_with_oracle(keys, vals, ALL_AGGS, include_null_keys=True, name=f"{ktype}/{vtype}/include")


@pytest.mark.parametrize("force_path", [0, 1])
@pytest.mark.parametrize("groups", [4, 1000, 200_000])
def test_random_int64_sum_count(groups, force_path):
    rng = np.random.default_rng(groups)
    n = 1_000_000
    keys = [H(rng.integers(-groups // 2, groups // 2, n).astype(np.int64))]
    vals = [H(rng.integers(-1000, 1001, n).astype(np.int64))]
    G, path = gpu_util.compare_with_oracle(keys, vals, [(0, "sum", 1), (0, "count_valid", 1)], options={"force_path": force_path},
                                           name=f"G={groups}")
    assert G == len(np.unique(keys[0].data))


def test_large_values_wrap_like_int64():
    rng = np.random.default_rng(5)
    n = 200_000
    keys = [H(rng.integers(0, 50, n).astype(np.int32))]
    vals = [H(rng.integers(np.iinfo(np.int64).min, np.iinfo(np.int64).max, n, dtype=np.int64))]
    gpu_util.compare_with_oracle(keys, vals, [(0, "sum", 1), (0, "min", 1), (0, "max", 1), (0, "sum_of_squares", 1)], name="wrap")


def test_multi_column_keys_packed_64_and_128():
    rng = np.random.default_rng(11)
    n = 300_000
    a = _rand_col(rng, n, "int8", 0, 3, 0.02)
    b = _rand_col(rng, n, "int8", 0, 2, 0.02)
    v1 = _rand_col(rng, n, "float64", 1, 51, 0.1)
    v2 = _rand_col(rng, n, "float64", 900, 105000, 0.0)
    aggs = [(0, "mean", 1), (0, "variance", 1), (0, "count_valid", 1), (1, "mean", 1), (1, "variance", 1), (1, "count_valid", 1)]
    gpu_util.compare_with_oracle([a, b], [v1, v2], aggs, name="q1-shape")
    gpu_util.compare_with_oracle([a, b], [v1, v2], aggs, include_null_keys=True, name="q1-shape/include")
    # two int64 keys (pandas config shape) -> 128-bit packed key
    k1 = _rand_col(rng, n, "int64", 0, 32, 0.01)
    k2 = _rand_col(rng, n, "int64", -16, 16, 0.0)
    w = H(rng.standard_normal(n))
    gpu_util.compare_with_oracle([k1, k2], [w], [(0, "mean", 1), (0, "sum", 1), (0, "min", 1)], name="2xint64")
    k1nn = H(k1.data)  # 128 value bits + a null flag would need 129 bits: only non-nullable 2xint64 packs
    gpu_util.compare_with_oracle([k1nn, k2], [w], [(0, "mean", 1), (0, "count_all", 1)], include_null_keys=True, name="2xint64/include")
    # mixed widths straddling words: int32 + int64 + int16 (+ null flags)
    k3 = _rand_col(rng, n, "int32", 0, 7, 0.03)
    k4 = _rand_col(rng, n, "int16", 0, 5, 0.03)
    gpu_util.compare_with_oracle([k3, k1, k4], [w], [(0, "sum", 1), (0, "max", 1)], include_null_keys=True, name="mixed-128")
    k5 = _rand_col(rng, n, "int8", -3, 3, 0.03)
    gpu_util.compare_with_oracle([k3, k1, k4, k5], [w], [(0, "sum", 1), (0, "count_valid", 1)], include_null_keys=False, name="mixed-120")


def test_special_keys_all_ones_and_float_canonicalisation():
    # the packed key equal to the EMPTY sentinel (int64 -1) must still form its own group
    keys = [H(np.array([-1, 5, -1, 5, 0, -1], dtype=np.int64))]
    vals = [H(np.array([1, 2, 3, 4, 5, 6], dtype=np.int64))]
    gpu_util.compare_with_oracle(keys, vals, [(0, "sum", 1), (0, "count_valid", 1)], name="all-ones key")
    k2 = [H(np.array([-1, -1, 3], dtype=np.int64)), H(np.array([-1, -1, -1], dtype=np.int64))]
    gpu_util.compare_with_oracle(k2, vals[:1] and [H(np.array([1, 2, 3], dtype=np.int64))], [(0, "sum", 1)], name="all-ones 128")
    # -0.0 == +0.0 and NaN == NaN as keys (row_operator/equality.cuh:59-89)
    fk = [H(np.array([0.0, -0.0, np.nan, np.nan, 1.5, 1.5], dtype=np.float64))]
    fv = [H(np.array([1, 2, 3, 4, 5, 6], dtype=np.int32))]
    G, kout, aout, _ = gpu_util.run_gpu(fk, fv, [(0, "sum", 1)])
    assert G == 3
    np.testing.assert_array_equal(aout[0][0], np.array([3, 11, 7]))  # 0, 1.5, NaN (NaN sorts last)


def test_nan_values_min_max_terminate():
    # max_tests.cpp:526-549: NaN,NaN MAX must terminate (result itself is order-dependent in the reference)
    keys = [H(np.array([1, 1], dtype=np.int32))]
    for dt in (np.float32, np.float64):
        vals = [H(np.array([np.nan, np.nan], dtype=dt))]
        G, _, _, _ = gpu_util.run_gpu(keys, vals, [(0, "max", 1), (0, "min", 1)])
        assert G == 1


def test_shared_memory_kernel_shapes_from_reference():
    # max_tests.cpp:552-597: 512 unique keys; 1.28M rows over 128 shuffled repeated keys
    k = np.arange(512, dtype=np.int32)
    gpu_util.compare_with_oracle([H(k)], [H(k)], [(0, "max", 1)], name="all_unique_keys")
    k = np.tile(np.arange(128, dtype=np.int32), 10000)
    np.random.default_rng(0).shuffle(k)
    gpu_util.compare_with_oracle([H(k)], [H(k)], [(0, "max", 1)], name="repeated_keys")


def test_duplicate_requests_and_zero_requests():
    import cudf_b200 as cb
    from cudf_b200 import aggregation as A
    k = cb.Column.from_numpy(np.array([1, 2, 3, 1, 2, 2, 1, 3, 3, 2], dtype=np.int32))
    v = cb.Column.from_numpy(np.arange(10, dtype=np.int32))
    gb = cb.GroupBy(cb.Table([k]))
    # keys_tests.cpp:355-409: duplicate aggregations in one request, same column in two requests
    kt, res = gb.aggregate([cb.GroupByRequest(v, [A.sum(), A.sum()]), cb.GroupByRequest(v, [A.sum()])])
    kk = kt.columns()[0].to_numpy()[0]
    o = np.argsort(kk)
    for col in (res[0].columns()[0], res[0].columns()[1], res[1].columns()[0]):
        np.testing.assert_array_equal(col.to_numpy()[0][o], [9, 19, 17])
    # compute_groupby.cu:113-130: no requests -> just the unique keys
    kt, res = gb.aggregate([])
    assert sorted(kt.columns()[0].to_numpy()[0].tolist()) == [1, 2, 3] and res == []


def test_error_behaviour_matches_reference():
    import cudf_b200 as cb
    from cudf_b200 import aggregation as A
    k = cb.Column.from_numpy(np.array([1, 2, 3], dtype=np.int32))
    v5 = cb.Column.from_numpy(np.arange(5, dtype=np.int32))
    with pytest.raises(cb.LogicError, match="Size mismatch"):  # keys_tests.cpp:215-229, groupby.cu:225-229
        cb.GroupBy(cb.Table([k])).aggregate([cb.GroupByRequest(v5, [A.count()])])
    ts = cb.Column.from_numpy(np.arange(3, dtype=np.int64), type_id=cb.TypeId.TIMESTAMP_SECONDS)
    with pytest.raises(cb.LogicError, match="Invalid type/aggregation"):  # groupby.cu:186-201
        cb.GroupBy(cb.Table([k])).aggregate([cb.GroupByRequest(ts, [A.sum()])])
    v3 = cb.Column.from_numpy(np.arange(3, dtype=np.int32))
    with pytest.raises((cb.UnsupportedError, cb.LogicError)):  # python/pylibcudf/tests/test_groupby.py: ANY/ALL raise
        cb.GroupBy(cb.Table([k])).aggregate([cb.GroupByRequest(v3, [A.any()])])
    assert not cb.can_use_hash_groupby(cb.Table([k]), [cb.GroupByRequest(v3, [A.median()])])
    assert cb.can_use_hash_groupby(cb.Table([k]), [cb.GroupByRequest(v3, [A.mean(), A.std()])])


def test_chrono_and_decimal_types_pass_through():
    import cudf_b200 as cb
    rng = np.random.default_rng(3)
    n = 10_000
    keys = [H(rng.integers(0, 20, n).astype(np.int64), None, cb.TypeId.TIMESTAMP_NANOSECONDS)]
    dur = [H(rng.integers(-500, 500, n).astype(np.int32), rng.random(n) > 0.1, cb.TypeId.DURATION_DAYS)]
    gpu_util.compare_with_oracle(keys, dur, [(0, "sum", 1), (0, "min", 1), (0, "max", 1), (0, "mean", 1), (0, "count_valid", 1)],
                                 name="duration")
    dec = [H(rng.integers(-10**6, 10**6, n).astype(np.int64), None, cb.TypeId.DECIMAL64)]
    gpu_util.compare_with_oracle(keys, dec, [(0, "sum", 1), (0, "min", 1), (0, "max", 1)], name="decimal64")


def test_sliced_columns_offsets():
    import cudf_b200 as cb
    from cudf_b200 import aggregation as A
    rng = np.random.default_rng(9)
    n = 5000
    kd = rng.integers(0, 9, n).astype(np.int32); vd = rng.integers(0, 100, n).astype(np.int64)
    vv = rng.random(n) > 0.2
    k = cb.Column.from_numpy(kd).sliced(37, 3001)
    v = cb.Column.from_numpy(vd, vv).sliced(37, 3001)
    kt, res = cb.GroupBy(cb.Table([k])).aggregate([cb.GroupByRequest(v, [A.sum(), A.count()])])
    Go, ko, ao = orc.groupby([H(kd[37:3038])], [H(vd[37:3038], vv[37:3038])], [(0, "sum", 1), (0, "count_valid", 1)])
    kk = kt.columns()[0].to_numpy()[0]
    o = np.argsort(kk)
    np.testing.assert_array_equal(kk[o], ko[0][0])
    np.testing.assert_array_equal(res[0].columns()[0].to_numpy()[0][o], ao[0][0])
    np.testing.assert_array_equal(res[0].columns()[1].to_numpy()[0][o], ao[1][0])


@pytest.mark.parametrize("vtype", ["int64", "uint64", "float64"])
def test_argmin_argmax_of_64bit_values(vtype):
    # 64-bit values take the two-pass form (MIN / MAX first, then the smallest row holding the extremum); ties, nulls and an
    # all-null group included. The reference serves these through its index CAS loop (device_aggregators.cuh:131-160).
    rng = np.random.default_rng(77)
    n = 150_000
    k = orc.HostColumn(rng.integers(0, 5_000, n).astype(np.int64))
    raw = rng.integers(-40, 40, n) if vtype != "uint64" else rng.integers(0, 80, n)  # few distinct values: many ties
    valid = rng.random(n) > 0.15
    valid[k.data == 7] = False  # a group without any valid value -> null result
    v = orc.HostColumn(raw.astype(vtype), valid)
    G, path = gpu_util.compare_with_oracle([k], [v], [(0, "argmin", 1), (0, "argmax", 1), (0, "min", 1), (0, "max", 1), (0, "sum", 1)],
                                           name=f"arg64/{vtype}")
    assert path == "global", path
    w = orc.HostColumn(rng.permutation(n).astype(vtype))  # no nulls, no ties
    gpu_util.compare_with_oracle([k], [w], [(0, "argmax", 1), (0, "argmin", 1)], name=f"arg64/{vtype}/unique")


@pytest.mark.parametrize("vtype", ["int8", "int16", "int32", "uint8", "uint16", "uint32", "float32"])
@pytest.mark.parametrize("nullable", [False, True])
def test_argmin_argmax_of_narrow_values_mixed_sign(vtype, nullable):
    # values of at most 32 bits travel as one packed (order image, row) cell: the image must order NEGATIVE values below
    # positive ones (round-1 advisor finding: the sign flip of the 64-bit image was shifted out of the cell)
    rng = np.random.default_rng(91)
    n = 120_000
    k = orc.HostColumn(rng.integers(0, 3_000, n).astype(np.int64))
    signed = vtype.startswith("int") or vtype == "float32"
    raw = rng.integers(-100, 100, n) if signed else rng.integers(0, 200, n)
    valid = None
    if nullable:
        valid = rng.random(n) > 0.2
        valid[k.data == 11] = False
    v = orc.HostColumn(raw.astype(vtype), valid)
    gpu_util.compare_with_oracle([k], [v], [(0, "argmin", 1), (0, "argmax", 1), (0, "min", 1), (0, "max", 1)], name=f"arg32/{vtype}")
    # two rows only, {-1, 1}: ARGMIN must be the row of -1
    if signed:
        k2 = orc.HostColumn(np.zeros(2, dtype=np.int64))
        v2 = orc.HostColumn(np.array([1, -1]).astype(vtype))
        G, kg, ag, _ = gpu_util.run_gpu([k2], [v2], [(0, "argmin", 1), (0, "argmax", 1)])
        assert G == 1 and ag[0][0][0] == 1 and ag[1][0][0] == 0
