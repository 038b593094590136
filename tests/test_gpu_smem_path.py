"""GPU parity of the shared-memory pre-aggregation kernel (cudf_b200/csrc/kernels_smem.cu) against the oracle.

Every case forces the kernel (options force_path=2) and also runs the automatic choice; integer results are bit-exact.
Covers: direct and packed key modes, 4- and 8-byte table keys, 4- and 8-byte SUM inputs, COUNT only, null keys
(EXCLUDE and INCLUDE), null values, all-null groups, the EMPTY-image key, values outside the split-accumulator range
(global-atomic bypass), more groups than the table holds (bypass), mid-kernel flushes, int64 wrap-around.
"""
import numpy as np
import pytest

from oracle import oracle as orc

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():  # pragma: no cover
    pytest.skip("needs a CUDA device", allow_module_level=True)

from tests import gpu_util  # noqa: E402

H = orc.HostColumn
SUMCOUNT = [(0, "sum", 1), (0, "count_valid", 1)]


def _check(keys, vals, aggs, include=False, name="", expect_path="smem", auto=True):
    G, path = gpu_util.compare_with_oracle(keys, vals, aggs, include_null_keys=include, options={"force_path": 2}, name=name + "/forced")
    assert path == expect_path, f"{name}: expected the {expect_path} kernel, got {path}"
    if auto:
        gpu_util.compare_with_oracle(keys, vals, aggs, include_null_keys=include, name=name + "/auto")
    return G


@pytest.mark.parametrize("ktype", ["int64", "int32", "uint32", "uint64", "int16", "float32", "float64"])
@pytest.mark.parametrize("vtype", ["int64", "int32", "uint32"])
def test_sum_count_types(ktype, vtype):
    import zlib
    rng = np.random.default_rng(zlib.crc32(f"s-{ktype}-{vtype}".encode()))
    n = 400_000
    keys = [H(rng.integers(0, 3000, n).astype(ktype), rng.random(n) > 0.03)]
    vals = [H(rng.integers(0 if vtype[0] == "u" else -1000, 1001, n).astype(vtype), rng.random(n) > 0.1)]
    _check(keys, vals, SUMCOUNT + [(0, "mean", 1)], name=f"{ktype}/{vtype}")
    # INCLUDE + a nullable 8-byte key needs 65 key bits -> 128-bit packed key -> served by the generic HBM-table kernel
    wide = np.dtype(ktype).itemsize == 8
    _check(keys, vals, SUMCOUNT, include=True, name=f"{ktype}/{vtype}/include", expect_path="global" if wide else "smem")


def test_config_a_shape_reduced():
    # BASELINE config A at reduced N: int64 key x % 1e4 (SplitMix64 generator), int64 values in [-1000, 1000]
    n = 4_000_000
    rows = np.arange(n, dtype=np.uint64)
    k = (orc.splitmix64(42, rows) % np.uint64(10_000)).astype(np.int64)
    v = (orc.splitmix64(43, rows) % np.uint64(2001)).astype(np.int64) - 1000
    G = _check([H(k)], [H(v)], SUMCOUNT, name="configA")
    assert G == 10_000


def test_count_only_and_keys_only():
    rng = np.random.default_rng(1)
    n = 300_000
    k = H(rng.integers(-500, 500, n).astype(np.int64))
    v = H(rng.integers(0, 9, n).astype(np.int32), rng.random(n) > 0.3)
    _check([k], [v], [(0, "count_valid", 1)], name="count_valid only")
    _check([k], [H(v.data)], [(0, "count_all", 1), (0, "count_valid", 1)], name="count_all no mask")
    import cudf_b200 as cb
    gb = cb.GroupBy(cb.Table([gpu_util.to_device(k)]), options={"force_path": 2})
    kt, res = gb.aggregate([])
    assert gb.last_path == "smem" and res == []
    np.testing.assert_array_equal(np.sort(kt.columns()[0].to_numpy()[0]), np.unique(k.data))


def test_all_null_values_group_still_exists():
    k = np.repeat(np.arange(100, dtype=np.int64), 1000)
    v = np.ones(len(k), dtype=np.int64)
    valid = k != 7  # every value of group 7 is null -> SUM null, COUNT 0, group present
    G = _check([H(k)], [H(v, valid)], SUMCOUNT, name="all-null group")
    assert G == 100


def test_empty_image_key_and_out_of_range_values_bypass():
    rng = np.random.default_rng(2)
    n = 500_000
    k = rng.integers(-2, 50, n).astype(np.int64)  # includes -1 == the EMPTY image of the 8-byte table key
    v = rng.integers(-1000, 1000, n).astype(np.int64)
    v[::97] = rng.integers(np.iinfo(np.int64).min, np.iinfo(np.int64).max, len(v[::97]), dtype=np.int64)  # far outside [-2^m, 2^m)
    _check([H(k)], [H(v)], SUMCOUNT, name="bypass rows")
    k32 = rng.integers(-2, 50, n).astype(np.int32)  # 0xFFFFFFFF == EMPTY image of the 4-byte table key
    _check([H(k32)], [H(v)], SUMCOUNT, name="bypass rows/int32 key")


def test_large_magnitudes_force_small_count_field_and_mid_kernel_flushes():
    # sampled |v| ~ 2^30 -> count field 16 bits -> every CTA flushes every < 65536 rows
    rng = np.random.default_rng(3)
    n = 30_000_000
    k = rng.integers(0, 2000, n).astype(np.int32)
    v = rng.integers(-2**30, 2**30, n).astype(np.int64)
    _check([H(k)], [H(v)], SUMCOUNT, name="flushes", auto=False)


def test_int64_wraparound_matches_reference_atomics():
    rng = np.random.default_rng(4)
    n = 300_000
    k = rng.integers(0, 20, n).astype(np.int64)
    v = rng.integers(np.iinfo(np.int64).min, np.iinfo(np.int64).max, n, dtype=np.int64)
    _check([H(k)], [H(v)], SUMCOUNT, name="wrap")


def test_more_groups_than_the_table_holds():
    rng = np.random.default_rng(5)
    n = 2_000_000
    k = rng.integers(0, 60_000, n).astype(np.int64)  # > 14.5K slots: later keys bypass to the HBM table
    v = rng.integers(-1000, 1000, n).astype(np.int64)
    _check([H(k)], [H(v)], SUMCOUNT, name="overflowing table", auto=True)


def test_packed_multi_column_keys():
    rng = np.random.default_rng(6)
    n = 500_000
    a = H(rng.integers(0, 30, n).astype(np.int16), rng.random(n) > 0.02)
    b = H(rng.integers(0, 20, n).astype(np.int8), rng.random(n) > 0.02)
    c = H(rng.integers(0, 5, n).astype(np.int32))
    v = H(rng.integers(-5, 5, n).astype(np.int64), rng.random(n) > 0.2)
    _check([a, b], [v], SUMCOUNT, name="int16+int8 (4-byte table key)")
    _check([a, b], [v], SUMCOUNT, include=True, name="int16+int8 include")
    _check([a, b, c], [v], SUMCOUNT, include=True, name="int16+int8+int32 (8-byte table key)")


def test_sliced_unaligned_columns_take_generic_key_mode_or_generic_path():
    import cudf_b200 as cb
    from cudf_b200 import aggregation as A
    rng = np.random.default_rng(7)
    n = 200_001
    kd = rng.integers(0, 99, n).astype(np.int64); vd = rng.integers(-9, 9, n).astype(np.int64)
    k = cb.Column.from_numpy(kd).sliced(1, n - 1)   # 8-byte offset: not 16-byte aligned
    v = cb.Column.from_numpy(vd).sliced(1, n - 1)
    kt, res = cb.GroupBy(cb.Table([k]), options={"force_path": 2}).aggregate([cb.GroupByRequest(v, [A.sum(), A.count()])])
    Go, ko, ao = orc.groupby([H(kd[1:])], [H(vd[1:])], SUMCOUNT)
    kk = kt.columns()[0].to_numpy()[0]
    o = np.argsort(kk)
    np.testing.assert_array_equal(kk[o], ko[0][0])
    np.testing.assert_array_equal(res[0].columns()[0].to_numpy()[0][o], ao[0][0])
    np.testing.assert_array_equal(res[0].columns()[1].to_numpy()[0][o], ao[1][0])


# ------------------------------------------------------------------ direct-address variant (dense integer key domains)
def _check_direct(keys, vals, aggs, include=False, name=""):
    G, path = gpu_util.compare_with_oracle(keys, vals, aggs, include_null_keys=include, options={"force_path": 4}, name=name)
    assert path == "smem-direct", f"{name}: expected the direct-address kernel, got {path}"
    return G


@pytest.mark.parametrize("ktype", ["int64", "int32", "uint32", "int16", "uint8"])
def test_direct_address_dense_integer_keys(ktype):
    rng = np.random.default_rng(11)
    n = 500_000
    lo, hi = (-3000, 3000) if ktype in ("int64", "int32", "int16") else (0, 200 if ktype == "uint8" else 6000)
    keys = [H(rng.integers(lo, hi, n).astype(ktype), rng.random(n) > 0.03)]
    vals = [H(rng.integers(-1000, 1001, n).astype(np.int64), rng.random(n) > 0.1)]
    _check_direct(keys, vals, SUMCOUNT + [(0, "mean", 1)], name=f"direct/{ktype}")
    # INCLUDE adds a null flag above the value bits: signed keys then span the whole unsigned image (not a dense domain)
    gpu_util.compare_with_oracle(keys, vals, SUMCOUNT, include_null_keys=True, options={"force_path": 4}, name=f"direct/{ktype}/include")


def test_direct_address_outliers_outside_the_sampled_range_bypass():
    rng = np.random.default_rng(12)
    n = 1_000_000
    k = rng.integers(100, 5100, n).astype(np.int64)
    out = rng.choice(n, 40, replace=False)  # 40 rows far outside: the 65536-row sample almost surely misses them
    k[out[:10]] = 10**12 + np.arange(10)
    k[out[10:20]] = -7
    k[out[20:30]] = np.iinfo(np.int64).min
    k[out[30:]] = 99  # just below the sampled minimum (inside the padding)
    v = rng.integers(-1000, 1001, n).astype(np.int64)
    _check_direct([H(k)], [H(v)], SUMCOUNT, name="direct/outliers")


def test_direct_address_all_null_group_and_count_only():
    k = np.repeat(np.arange(300, dtype=np.int32), 400)
    v = np.ones(len(k), dtype=np.int64)
    valid = k != 17
    G = _check_direct([H(k)], [H(v, valid)], SUMCOUNT, name="direct/all-null group")
    assert G == 300
    _check_direct([H(k)], [H(v, valid)], [(0, "count_valid", 1)], name="direct/count only")


def test_direct_address_packed_two_small_keys_and_config_a():
    rng = np.random.default_rng(13)
    n = 400_000
    a = H(rng.integers(0, 3, n).astype(np.int8))
    b = H(rng.integers(0, 2, n).astype(np.int8), rng.random(n) > 0.05)
    v = H(rng.integers(-50, 50, n).astype(np.int32))
    _check_direct([a, b], [v], SUMCOUNT, name="direct/2xint8")
    # INCLUDE puts b's null flag at bit 16: the packed image spans 65K values, more than the direct table holds -> hashed
    gpu_util.compare_with_oracle([a, b], [v], SUMCOUNT, include_null_keys=True, options={"force_path": 4}, name="direct/2xint8/include")
    rows = np.arange(2_000_000, dtype=np.uint64)
    k = (orc.splitmix64(42, rows) % np.uint64(10_000)).astype(np.int64)
    vv = (orc.splitmix64(43, rows) % np.uint64(2001)).astype(np.int64) - 1000
    assert _check_direct([H(k)], [H(vv)], SUMCOUNT, name="direct/configA") == 10_000
