"""GPU parity of the bucketed high-cardinality path (cudf_b200/csrc/kernels_bucket.cu): one or two radix-scatter levels
into fixed-capacity buckets, one shared-memory table slice per bucket, overflow rows through HBM atomics. Forced with
options force_path=7 (bucket_slice_log2 shrinks the slices so that small inputs take two levels), against the oracle."""
import numpy as np
import pytest

from oracle import oracle as orc

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():  # pragma: no cover
    pytest.skip("needs a CUDA device", allow_module_level=True)

from tests import gpu_util  # noqa: E402

H = orc.HostColumn


def _check(keys, vals, aggs, include=False, name="", capacity_log2=0, slice_log2=0, expect="bucketed"):
    opts = {"force_path": 7}
    if capacity_log2:
        opts["capacity_log2"] = capacity_log2
    if slice_log2:
        opts["bucket_slice_log2"] = slice_log2
    G, path = gpu_util.compare_with_oracle(keys, vals, aggs, include_null_keys=include, options=opts, name=name)
    if expect:
        assert path == expect, f"{name}: expected the {expect} path, got {path}"
    return G


def _a9(n, g, seed=42):
    rows = np.arange(n, dtype=np.uint64)
    k = (orc.splitmix64(seed, rows) % np.uint64(g)).astype(np.int64)
    v = (orc.splitmix64(seed + 1, rows) % np.uint64(2001)).astype(np.int64) - 1000
    return k, v


@pytest.mark.parametrize("slice_log2", [0, 8, 5])  # 0: automatic slice (one level here); 8: 2 levels; 5: 9+8 bits
def test_config_a9_shape_reduced(slice_log2):
    k, v = _a9(4_000_000, 1_000_000)
    G = _check([H(k)], [H(v)], [(0, "sum", 1), (0, "count_valid", 1)], name=f"A9/slice{slice_log2}", slice_log2=slice_log2)
    assert G == len(np.unique(k))


@pytest.mark.parametrize("slice_log2", [0, 7])
def test_config_b_shape_reduced(slice_log2):
    # int32 key, float64 value with 10% nulls, SUM + MIN + MAX: the validity word rides in the upper half of the key word
    n, g = 3_000_000, 400_000
    rows = np.arange(n, dtype=np.uint64)
    k = (orc.splitmix64(42, rows) % np.uint64(g)).astype(np.int32)
    v = (orc.splitmix64(43, rows) % np.uint64(1_000_000)).astype(np.float64)
    valid = (orc.splitmix64(44, rows) % np.uint64(10)) != 0
    _check([H(k)], [H(v, valid)], [(0, "sum", 1), (0, "min", 1), (0, "max", 1)], name="configB", slice_log2=slice_log2)


@pytest.mark.parametrize("ktype", ["int8", "int16", "int32", "int64", "uint64", "float32", "float64"])
def test_key_types_nulls_and_all_aggs(ktype):
    rng = np.random.default_rng(21)
    n = 300_000
    keys = [H(rng.integers(0, 100 if ktype == "int8" else 20_000, n).astype(ktype), rng.random(n) > 0.04)]
    vals = [H(rng.integers(-100, 100, n).astype(np.int32), rng.random(n) > 0.1), H(rng.integers(0, 1000, n).astype(np.float64), rng.random(n) > 0.2)]
    aggs = [(0, "sum", 1), (0, "count_valid", 1), (0, "count_all", 1), (0, "min", 1), (1, "max", 1), (1, "mean", 1), (1, "variance", 1),
            (1, "sum", 1), (0, "product", 1), (0, "sum_of_squares", 1)]
    # int8 keys: 100 groups over 256+ buckets is heavily skewed -> the overflow region may fill and the call reruns partitioned
    expect = None if ktype == "int8" else "bucketed"
    _check(keys, vals, aggs, name=f"bucket/{ktype}", slice_log2=6, expect=expect)
    # INCLUDE adds a null flag bit: 64-bit key types then need a 128-bit image, which this path leaves to the partitioned one
    wide = ktype in ("int64", "uint64", "float64")
    _check(keys, vals, aggs, include=True, name=f"bucket/{ktype}/include", slice_log2=6, expect="partitioned" if wide else expect)


def test_many_masked_value_columns_separate_validity_column():
    # 64-bit key image + 9 masked value columns: the validity word gets its own 4-byte column; with 3: one byte
    rng = np.random.default_rng(5)
    n = 200_000
    k = H(rng.integers(-(1 << 40), 1 << 40, 30_000).astype(np.int64)[rng.integers(0, 30_000, n)])
    for ncols in (3, 9):
        vals = [H(rng.integers(-1000, 1000, n).astype(np.int64), rng.random(n) > 0.1 * (c + 1) / ncols) for c in range(ncols)]
        aggs = [(c, "sum", 1) for c in range(ncols)] + [(c, "count_valid", 1) for c in range(ncols)]
        _check([k], vals, aggs, name=f"{ncols} masked columns", slice_log2=7)


def test_mixed_width_values_two_levels():
    rng = np.random.default_rng(6)
    n = 400_000
    k1 = H(rng.integers(0, 3000, n).astype(np.int32), rng.random(n) > 0.02)
    k2 = H(rng.integers(0, 20, n).astype(np.int16))
    vals = [H(rng.integers(0, 255, n).astype(np.uint8), rng.random(n) > 0.5), H(rng.integers(-30000, 30000, n).astype(np.int16)),
            H((rng.random(n) + 0.5).astype(np.float32), rng.random(n) > 0.1), H(rng.standard_normal(n))]
    aggs = [(0, "sum", 1), (0, "max", 1), (1, "min", 1), (1, "sum", 1), (2, "sum", 1), (2, "min", 1), (3, "mean", 1), (3, "std", 1), (0, "count_all", 1)]
    _check([k1, k2], vals, aggs, name="mixed widths", slice_log2=5)
    _check([k1, k2], vals, aggs, include=True, name="mixed widths/include", slice_log2=5)


def test_skewed_keys_overflow_region_and_fallback():
    # one key holds 30 % of the rows: its bucket overflows. A small share is folded in through the overflow pass ...
    rng = np.random.default_rng(7)
    n = 2_000_000
    k = rng.integers(0, 500_000, n).astype(np.int64)
    heavy = rng.random(n) < 0.04
    k[heavy] = 123456
    v = rng.integers(-1000, 1000, n).astype(np.int64)
    _check([H(k)], [H(v)], [(0, "sum", 1), (0, "count_valid", 1), (0, "min", 1)], name="skew 4%", slice_log2=8)
    # ... a large share fills the overflow region (N/16 rows) and the call reruns on the exact-offset partitioned path
    heavy = rng.random(n) < 0.3
    k[heavy] = 123456
    _check([H(k)], [H(v)], [(0, "sum", 1), (0, "count_valid", 1), (0, "min", 1)], name="skew 30%", slice_log2=8, expect="partitioned")


def test_all_ones_key_reserved_slot():
    rng = np.random.default_rng(8)
    n = 300_000
    k = rng.integers(0, 50_000, n).astype(np.int64)
    k[rng.random(n) < 0.01] = -1  # image == EMPTY sentinel
    v = rng.integers(-5, 5, n).astype(np.int64)
    _check([H(k)], [H(v)], [(0, "sum", 1), (0, "count_valid", 1), (0, "max", 1)], name="all-ones key", slice_log2=7)


def test_table_too_small_retries_bigger():
    # explicit capacity smaller than the group count: the slices overflow, the host retries at full size
    k, v = _a9(500_000, 200_000)
    G = _check([H(k)], [H(v)], [(0, "sum", 1), (0, "count_valid", 1)], name="undersized", capacity_log2=16, slice_log2=8, expect=None)
    assert G == len(np.unique(k))


def test_all_keys_null_and_tiny_inputs():
    n = 70_000
    k = H(np.arange(n, dtype=np.int32), np.zeros(n, dtype=bool))
    v = H(np.ones(n, dtype=np.int64))
    G, _, _, path = gpu_util.run_gpu([k], [v], [(0, "sum", 1)], options={"force_path": 7})
    assert G == 0
    _check([H(np.array([5, 5, 7], dtype=np.int64))], [H(np.array([1, 2, 3], dtype=np.int64))], [(0, "sum", 1), (0, "count_all", 1)], name="tiny")


def test_automatic_choice_takes_the_bucketed_path_for_large_tables():
    # table > 96 MB and >= 2^22 rows: the automatic plan buckets
    k, v = _a9(6_000_000, 3_000_000)
    G, path = gpu_util.compare_with_oracle([H(k)], [H(v)], [(0, "sum", 1), (0, "count_valid", 1)], name="auto")
    assert path == "bucketed", path
    assert G == len(np.unique(k))


def test_long_buckets_are_walked_in_chunks():
    # mid cardinality: few slices with > 65535 rows each -> chunks of 16384 rows, digit sums folded into 64-bit totals
    rng = np.random.default_rng(9)
    n, g = 6_000_000, 20_000
    k = rng.integers(0, g, n).astype(np.int64)
    v = rng.integers(-(1 << 40), 1 << 40, n).astype(np.int64)          # all three digit arrays in use
    w = rng.integers(-30_000, 30_000, n).astype(np.int32)
    x = H(rng.integers(0, 1000, n).astype(np.float64), rng.random(n) > 0.1)
    aggs = [(0, "sum", 1), (0, "count_valid", 1), (0, "min", 1), (1, "sum_of_squares", 1), (1, "sum", 1), (2, "sum", 1), (2, "max", 1), (2, "mean", 1)]
    _check([H(k)], [H(v), H(w), x], aggs, name="long buckets", capacity_log2=16, slice_log2=10)
    # the same rows with every value at the extreme of its digit: 65535+ addends of -32768 / 32767 per group and digit
    k2 = (np.arange(n) % 8).astype(np.int64)
    v2 = np.where(np.arange(n) % 2 == 0, np.int64(-32768), np.int64(32767)).astype(np.int64) * np.int64(65537)
    _check([H(k2)], [H(v2)], [(0, "sum", 1), (0, "count_valid", 1)], name="digit extremes", capacity_log2=12, slice_log2=8, expect=None)


@pytest.mark.parametrize("what", ["key", "value", "negative-value", "narrow"])
def test_packed_8_byte_records_and_the_row_that_does_not_fit(what):
    """SUM / COUNT of one integer column whose sampled key and value ranges fit 32 bits travel as 8-byte records {key32, value32}
    (groupby.cu: pack32). The sample only predicts: a single row outside the sample whose key or value needs more bits raises
    status[7] in level 1 and the call reruns with 16-byte records -- same answer either way."""
    n = 400_000
    rng = np.random.default_rng(5)
    k = rng.integers(0, 60_000, n).astype(np.int64)
    v = rng.integers(-1000, 1001, n).astype(np.int64)
    odd = 100_001  # (the estimator samples every 6th row of 400 000: row 100 001 is not one of them)
    if what == "key":
        k[odd] = (1 << 40) + 5
    elif what == "value":
        v[odd] = (1 << 40)
    elif what == "negative-value":
        v[odd] = -(1 << 33) - 7
    elif what == "narrow":
        v = v.astype(np.int16)  # narrow value types always fit; the sum still promotes to int64
    G = _check([H(k)], [H(v)], [(0, "sum", 1), (0, "count_valid", 1), (0, "count_all", 1)], name=f"pack32/{what}")
    assert G == len(np.unique(k))
