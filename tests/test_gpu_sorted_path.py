"""GPU parity of the tile-sorting mid-cardinality kernel (cudf_b200/csrc/kernels_sorted.cu): dense integer key range of a few
thousand ids, any accumulator set but ARGMIN/ARGMAX; rows of a tile are counting-sorted by group in shared memory and every
thread folds the groups it owns without atomics. Forced with options force_path=8, and taken automatically for shapes the
SUM/COUNT shared-memory kernel does not serve. Compared key-sorted with the oracle (floats within 1e-9 relative)."""
import numpy as np
import pytest

from oracle import oracle as orc

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():  # pragma: no cover
    pytest.skip("needs a CUDA device", allow_module_level=True)

from tests import gpu_util  # noqa: E402

H = orc.HostColumn
N = 400_000


def _check(keys, vals, aggs, include=False, name="", force=True, expect="sorted"):
    opts = {"force_path": 8} if force else None
    G, path = gpu_util.compare_with_oracle(keys, vals, aggs, include_null_keys=include, options=opts, name=name)
    if expect:
        assert path == expect, f"{name}: expected the {expect} path, got {path}"
    return G


def test_baseline_config_1_shape_is_repacked_and_sorted():
    # BASELINE configs[0]: df.groupby(['A','B']).mean(), int64 keys in [0, 32), float64 values -> 1024 groups
    rng = np.random.default_rng(42)
    a, b = rng.integers(0, 32, N).astype(np.int64), rng.integers(0, 32, N).astype(np.int64)
    c, d = rng.standard_normal(N), rng.standard_normal(N)
    G = _check([H(a), H(b)], [H(c), H(d)], [(0, "mean", 1), (1, "mean", 1)], name="cfg1", force=False, expect="sorted+repack")
    assert G == 1024


@pytest.mark.parametrize("ktype", ["int8", "int16", "int32", "int64", "uint8", "uint32", "uint64"])
def test_key_types_every_accumulator_kind(ktype):
    rng = np.random.default_rng(3)
    lo = -60 if ktype.startswith("int") else 0
    k = H(rng.integers(lo, lo + 120, N).astype(ktype))
    vals = [H(rng.integers(-100, 100, N).astype(np.int32), rng.random(N) > 0.1), H(rng.integers(0, 1000, N).astype(np.float64), rng.random(N) > 0.2),
            H((rng.random(N) + 0.5).astype(np.float32)), H(rng.integers(-3, 4, N).astype(np.int64))]
    aggs = [(0, "sum", 1), (0, "count_valid", 1), (0, "count_all", 1), (0, "min", 1), (0, "sum_of_squares", 1), (1, "max", 1), (1, "mean", 1),
            (1, "variance", 1), (1, "std", 0), (1, "sum", 1), (2, "sum", 1), (2, "min", 1), (2, "mean", 1), (3, "sum", 1), (3, "max", 1)]
    assert _check([k], vals, aggs, name=f"sorted/{ktype}") == 120


def test_products_stay_exact_for_small_factors():
    rng = np.random.default_rng(4)
    n = 70_000
    k = H(rng.integers(0, 3000, n).astype(np.int32))
    v = H(rng.choice(np.array([-1, 1, 2], dtype=np.int64), n), rng.random(n) > 0.3)
    w = H(rng.choice(np.array([0.5, 1.0, 2.0, -1.0]), n))
    _check([k], [v, w], [(0, "product", 1), (1, "product", 1), (0, "count_valid", 1)], name="products")


@pytest.mark.parametrize("include", [False, True])
def test_masked_keys_several_key_columns(include):
    rng = np.random.default_rng(5)
    k1 = H(rng.integers(0, 50, N).astype(np.int16), rng.random(N) > 0.05)
    k2 = H(rng.integers(0, 7, N).astype(np.int8))
    v = H(rng.standard_normal(N), rng.random(N) > 0.1)
    w = H(rng.integers(-1000, 1000, N).astype(np.int64))
    # int16 + int8 fields leave holes in the packed image (a 24-bit range): not dense, so the ids come from the CTA's hash table
    _check([k1, k2], [v, w], [(0, "mean", 1), (0, "min", 1), (1, "sum", 1), (1, "count_all", 1)], include=include, name="2 keys", expect="sorted-hashed")


def test_single_masked_key_column_exclude_and_include():
    rng = np.random.default_rng(6)
    k = H(rng.integers(-500, 500, N).astype(np.int32), rng.random(N) > 0.1)
    v = H(rng.standard_normal(N))
    _check([k], [v], [(0, "sum", 1), (0, "mean", 1), (0, "max", 1)], include=False, name="masked key/exclude")
    # INCLUDE puts the null flag at bit 32 of the image: the range no longer fits, the automatic plan takes another path
    G, path = gpu_util.compare_with_oracle([k], [v], [(0, "sum", 1)], include_null_keys=True, name="masked key/include")
    assert G == 1001 and path == "sorted-hashed", path


def test_skewed_keys_heavy_groups_take_a_warp():
    # one key holds 60 % of the rows, another 10 %: their tile segments exceed 64 rows and are folded by whole warps
    rng = np.random.default_rng(7)
    k = rng.integers(0, 2000, N).astype(np.int64)
    u = rng.random(N)
    k[u < 0.6] = 77
    k[(u >= 0.6) & (u < 0.7)] = 1999
    v = H(rng.integers(0, 1000, N).astype(np.float64), rng.random(N) > 0.25)
    w = H(rng.integers(-(1 << 40), 1 << 40, N).astype(np.int64))
    aggs = [(0, "sum", 1), (0, "count_valid", 1), (0, "min", 1), (0, "max", 1), (0, "variance", 1), (1, "sum", 1), (1, "sum_of_squares", 1), (1, "count_all", 1)]
    _check([H(k)], [v, w], aggs, name="skew")
    # every row in one group
    _check([H(np.full(N, 5, dtype=np.int32))], [v, w], aggs, name="one group", expect=None)


def test_keys_outside_the_sampled_range_take_the_bypass():
    # 65536 strided sample rows see [0, 1000); a few rows far outside must still land in the right groups
    rng = np.random.default_rng(8)
    n = 2_000_000
    k = rng.integers(0, 1000, n).astype(np.int64)
    far = rng.choice(n // 30 - 1, 37, replace=False) * 30 + 7  # the estimator samples rows 0, 30, 60, ...: these are never seen
    k[far] = rng.integers(-(1 << 50), 1 << 50, 37)
    k[far[:3]] = -1  # image of the EMPTY sentinel
    v = H(rng.standard_normal(n), rng.random(n) > 0.1)
    G = _check([H(k)], [v], [(0, "sum", 1), (0, "count_valid", 1), (0, "count_all", 1), (0, "min", 1)], name="bypass")
    assert G == len(np.unique(k))


def test_largest_group_range_and_ragged_tail():
    rng = np.random.default_rng(9)
    n = 300_001  # not a multiple of the tile
    k = H(rng.integers(0, 3900, n).astype(np.int32))
    v = H(rng.standard_normal(n))
    _check([k], [v], [(0, "mean", 1), (0, "std", 1)], name="3900 ids")


def test_automatic_choice_float_mean_dense_single_key():
    rng = np.random.default_rng(10)
    k = H(rng.integers(100, 2100, N).astype(np.int64))
    v = H(rng.standard_normal(N), rng.random(N) > 0.1)
    _check([k], [v], [(0, "mean", 1), (0, "max", 1)], name="auto", force=False)
    # int64 SUM + COUNT stays on the split-lane shared-memory kernel
    w = H(rng.integers(-9, 9, N).astype(np.int64))
    G, path = gpu_util.compare_with_oracle([k], [w], [(0, "sum", 1), (0, "count_valid", 1)], name="auto/sumcount")
    assert path.startswith("smem"), path


def test_argmin_is_not_served():
    import cudf_b200 as cb
    rng = np.random.default_rng(11)
    k = H(rng.integers(0, 100, N).astype(np.int32))
    v = H(rng.standard_normal(N))
    with pytest.raises(cb.UnsupportedError):
        gpu_util.run_gpu([k], [v], [(0, "argmin", 1)], options={"force_path": 8})


@pytest.mark.parametrize("include", [False, True])
def test_repacked_nullable_key_columns(include):
    # three nullable key columns of different widths whose exact ranges need 7 bits (+ 3 null flags under INCLUDE): repacked
    # into one dense 64-bit image, which the tile-sorting kernel indexes directly
    rng = np.random.default_rng(12)
    k1 = H(rng.integers(-7_000_000_000, -6_999_999_992, N).astype(np.int64), rng.random(N) > 0.05)
    k2 = H(rng.integers(-3, 5, N).astype(np.int64), rng.random(N) > 0.1)
    k3 = H(rng.integers(200, 202, N).astype(np.uint8), rng.random(N) > 0.02)
    v = H(rng.standard_normal(N), rng.random(N) > 0.1)
    w = H(rng.integers(-50, 50, N).astype(np.int16))
    aggs = [(0, "mean", 1), (0, "min", 1), (1, "sum", 1), (1, "max", 1)]
    _check([k1, k2, k3], [v, w], aggs, include=include, name=f"repacked/include={include}", force=False, expect="sorted+repack")


def test_repacked_keys_all_null_column():
    rng = np.random.default_rng(13)
    k1 = H(rng.integers(0, 100, N).astype(np.int64))
    k2 = H(np.zeros(N, dtype=np.int64), np.zeros(N, dtype=bool))  # every element null
    v = H(rng.standard_normal(N))
    G = _check([k1, k2], [v], [(0, "mean", 1)], include=True, name="repacked/all-null column", force=False, expect="sorted+repack")
    assert G == 100
    G, path = gpu_util.compare_with_oracle([k1, k2], [v], [(0, "mean", 1)], include_null_keys=False, name="repacked/all-null column/exclude")
    assert G == 0


# ---- hashed variant: the group id is the key's slot in a small shared-memory hash table of the CTA

def test_hashed_sparse_integer_keys_automatic():
    rng = np.random.default_rng(20)
    pool = rng.integers(-(1 << 60), 1 << 60, 500).astype(np.int64)
    k = H(pool[rng.integers(0, 500, N)])
    v = H(rng.standard_normal(N), rng.random(N) > 0.1)
    w = H(rng.integers(-1000, 1000, N).astype(np.int32))
    aggs = [(0, "mean", 1), (0, "std", 1), (0, "min", 1), (1, "sum", 1), (1, "max", 1)]
    assert _check([k], [v, w], aggs, name="hashed/sparse", force=False, expect="sorted-hashed") == 500


@pytest.mark.parametrize("include", [False, True])
def test_hashed_float_keys_with_nan_and_negative_zero(include):
    # float32 keys: -0.0 == +0.0 and every NaN payload is one key (the image is canonicalised before it is hashed)
    rng = np.random.default_rng(21)
    pool = np.concatenate([rng.standard_normal(300), [0.0, -0.0, np.nan, np.inf, -np.inf]]).astype(np.float32)
    kd = pool[rng.integers(0, len(pool), N)]
    other_nan = np.frombuffer(np.uint32(0xFFC00001).tobytes(), dtype=np.float32)[0]
    kd[np.isnan(kd) & (rng.random(N) < 0.5)] = other_nan
    k = H(kd, rng.random(N) > 0.03)
    v = H(rng.integers(0, 1000, N).astype(np.float64))
    _check([k], [v], [(0, "sum", 1), (0, "max", 1), (0, "count_all", 1)], include=include, name=f"hashed/float keys/{include}", force=False, expect="sorted-hashed")


@pytest.mark.parametrize("include", [False, True])
def test_hashed_two_key_columns_with_masks(include):
    rng = np.random.default_rng(22)
    k1 = H(rng.integers(0, 50, N).astype(np.int16), rng.random(N) > 0.05)
    k2 = H(rng.integers(0, 7, N).astype(np.int8), rng.random(N) > 0.5)
    v = H(rng.standard_normal(N), rng.random(N) > 0.1)
    w = H(rng.integers(-1000, 1000, N).astype(np.int64))
    _check([k1, k2], [v, w], [(0, "mean", 1), (0, "min", 1), (1, "sum", 1), (1, "count_all", 1)], include=include, name="hashed/2 keys", expect="sorted-hashed")


def test_hashed_more_groups_than_slots_overflow_to_the_hbm_table():
    # 20 000 distinct sparse keys against at most 4096 slots per CTA: most rows bypass to the HBM table; still exact
    rng = np.random.default_rng(23)
    pool = rng.integers(-(1 << 60), 1 << 60, 20_000).astype(np.int64)
    pool[:2] = [-1, 0]  # -1: image of the EMPTY sentinel
    k = H(pool[rng.integers(0, 20_000, N)])
    v = H(rng.integers(-100, 100, N).astype(np.int64), rng.random(N) > 0.2)
    G = _check([k], [v], [(0, "sum", 1), (0, "count_valid", 1), (0, "min", 1), (0, "sum_of_squares", 1)], name="hashed/overflow", expect="sorted-hashed")
    assert G == len(np.unique(k.data))


def test_hashed_skewed_keys():
    rng = np.random.default_rng(24)
    pool = rng.integers(-(1 << 40), 1 << 40, 500).astype(np.int64)
    kd = pool[rng.integers(0, 500, N)]
    kd[rng.random(N) < 0.7] = pool[3]
    v = H(rng.integers(0, 1000, N).astype(np.float64), rng.random(N) > 0.25)
    _check([H(kd)], [v], [(0, "sum", 1), (0, "variance", 1), (0, "max", 1)], name="hashed/skew", force=False, expect="sorted-hashed")
