"""GPU parity of the tiny-cardinality kernel (cudf_b200/csrc/kernels_tiny.cu): per-thread private shared-memory
accumulators for a handful of groups. Forced with options force_path=5, plus the automatic choice and the fall-back
when the data holds more groups than the kernel has rows for. float64 SUM/MEAN/VAR within 1e-9 relative."""
import numpy as np
import pytest

from oracle import oracle as orc

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():  # pragma: no cover
    pytest.skip("needs a CUDA device", allow_module_level=True)

from tests import gpu_util  # noqa: E402

H = orc.HostColumn
Q1_AGGS = [(0, "mean", 1), (0, "variance", 1), (0, "count_valid", 1), (1, "mean", 1), (1, "variance", 1), (1, "count_valid", 1)]


def _q1_columns(n, seed=42):
    rows = np.arange(n, dtype=np.uint64)
    x = (orc.splitmix64(seed, rows) & np.uint64(0xFFFF)).astype(np.int64)
    cum = np.array([int(0.247 * 65536), int(0.253 * 65536), int(0.753 * 65536), 65536])
    cat = np.searchsorted(cum, x, side="right").clip(0, 3)
    rf = np.array([0, 1, 1, 2], dtype=np.int8)[cat]
    ls = np.array([0, 0, 1, 0], dtype=np.int8)[cat]
    q = (orc.splitmix64(seed + 1, rows) % np.uint64(50)).astype(np.float64) + 1.0
    p = (orc.splitmix64(seed + 2, rows) % np.uint64(104_100)).astype(np.float64) + 900.0
    return [H(rf), H(ls)], [H(q), H(p)]


def test_config_q1_shape_reduced_forced_and_auto():
    keys, vals = _q1_columns(3_000_000)
    G, path = gpu_util.compare_with_oracle(keys, vals, Q1_AGGS, options={"force_path": 5}, name="Q1/forced")
    assert path == "tiny" and G == 4
    G, path = gpu_util.compare_with_oracle(keys, vals, Q1_AGGS, name="Q1/auto")
    assert path == "tiny-reg" and G == 4  # 4 groups, two float64 columns: accumulators live in registers
    G, path = gpu_util.compare_with_oracle(keys, vals, Q1_AGGS, options={"force_path": 6}, name="Q1/forced-reg8")
    assert path == "tiny-reg" and G == 4


@pytest.mark.parametrize("vtype", ["int64", "uint64", "float64"])
@pytest.mark.parametrize("ncols", [1, 2])
def test_register_variant_with_nulls(vtype, ncols):
    rng = np.random.default_rng(41 + ncols)
    n = 500_001  # odd: the last pair of the last tile is half outside the column
    keys = [H(rng.integers(0, 7, n).astype(np.int16), rng.random(n) > 0.05), H(rng.integers(0, 1, n).astype(np.int16))]
    vals = [H(rng.integers(0, 1000, n).astype(vtype), rng.random(n) > 0.2) for _ in range(ncols)]
    aggs = []
    for c in range(ncols):
        aggs += [(c, "sum", 1), (c, "count_valid", 1), (c, "mean", 1), (c, "variance", 1), (c, "sum_of_squares", 1)]
    aggs.append((0, "count_all", 1))
    for include in (False, True):
        G, path = gpu_util.compare_with_oracle(keys, vals, aggs, include_null_keys=include, options={"force_path": 6},
                                               name=f"tiny-reg/{vtype}/{ncols}/{include}")
        assert path == "tiny-reg" and G == (8 if include else 7)
    G, path = gpu_util.compare_with_oracle(keys, vals, aggs, name=f"tiny-reg/{vtype}/{ncols}/auto")
    assert path == "tiny-reg"


def test_register_variant_more_than_8_groups_falls_back():
    rng = np.random.default_rng(43)
    n = 300_000
    k = rng.integers(0, 12, n).astype(np.int64)
    keys, vals = [H(k)], [H(rng.standard_normal(n))]
    aggs = [(0, "sum", 1), (0, "variance", 1), (0, "count_valid", 1)]
    G, path = gpu_util.compare_with_oracle(keys, vals, aggs, options={"force_path": 6}, name="tiny-reg/overflow")
    assert path != "tiny-reg" and G == 12
    # negative zero and NaN values go through the same adds as the reference's
    v = rng.standard_normal(n)
    v[::7] = -0.0
    k2 = rng.integers(0, 3, n).astype(np.int64)
    gpu_util.compare_with_oracle([H(k2)], [H(v)], [(0, "sum", 1), (0, "mean", 1)], options={"force_path": 6}, name="tiny-reg/negzero")


@pytest.mark.parametrize("ktype", ["int8", "int16", "int32", "int64", "uint32"])
def test_register_variant_key_widths(ktype):
    rng = np.random.default_rng(45)
    n = 200_003
    keys = [H(rng.integers(0, 5, n).astype(ktype))]
    vals = [H(rng.standard_normal(n), rng.random(n) > 0.1)]
    G, path = gpu_util.compare_with_oracle(keys, vals, [(0, "sum", 1), (0, "std", 1), (0, "count_valid", 1), (0, "count_all", 1)],
                                           options={"force_path": 6}, name=f"tiny-reg/key {ktype}")
    assert path == "tiny-reg" and G == 5


def test_register_variant_declines_min_max_and_narrow_values():
    import cudf_b200 as cb
    rng = np.random.default_rng(44)
    n = 100_000
    keys = [H(rng.integers(0, 3, n).astype(np.int32))]
    with pytest.raises(cb.UnsupportedError):  # key columns of different widths
        gpu_util.run_gpu([keys[0], H(rng.integers(0, 2, n).astype(np.int8))], [H(rng.standard_normal(n))], [(0, "sum", 1)], options={"force_path": 6})
    with pytest.raises(cb.UnsupportedError):
        gpu_util.run_gpu(keys, [H(rng.standard_normal(n))], [(0, "min", 1)], options={"force_path": 6})
    with pytest.raises(cb.UnsupportedError):
        gpu_util.run_gpu(keys, [H(rng.integers(0, 9, n).astype(np.int32))], [(0, "sum", 1)], options={"force_path": 6})
    # without forcing, the same calls are served by the shared-memory variant
    G, path = gpu_util.compare_with_oracle(keys, [H(rng.standard_normal(n))], [(0, "min", 1), (0, "sum", 1)], name="tiny/minmax-auto")
    assert path == "tiny"


@pytest.mark.parametrize("vtype", ["int8", "int32", "int64", "uint16", "float32", "float64"])
def test_all_supported_accumulators_with_nulls(vtype):
    rng = np.random.default_rng(31)
    n = 400_000
    keys = [H(rng.integers(0, 5, n).astype(np.int32), rng.random(n) > 0.05)]
    # float32 sums accumulate in float32 like the reference's: small integers keep them exact whatever the order
    vals = [H(rng.integers(0, 10 if vtype == "float32" else 100, n).astype(vtype), rng.random(n) > 0.15)]
    aggs = [(0, "sum", 1), (0, "count_valid", 1), (0, "count_all", 1), (0, "min", 1), (0, "max", 1), (0, "mean", 1), (0, "std", 1),
            (0, "sum_of_squares", 1)]
    for include in (False, True):
        G, path = gpu_util.compare_with_oracle(keys, vals, aggs, include_null_keys=include, options={"force_path": 5}, name=f"tiny/{vtype}/{include}")
        assert path == "tiny"


def test_more_groups_than_rows_falls_back_to_the_general_path():
    rng = np.random.default_rng(32)
    n = 300_000
    k = rng.integers(0, 3, n).astype(np.int64)
    k[rng.choice(n, 200, replace=False)] = rng.integers(1000, 1200, 200)  # ~170 rare keys the 65536-row sample mostly misses
    keys, vals = [H(k)], [H(rng.standard_normal(n)), H(rng.integers(0, 9, n).astype(np.int32))]
    aggs = [(0, "sum", 1), (0, "min", 1), (0, "max", 1), (0, "variance", 1), (1, "sum", 1), (1, "count_valid", 1)]
    G, path = gpu_util.compare_with_oracle(keys, vals, aggs, options={"force_path": 5}, name="tiny/overflow")
    assert path != "tiny" and G == len(np.unique(k))
    gpu_util.compare_with_oracle(keys, vals, aggs, name="tiny/overflow/auto")


def test_register_variant_all_ones_key_image():
    rng = np.random.default_rng(46)
    n = 200_000
    k = rng.integers(-2, 2, n).astype(np.int64)  # -1 packs to the all-ones image, which the kernels reserve: general path
    v = rng.integers(-1000, 1001, n).astype(np.int64)
    for fp in (0, 6):
        G, path = gpu_util.compare_with_oracle([H(k)], [H(v)], [(0, "sum", 1), (0, "count_valid", 1)], options={"force_path": fp}, name=f"tiny-reg/all-ones/{fp}")
        assert G == 4 and path != "tiny-reg"


def test_all_ones_key_and_single_group():
    k = np.full(100_000, -1, dtype=np.int64)  # packed image == EMPTY: must still be aggregated (general path)
    v = np.arange(100_000, dtype=np.int64)
    gpu_util.compare_with_oracle([H(k)], [H(v)], [(0, "sum", 1), (0, "max", 1)], options={"force_path": 5}, name="tiny/all-ones")
    k[:] = 7
    G, path = gpu_util.compare_with_oracle([H(k)], [H(v.astype(np.float64))], [(0, "mean", 1), (0, "std", 1)], options={"force_path": 5}, name="tiny/one group")
    assert path == "tiny" and G == 1
