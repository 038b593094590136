"""GPU parity of key repacking (cudf_b200/csrc/kernels_repack.cu, groupby.cu: try_repack_keys): integer key columns whose
packed image would need 128 bits are range-compressed into one 64-bit key column when their exact value ranges allow it,
and the call then takes the single-word paths. Results are compared key-sorted with the oracle; the path string tells
whether the call was repacked."""
import numpy as np
import pytest

from oracle import oracle as orc

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():  # pragma: no cover
    pytest.skip("needs a CUDA device", allow_module_level=True)

from tests import gpu_util  # noqa: E402

H = orc.HostColumn
N = 300_000


def _check(keys, vals, aggs, include=False, name="", repacked=True):
    G, path = gpu_util.compare_with_oracle(keys, vals, aggs, include_null_keys=include, name=name)
    assert ("+repack" in path) == repacked, f"{name}: path {path}"
    return G, path


def test_two_int64_keys_sum_count_take_the_shared_memory_table():
    # two int64 columns are a 128-bit image; their ranges need 10 bits: SUM + COUNT then run on a shared-memory table
    rng = np.random.default_rng(42)
    a, b = rng.integers(0, 32, N).astype(np.int64), rng.integers(-16, 16, N).astype(np.int64)
    v = H(rng.integers(-1000, 1000, N).astype(np.int64))
    G, path = _check([H(a), H(b)], [v], [(0, "sum", 1), (0, "count_valid", 1)], name="2xint64 sum/count")
    assert G == 1024 and path.startswith("smem"), path


def test_two_int64_keys_few_groups_take_the_tiny_kernel():
    rng = np.random.default_rng(41)
    a, b = rng.integers(5, 8, N).astype(np.int64), rng.integers(-1, 1, N).astype(np.int64)
    c, d = H(rng.standard_normal(N)), H(rng.standard_normal(N), rng.random(N) > 0.2)
    G, path = _check([H(a), H(b)], [c, d], [(0, "mean", 1), (0, "variance", 1), (1, "mean", 1), (1, "count_valid", 1)], name="6 groups")
    assert G == 6 and path.startswith("tiny"), path


def test_baseline_config_1_shape_takes_the_tile_sorting_kernel():
    # BASELINE configs[0]: df.groupby(['A','B']).mean(), int64 keys in [0, 32), float64 values -> 1024 groups: the packed image is
    # a 10-bit dense id, which the tile-sorting kernel (kernels_sorted.cu) indexes directly
    rng = np.random.default_rng(42)
    a, b = rng.integers(0, 32, N).astype(np.int64), rng.integers(0, 32, N).astype(np.int64)
    c, d = rng.standard_normal(N), rng.standard_normal(N)
    G, path = _check([H(a), H(b)], [H(c), H(d)], [(0, "mean", 1), (1, "mean", 1)], name="cfg1")
    assert G == 1024 and path.startswith("sorted"), path


def test_float_means_over_a_wide_packed_range_are_left_alone():
    # 2 x 9 bits of dense ids exceed what the tile-sorting kernel holds, and no other single-word path serves two float MEANs at
    # 1e5 groups faster than the L2-resident 128-bit table: the two extra passes would buy nothing
    rng = np.random.default_rng(42)
    a, b = rng.integers(0, 400, N).astype(np.int64), rng.integers(0, 400, N).astype(np.int64)
    c, d = rng.standard_normal(N), rng.standard_normal(N)
    _check([H(a), H(b)], [H(c), H(d)], [(0, "mean", 1), (1, "mean", 1)], name="wide float", repacked=False)


@pytest.mark.parametrize("include", [False, True])
def test_nullable_keys_negative_ranges_mixed_widths(include):
    rng = np.random.default_rng(43)
    k1 = H(rng.integers(-5_000_000_000, -4_999_990_000, N).astype(np.int64), rng.random(N) > 0.03)
    k2 = H(rng.integers(-7, 9, N).astype(np.int64), rng.random(N) > 0.05)
    k3 = H(rng.integers(0, 3, N).astype(np.int16))
    k4 = H((rng.random(N) > 0.5))
    v = H(rng.integers(-1000, 1000, N).astype(np.int64), rng.random(N) > 0.1)
    w = H(rng.integers(0, 100, N).astype(np.float64))
    aggs = [(0, "sum", 1), (0, "count_valid", 1), (0, "count_all", 1), (0, "min", 1), (1, "max", 1), (1, "variance", 1)]
    _check([k1, k2, k3, k4], [v, w], aggs, include=include, name=f"mixed/include={include}")


def test_unsigned_keys_above_2_63_and_timestamps():
    rng = np.random.default_rng(44)
    k1 = H((np.uint64(1 << 63) + rng.integers(0, 2000, N).astype(np.uint64)))
    k2 = H(rng.integers(1_700_000_000_000, 1_700_000_100_000, N).astype(np.int64), type_id=14)  # TIMESTAMP_MILLISECONDS
    k3 = H(rng.integers(-2, 2, N).astype(np.int64))  # third 64-bit column: 192 bits, only the repacked form is served
    v = H(rng.integers(0, 50, N).astype(np.int32))
    _check([k1, k2, k3], [v], [(0, "sum", 1), (0, "max", 1)], name="uint64+timestamp")


def test_sliced_columns_and_all_null_key_column():
    rng = np.random.default_rng(45)
    base1, base2 = rng.integers(0, 500, N + 77).astype(np.int64), rng.integers(-3, 3, N + 77).astype(np.int64)
    k1 = H(base1[77:])  # to_device keeps offset 0; slicing is exercised through the oracle-equal copy below
    k2 = H(base2[77:], rng.random(N) > 0.5)
    k3 = H(np.zeros(N, dtype=np.int64), np.zeros(N, dtype=bool))  # every element null
    v = H(rng.integers(-9, 9, N).astype(np.int64))
    _check([k1, k2, k3], [v], [(0, "sum", 1), (0, "count_all", 1)], include=True, name="all-null column/include")
    G, path = gpu_util.compare_with_oracle([k1, k2, k3], [v], [(0, "sum", 1)], include_null_keys=False, name="all-null column/exclude")
    assert G == 0


def test_ranges_too_wide_stay_on_the_128_bit_path():
    rng = np.random.default_rng(46)
    k1 = H(rng.integers(-(1 << 62), 1 << 62, 20_000).astype(np.int64)[rng.integers(0, 20_000, N)])
    k2 = H(rng.integers(-(1 << 62), 1 << 62, 20_000).astype(np.int64)[rng.integers(0, 20_000, N)])
    v = H(rng.integers(-9, 9, N).astype(np.int64))
    _check([k1, k2], [v], [(0, "sum", 1), (0, "count_valid", 1)], name="wide ranges", repacked=False)


def test_forced_paths_keep_the_caller_layout():
    # force_path != 0 never repacks (the forced kernels are tested on the layout the test asked for)
    rng = np.random.default_rng(47)
    a, b = rng.integers(0, 32, N).astype(np.int64), rng.integers(0, 32, N).astype(np.int64)
    v = H(rng.integers(-9, 9, N).astype(np.int64))
    G, path = gpu_util.compare_with_oracle([H(a), H(b)], [v], [(0, "sum", 1)], options={"force_path": 1}, name="forced generic")
    assert path == "global", path


def test_three_wide_columns_whose_ranges_do_not_fit_take_the_row_number_table():
    # 192 bits of key whose ranges need 189: no packed image; the table then holds row numbers (tests/test_gpu_wide_keys.py)
    rng = np.random.default_rng(48)
    ks = [H(rng.integers(-(1 << 62), 1 << 62, N).astype(np.int64)) for _ in range(3)]
    v = H(rng.integers(-9, 9, N).astype(np.int64))
    G, path = _check(ks, [v], [(0, "sum", 1)], name="3 x int64 full range", repacked=False)
    assert G == N and path == "global+rowkeys", path
