"""GPU parity of keys whose packed image is wider than 128 bits: the group table holds the ROW NUMBER
of each group's first row and rows are compared column by column in the caller's key columns (device_utils.cuh:
find_or_insert_row), the scheme the reference uses for every key (/root/reference/cpp/src/groupby/hash/compute_groupby.cu:79-104).
Results are compared key-sorted with the oracle."""
import numpy as np
import pytest

from oracle import oracle as orc

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():  # pragma: no cover
    pytest.skip("needs a CUDA device", allow_module_level=True)

from tests import gpu_util  # noqa: E402

H = orc.HostColumn
N = 200_000
BIG = 1 << 62


def _wide(rng, n, groups, cols=4, dtype=np.int64):
    """`cols` full-range columns that together name `groups` distinct keys"""
    pool = [rng.integers(-BIG, BIG, groups).astype(dtype) for _ in range(cols)]
    pick = rng.integers(0, groups, n)
    return [p[pick] for p in pool], pick


def _check(keys, vals, aggs, include=False, name="", path_want="global+rowkeys"):
    G, path = gpu_util.compare_with_oracle(keys, vals, aggs, include_null_keys=include, name=name)
    assert path == path_want, f"{name}: path {path}"
    return G


@pytest.mark.parametrize("groups", [1, 37, 5000, 150_000])
def test_four_int64_keys_every_kind(groups):
    rng = np.random.default_rng(groups)
    ks, _ = _wide(rng, N, groups)
    v = H(rng.integers(-1000, 1000, N).astype(np.int64), rng.random(N) > 0.2)
    f = H(rng.standard_normal(N))
    aggs = [(0, "sum", 1), (0, "min", 1), (0, "max", 1), (0, "count_valid", 1), (0, "count_all", 1), (0, "mean", 1),
            (1, "sum", 1), (1, "variance", 1), (1, "std", 0), (0, "argmin", 1), (1, "argmax", 1), (0, "sum_of_squares", 1), (0, "product", 1)]
    # (one distinct key: every column's range is a single value, so the image repacks into one word after all)
    G = _check([H(k) for k in ks], [v, f], aggs, name=f"4 x int64, {groups} groups", path_want="global+repack" if groups == 1 else "global+rowkeys")
    assert G == len(np.unique(np.stack(ks), axis=1).T)


def test_columns_that_differ_only_in_the_last_column():
    # equal hashes of a prefix must not merge groups: rows agree in three columns and differ in the fourth
    rng = np.random.default_rng(3)
    a = np.full(N, 7, np.int64)
    b = np.full(N, -BIG, np.int64)
    c = np.full(N, BIG - 1, np.int64)
    d = rng.integers(-BIG, BIG, 900).astype(np.int64)[rng.integers(0, 900, N)]
    v = H(rng.integers(-5, 5, N).astype(np.int64))
    assert _check([H(a), H(b), H(c), H(d)], [v], [(0, "sum", 1), (0, "count_all", 1)], name="last column decides") == len(np.unique(d))


@pytest.mark.parametrize("include", [False, True])
def test_null_keys(include):
    rng = np.random.default_rng(11)
    ks, _ = _wide(rng, N, 300, cols=3)
    masks = [rng.random(N) > 0.05, None, rng.random(N) > 0.5]
    keys = [H(k, m) for k, m in zip(ks, masks)]
    v = H(rng.integers(-100, 100, N).astype(np.int32), rng.random(N) > 0.3)
    _check(keys, [v], [(0, "sum", 1), (0, "max", 1), (0, "count_valid", 1), (0, "mean", 1)], include=include, name=f"nulls include={include}")


def test_float_and_mixed_width_keys():
    # float64 keys are compared canonicalised (-0.0 == +0.0, NaN == NaN) like every other path
    rng = np.random.default_rng(5)
    f = rng.integers(-3, 3, N).astype(np.float64) * 0.5
    f[rng.random(N) < 0.05] = np.nan
    f[rng.random(N) < 0.05] = -0.0
    a = rng.integers(-BIG, BIG, 50).astype(np.int64)[rng.integers(0, 50, N)]
    b = rng.integers(0, 1 << 63, 7).astype(np.uint64)[rng.integers(0, 7, N)]
    c = rng.integers(-100, 100, N).astype(np.int16)
    keys = [H(f), H(a), H(b), H(c)]
    v = H((rng.random(N) + 0.5).astype(np.float32))  # (positive: a float32 sum that cancels has no relative tolerance to speak of)
    _check(keys, [v], [(0, "sum", 1), (0, "min", 1), (0, "count_valid", 1)], name="float + mixed widths")


def test_eight_key_columns_and_sum_overflow():
    rng = np.random.default_rng(8)
    ks, _ = _wide(rng, N, 1000, cols=8)
    v = H(rng.integers(-(1 << 62), 1 << 62, N).astype(np.int64))
    _check([H(k) for k in ks], [v], [(0, "sum", 1), (0, "count_all", 1)], name="8 x int64")


def test_forced_fast_path_is_refused():
    import cudf_b200 as cb
    rng = np.random.default_rng(9)
    ks, _ = _wide(rng, 10_000, 10, cols=3)
    with pytest.raises(cb.UnsupportedError):
        gpu_util.run_gpu([H(k) for k in ks], [H(np.ones(10_000, np.int64))], [(0, "sum", 1)], options={"force_path": 7})
