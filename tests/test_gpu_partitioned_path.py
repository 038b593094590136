"""GPU parity of the partitioned (high-cardinality) path: radix partition by table region + aggregation over the
partition-ordered copy (cudf_b200/csrc/kernels_partition.cu), forced with options force_path=3, against the oracle."""
import numpy as np
import pytest

from oracle import oracle as orc

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():  # pragma: no cover
    pytest.skip("needs a CUDA device", allow_module_level=True)

from tests import gpu_util  # noqa: E402

H = orc.HostColumn


def _check(keys, vals, aggs, include=False, name="", capacity_log2=0):
    opts = {"force_path": 3}
    if capacity_log2:
        opts["capacity_log2"] = capacity_log2
    G, path = gpu_util.compare_with_oracle(keys, vals, aggs, include_null_keys=include, options=opts, name=name)
    assert path == "partitioned", f"{name}: expected the partitioned path, got {path}"
    return G


def test_config_b_shape_reduced():
    # int32 key, float64 value with 10% nulls, SUM + MIN + MAX (BASELINE config B at reduced N, high cardinality)
    n, g = 3_000_000, 400_000
    rows = np.arange(n, dtype=np.uint64)
    k = (orc.splitmix64(42, rows) % np.uint64(g)).astype(np.int32)
    v = (orc.splitmix64(43, rows) % np.uint64(1_000_000)).astype(np.float64)
    valid = (orc.splitmix64(44, rows) % np.uint64(10)) != 0
    _check([H(k)], [H(v, valid)], [(0, "sum", 1), (0, "min", 1), (0, "max", 1)], name="configB")


def test_config_a9_shape_reduced_high_cardinality():
    n, g = 4_000_000, 1_000_000
    rows = np.arange(n, dtype=np.uint64)
    k = (orc.splitmix64(42, rows) % np.uint64(g)).astype(np.int64)
    v = (orc.splitmix64(43, rows) % np.uint64(2001)).astype(np.int64) - 1000
    G = _check([H(k)], [H(v)], [(0, "sum", 1), (0, "count_valid", 1)], name="A9 high cardinality")
    assert G == len(np.unique(k))


@pytest.mark.parametrize("ktype", ["int8", "int16", "int32", "int64", "uint64", "float32", "float64"])
def test_key_types_nulls_and_all_aggs(ktype):
    rng = np.random.default_rng(21)
    n = 300_000
    keys = [H(rng.integers(0, 100 if ktype == "int8" else 20_000, n).astype(ktype), rng.random(n) > 0.04)]
    vals = [H(rng.integers(-100, 100, n).astype(np.int32), rng.random(n) > 0.1), H(rng.integers(0, 1000, n).astype(np.float64), rng.random(n) > 0.2)]
    aggs = [(0, "sum", 1), (0, "count_valid", 1), (0, "count_all", 1), (0, "min", 1), (1, "max", 1), (1, "mean", 1), (1, "variance", 1), (1, "sum", 1)]
    _check(keys, vals, aggs, name=f"part/{ktype}")
    _check(keys, vals, aggs, include=True, name=f"part/{ktype}/include")


def test_wide_keys_and_many_partitions():
    rng = np.random.default_rng(22)
    n = 500_000
    k1 = H(rng.integers(0, 300, n).astype(np.int64), rng.random(n) > 0.01)
    k2 = H(rng.integers(-50, 50, n).astype(np.int64))
    k3 = H(rng.integers(0, 4, n).astype(np.int16), rng.random(n) > 0.05)
    v = H(rng.standard_normal(n))
    w = H(rng.integers(0, 255, n).astype(np.uint8), rng.random(n) > 0.5)
    _check([k2, H(k1.data)], [v, w], [(0, "sum", 1), (0, "min", 1), (1, "sum", 1), (1, "max", 1)], name="2xint64 (128-bit key)")
    _check([k3, k1], [v, w], [(0, "mean", 1), (1, "count_valid", 1)], include=True, name="int16+int64 include")
    # a large table forces the maximum partition count (4096 regions)
    _check([H(k1.data)], [w], [(0, "sum", 1)], name="4096 partitions", capacity_log2=24)


def test_all_keys_null_and_tiny_inputs():
    n = 70_000
    k = H(np.arange(n, dtype=np.int32), np.zeros(n, dtype=bool))
    v = H(np.ones(n, dtype=np.int64))
    G, _, _, path = gpu_util.run_gpu([k], [v], [(0, "sum", 1)], options={"force_path": 3})
    assert G == 0
    _check([H(np.array([5, 5, 7], dtype=np.int64))], [H(np.array([1, 2, 3], dtype=np.int64))], [(0, "sum", 1), (0, "count_all", 1)], name="tiny")


@pytest.mark.parametrize("force_path", [3, 7])
def test_argmin_argmax_keep_source_row_numbers(force_path):
    # ARGMIN/ARGMAX report row numbers of the INPUT: a partition- or bucket-ordered copy cannot serve them, the call must
    # take a path that sees the original rows
    rng = np.random.default_rng(33)
    n = 200_000
    k = H(rng.integers(0, 30_000, n).astype(np.int64))
    v = H(rng.permutation(n).astype(np.int32))  # unique values: no ties
    G, path = gpu_util.compare_with_oracle([k], [v], [(0, "argmin", 1), (0, "argmax", 1), (0, "min", 1)], options={"force_path": force_path},
                                           name=f"arg/{force_path}")
    assert path == "global", path
