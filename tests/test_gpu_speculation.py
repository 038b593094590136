"""One host synchronisation per call: bgb_groupby_build starts on the estimate remembered for the call's SHAPE while a fresh sample
runs beside the kernels (include/bgb.h, "tuning"; groupby.cu: EstimateEntry). No result may depend on the estimate: these tests feed
the same shape with data of very different cardinalities back to back, so every call but the first of a pair runs on a WRONG
remembered estimate, and compare every result with the oracle."""
import numpy as np
import pytest

from oracle import oracle as orc

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():  # pragma: no cover
    pytest.skip("needs a CUDA device", allow_module_level=True)

from cudf_b200 import _lib  # noqa: E402
from tests import gpu_util  # noqa: E402

H = orc.HostColumn
N = 300_000


@pytest.fixture()
def speculation_on():
    L = _lib.lib()
    prev = L.bgb_set_estimate_speculation(1)
    yield
    L.bgb_set_estimate_speculation(prev)


def _keys(rng, groups, sparse=False):
    k = rng.integers(0, groups, N).astype(np.int64)
    if sparse:
        k = (k * np.int64(0x9E3779B97F4A7C15 - (1 << 64))) ^ np.int64(0x5555555555555555)
    return k


def test_cardinality_swings_between_calls_of_one_shape(speculation_on):
    rng = np.random.default_rng(7)
    v = H(rng.integers(-1000, 1000, N).astype(np.int64))
    aggs = [(0, "sum", 1), (0, "count_valid", 1)]
    paths = []
    # tiny -> shared-memory direct -> sparse keys -> one group per row -> tiny again: each call inherits the previous call's estimate
    for groups, sparse in [(4, False), (4, False), (9000, False), (9000, False), (9000, True), (N * 50, True), (N * 50, True), (3, False), (3, False)]:
        G, path = gpu_util.compare_with_oracle([H(_keys(rng, groups, sparse))], [v], aggs, name=f"{groups} groups sparse={sparse}")
        paths.append(path)
    # the second call of a pair sees the right remembered estimate again (the drift check dropped the wrong one)
    assert paths[1].startswith("tiny") and paths[3].startswith("smem") and paths[-1].startswith("tiny"), paths


def test_float_moments_with_nulls_under_a_stale_estimate(speculation_on):
    rng = np.random.default_rng(8)
    aggs = [(0, "mean", 1), (0, "variance", 1), (0, "min", 1), (0, "max", 1), (0, "count_valid", 1)]
    for groups in (2, 40_000, 5, 150_000, 1200, 1200):
        k = H(rng.integers(0, groups, N).astype(np.int32), rng.random(N) > 0.05)
        v = H(rng.standard_normal(N), rng.random(N) > 0.2)
        gpu_util.compare_with_oracle([k], [v], aggs, include_null_keys=bool(groups % 2), name=f"{groups} groups")


def test_same_data_twice_takes_the_same_path_with_one_sync(speculation_on):
    rng = np.random.default_rng(9)
    k, v = H(_keys(rng, 5000)), H(rng.integers(-9, 9, N).astype(np.int64))
    first = gpu_util.run_gpu([k], [v], [(0, "sum", 1)])
    second = gpu_util.run_gpu([k], [v], [(0, "sum", 1)])
    assert first[3] == second[3]
    np.testing.assert_array_equal(first[2][0][0], second[2][0][0])
