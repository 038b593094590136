"""Thread safety of the C ABI: 4 host threads, each with its own CUDA stream, run different groupby shapes at the same time and
repeatedly (SURVEY.md section 8 row b: the library is called from executor threads). Every thread's results are compared with the
oracle. What is shared between the calls: the library's memory pool (a mutex + per-stream block cache), the kernel-launch counter,
the per-thread error string -- and nothing else; a handle belongs to the call that built it."""
import threading

import numpy as np
import pytest

from oracle import oracle as orc

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
if not torch.cuda.is_available():  # pragma: no cover
    pytest.skip("needs a CUDA device", allow_module_level=True)

from tests import gpu_util  # noqa: E402

H = orc.HostColumn


def _shape(i, rng):
    """four different paths: tiny, shared-memory, bucketed (forced), generic with a two-word key"""
    n = 150_000 + 10_000 * i
    if i % 4 == 0:
        return [H(rng.integers(0, 5, n).astype(np.int32))], [H(rng.standard_normal(n))], [(0, "sum", 1), (0, "mean", 1)], None
    if i % 4 == 1:
        return [H(rng.integers(0, 3000, n).astype(np.int64))], [H(rng.integers(-50, 50, n).astype(np.int64))], [(0, "sum", 1), (0, "count_valid", 1)], None
    if i % 4 == 2:
        return ([H(rng.integers(0, 60_000, n).astype(np.int64))], [H(rng.integers(-50, 50, n).astype(np.int64), rng.random(n) > 0.1)],
                [(0, "sum", 1), (0, "max", 1), (0, "count_valid", 1)], {"force_path": 7})
    return ([H(rng.integers(-(1 << 62), 1 << 62, 500).astype(np.int64)[rng.integers(0, 500, n)]), H(rng.integers(0, 1 << 40, n).astype(np.int64) >> 30)],
            [H(rng.standard_normal(n))], [(0, "min", 1), (0, "variance", 1)], None)


def test_four_threads_four_streams():
    errors, done = [], [0] * 4
    start = threading.Barrier(4)

    def work(t):
        try:
            rng = np.random.default_rng(100 + t)
            with torch.cuda.stream(torch.cuda.Stream()):
                start.wait()
                for rep in range(6):
                    keys, vals, aggs, opts = _shape(t + rep, rng)
                    gpu_util.compare_with_oracle(keys, vals, aggs, options=opts, name=f"thread {t} rep {rep}")
                    done[t] += 1
        except BaseException as e:  # noqa: BLE001
            errors.append(e)
            start.abort()

    threads = [threading.Thread(target=work, args=(t,)) for t in range(4)]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    if errors:
        raise errors[0]
    assert done == [6] * 4
