import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(autouse=True, scope="session")
def _deterministic_kernel_choice():
    """Tests assert which kernel family a call takes; with the estimate speculation of bgb_groupby_build (include/bgb.h, "tuning")
    that can depend on the calls before it. Off for the session; tests/test_gpu_speculation.py turns it on for its own cases."""
    try:
        import torch
        if not torch.cuda.is_available():
            yield
            return
        from cudf_b200 import _lib
        prev = _lib.lib().bgb_set_estimate_speculation(0)
    except Exception:  # the CPU-only run never loads a GPU context
        yield
        return
    yield
    _lib.lib().bgb_set_estimate_speculation(prev)
