"""Runs the C++ host-layer test binary (include/cudf_b200/groupby.hpp over libbgb.so; tests/cpp/test_groupby_hpp.cpp,
built by __graft_entry__.build()) on the GPU box."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "tests", "cpp", "_build", "test_groupby_hpp")


@pytest.mark.gpu
def test_cpp_groupby_layer():
    assert os.path.exists(BIN), "tests/cpp/_build/test_groupby_hpp is missing: run __graft_entry__.build()"
    p = subprocess.run([BIN], capture_output=True, text=True, timeout=120)
    assert p.returncode == 0, p.stdout + p.stderr
    assert "all C++ host-layer checks passed" in p.stdout


def test_cpp_layer_refuses_to_run_without_a_gpu():
    """No CPU fallback: without a device the binary exits 2 (skipped when a GPU is present or the binary is not built)."""
    import torch
    if torch.cuda.is_available() or not os.path.exists(BIN):
        pytest.skip("needs the built binary and no GPU")
    p = subprocess.run([BIN], capture_output=True, text=True, timeout=60)
    assert p.returncode == 2 and "no CPU fallback" in p.stdout
