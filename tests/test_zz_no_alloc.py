"""bgb_bytes_to_bitmask_noalloc (include/bgb.h): the allocation-free variant of bgb_bytes_to_bitmask, whose scratch is a
file-scope __device__ array. No reference counterpart. Checked through the C
ABI against the existing entry point, bit for bit, on two calls in a row (the second call meets the first call's scratch
contents). Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _both(L, torch, valid_np):
    n = int(valid_np.size)
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    b = torch.from_numpy(valid_np.astype(np.uint8)).cuda()
    words = (n + 31) // 32
    out = []
    for fn in (L.bgb_bytes_to_bitmask, L.bgb_bytes_to_bitmask_noalloc):
        m = torch.full((max(words, 1),), -1, dtype=torch.int32, device="cuda")
        nc = C.c_int32(-7)
        rc = fn(C.c_void_p(b.data_ptr()), n, C.c_void_p(m.data_ptr()), C.byref(nc), s)
        assert rc == 0, L.bgb_last_error().decode()
        torch.cuda.synchronize()
        out.append((m[:words].cpu().numpy().view(np.uint32).copy(), nc.value))
    return out


def test_noalloc_matches_existing_entry_point_twice_in_a_row():
    import torch

    from cudf_b200 import _lib
    L = _lib.lib()
    rng = np.random.default_rng(5)
    # two calls in a row with different sizes and null densities: the second call starts from the first call's scratch
    for n, p in [(1_000_003, 0.3), (70_001, 0.9), (8192 * 3, 0.5), (33, 0.1), (1, 0.0)]:
        valid = rng.random(n) > p
        (w0, c0), (w1, c1) = _both(L, torch, valid)
        np.testing.assert_array_equal(w1, w0)
        assert c1 == c0 == int(n - np.count_nonzero(valid))
        # and against numpy directly
        bits = np.zeros(((n + 31) // 32) * 32, dtype=np.uint8)
        bits[:n] = valid
        np.testing.assert_array_equal(w1, np.packbits(bits, bitorder="little").view(np.uint32))


def test_noalloc_async_without_host_count_and_limit():
    import torch

    from cudf_b200 import _lib
    L = _lib.lib()
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    n = 300_000
    valid = np.random.default_rng(6).random(n) > 0.5
    b = torch.from_numpy(valid.astype(np.uint8)).cuda()
    m = torch.zeros((n + 31) // 32, dtype=torch.int32, device="cuda")
    assert L.bgb_bytes_to_bitmask_noalloc(C.c_void_p(b.data_ptr()), n, C.c_void_p(m.data_ptr()), None, s) == 0
    torch.cuda.synchronize()
    bits = np.zeros(((n + 31) // 32) * 32, dtype=np.uint8)
    bits[:n] = valid
    np.testing.assert_array_equal(m.cpu().numpy().view(np.uint32), np.packbits(bits, bitorder="little").view(np.uint32))
    # beyond the static scratch: a clear error, nothing launched
    too_many = L.bgb_bytes_to_bitmask_noalloc_max_rows() + 1
    nc = C.c_int32(0)
    rc = L.bgb_bytes_to_bitmask_noalloc(C.c_void_p(b.data_ptr()), too_many, C.c_void_p(m.data_ptr()), C.byref(nc), s)
    assert rc == 1 and b"exceed" in L.bgb_last_error()
    assert L.bgb_bytes_to_bitmask_noalloc(C.c_void_p(b.data_ptr()), 0, None, C.byref(nc), s) == 0 and nc.value == 0
