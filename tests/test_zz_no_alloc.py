"""matmul_cuda_device_noalloc -- the entry point for callers that may not allocate device memory
(applications with a fixed device-memory budget, latency-bound loops).  Its scratch is a static
array inside the library, shared by all calls and never cleared between them, so the test makes
calls in a row with different sizes and modes and requires each to match matmul_cuda_device --
bit for bit, because it is the same code on different scratch -- and the float64 oracle within the
matmul tolerance; it also checks the property that gives the entry point its name: the free device
memory does not move across a call.
Collected last on purpose: a failure here must not hide the other tests behind `-x`."""
import ctypes as C

import numpy as np
import pytest

from _pkg import pkg

pytestmark = pytest.mark.gpu
TOL = 1e-4          # the tolerance of tests/test_gpu_matmul.py


def _both(n, mode, seed):
    import torch
    rng = np.random.default_rng(seed)
    A = rng.uniform(0.5, 1.5, (n, n)).astype(np.float32)
    B = rng.uniform(0.5, 1.5, (n, n)).astype(np.float32)
    dA, dB = torch.from_numpy(A).cuda(), torch.from_numpy(B).cuda()
    d_ref, d_new = torch.zeros_like(dA), torch.zeros_like(dA)
    pkg.matmul_cuda_device(dA.data_ptr(), dB.data_ptr(), d_ref.data_ptr(), n, mode=mode)
    pkg.matmul_cuda_device_noalloc(dA.data_ptr(), dB.data_ptr(), d_new.data_ptr(), n, mode=mode)
    torch.cuda.synchronize()
    C64 = A.astype(np.float64) @ B.astype(np.float64)
    return d_ref.cpu().numpy(), d_new.cpu().numpy(), C64


def test_two_calls_in_a_row_match_the_allocating_entry():
    # second call is smaller, of odd size and in the other mode: stale scratch from the first must not show
    for n, mode, seed in ((640, pkg.MODE_3XTF32, 1), (301, pkg.MODE_3XTF32, 2), (301, pkg.MODE_FFMA, 3), (1024, pkg.MODE_3XTF32, 4)):
        ref, new, C64 = _both(n, mode, seed)
        assert np.max(np.abs(new - ref) / np.abs(ref)) <= TOL, (n, mode)
        assert np.array_equal(new, ref), (n, mode)                       # same kernels, same order: identical bits
        assert np.max(np.abs(new - C64) / np.abs(C64)) <= TOL, (n, mode)   # and right, not merely equal


def test_a_warm_call_allocates_nothing():
    # the defining property: once the per-device set-up has run (first call), a call leaves the amount of free
    # device memory exactly where it was -- no cudaMalloc / cudaFree behind the caller's back
    import torch
    n = 512
    dA = torch.rand((n, n), dtype=torch.float32, device="cuda") + 0.5
    dB = torch.rand((n, n), dtype=torch.float32, device="cuda") + 0.5
    dC = torch.zeros_like(dA)
    for mode in (pkg.MODE_3XTF32, pkg.MODE_FFMA):
        pkg.matmul_cuda_device_noalloc(dA.data_ptr(), dB.data_ptr(), dC.data_ptr(), n, mode=mode)      # set-up happens here
        pkg.matmul_cuda_device_noalloc(dA.data_ptr(), dB.data_ptr(), dC.data_ptr(), 333, mode=mode)    # (and the padded paths' kernels load)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for mode in (pkg.MODE_3XTF32, pkg.MODE_FFMA, pkg.MODE_3XTF32):
        pkg.matmul_cuda_device_noalloc(dA.data_ptr(), dB.data_ptr(), dC.data_ptr(), n, mode=mode)
        pkg.matmul_cuda_device_noalloc(dA.data_ptr(), dB.data_ptr(), dC.data_ptr(), 333, mode=mode)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0, (free0, free1)


def test_size_limit_is_enforced_with_a_message():
    import torch
    n = pkg.NOALLOC_MAX_N + 1
    d = torch.zeros((n, n), dtype=torch.float32, device="cuda")
    st = pkg.lib().matmul_cuda_device_noalloc(C.c_void_p(d.data_ptr()), C.c_void_p(d.data_ptr()), C.c_void_p(d.data_ptr()),
                                              n, pkg.MODE_3XTF32, None)
    assert st == pkg.BAD_ARG
    assert b"exceeds the built-in scratch" in pkg.lib().gauss_cuda_last_error()
