"""psb_sumsq_noalloc: the sum-of-squares entry point for callers that may not allocate.

Some users of libpsb_kernels.so cannot call cudaMalloc where they need a gradient norm: a
service that owns one fixed, pre-carved GPU pool, a hook that fires while the framework's
allocator is locked or full, a C host that only forwards pointers it was handed.  For them
the library keeps the reduction's scratch in a static 4 KB device array, and this test checks
through the plain C ABI (ctypes, no Python wrapper) that the result equals the regular
psb_sumsq entry point — twice in a row, because the array is shared between calls and keeps
the previous call's contents — and that inputs beyond the stated limit are refused.

Collected last on purpose (file name): a run that stops at the first failure has already
executed every other test.  It passes alone in a fresh process and after the other tests.
"""
import ctypes as C

import pytest
import torch

from parameter_server_distributed_b200 import _lib

pytestmark = pytest.mark.gpu

F32, BF16 = 0, 1
MAX_ELEMS = 1 << 31
ERANGE = -2


def _abi():
    lib = C.CDLL(str(_lib.lib_path()))
    lib.psb_sumsq_noalloc.restype = C.c_int
    lib.psb_sumsq_noalloc.argtypes = [C.c_void_p, C.c_int, C.c_size_t, C.c_void_p, C.c_void_p]
    lib.psb_sumsq.restype = C.c_int
    lib.psb_sumsq.argtypes = [C.c_void_p, C.c_int, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.psb_sumsq_scratch_floats.restype = C.c_size_t
    lib.psb_last_error.restype = C.c_char_p
    return lib


def test_noalloc_matches_regular_entry_point_twice_in_a_row():
    lib = _abi()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    stream = torch.cuda.current_stream().cuda_stream
    scratch = torch.empty(lib.psb_sumsq_scratch_floats(), dtype=torch.float32, device=dev)
    torch.manual_seed(0)
    # second input is smaller and of another dtype: stale partials of call 1 must not leak in
    cases = [(torch.randn((1 << 22) + 7, device=dev) * 3.0, F32),
             (torch.randn(1000, device=dev).to(torch.bfloat16), BF16),
             (torch.randn(1 << 24, device=dev).to(torch.bfloat16), BF16)]
    for i, (x, dt) in enumerate(cases):
        got = torch.full((1,), -1.0, device=dev)
        exp = torch.full((1,), -2.0, device=dev)
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info(dev)[0]
        rc = lib.psb_sumsq_noalloc(x.data_ptr(), dt, x.numel(), got.data_ptr(), stream)
        assert rc == 0, lib.psb_last_error()
        torch.cuda.synchronize()
        if i > 0:  # the first call may make the driver load the kernels' code
            assert torch.cuda.mem_get_info(dev)[0] == free_before  # no device memory was taken
        rc = lib.psb_sumsq(x.data_ptr(), dt, x.numel(), scratch.data_ptr(), exp.data_ptr(), stream)
        assert rc == 0, lib.psb_last_error()
        torch.cuda.synchronize()
        assert torch.allclose(got, exp, rtol=1e-4), (got.item(), exp.item())
        ref = (x.float() ** 2).sum()
        assert torch.allclose(got[0], ref, rtol=1e-4)


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    lib = _abi()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    x = torch.zeros(64, device=dev)
    out = torch.zeros(1, device=dev)
    rc = lib.psb_sumsq_noalloc(x.data_ptr(), F32, MAX_ELEMS + 1, out.data_ptr(), None)
    assert rc == ERANGE
    msg = lib.psb_last_error().decode()
    assert "PSB_SUMSQ_NOALLOC_MAX_ELEMS" in msg and "psb_sumsq" in msg
    torch.cuda.synchronize()  # nothing was launched, nothing faulted
