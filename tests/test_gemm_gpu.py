"""K2f numerics: the hand-written tcgen05/TMEM/TMA GEMM against a PyTorch fp32 matmul, and the
fused all-gather variant with N emulated ranks on one GPU (the gather CTAs then copy between
buffers of the same device; peer traffic is covered by tests/test_multigpu.py)."""
import ctypes as C

import pytest
import torch

from parameter_server_distributed_b200 import _lib
from parameter_server_distributed_b200.ops import native, reference
from parameter_server_distributed_b200.parallel.fabric import Fabric

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda:0")


@pytest.fixture(autouse=True, params=[1, 0], ids=["cta_pair", "single_cta"])
def gemm_variant(request):
    """Every test runs on both kernels: cta_group::2 (default) and cta_group::1."""
    _lib.check(_lib.load().psb_gemm_select(request.param))
    yield request.param
    _lib.check(_lib.load().psb_gemm_select(1))


def _check(y, x, w, bias=None):
    exp = x.float() @ w.float().t()
    if bias is not None:
        exp = exp + bias
    err = (y.float() - exp).abs().max().item()
    scale = exp.abs().max().item() + 1e-6
    assert err / scale < 1.5e-2, f"max err {err} (scale {scale})"


@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (128, 256, 256), (256, 512, 128),
                                   (1024, 1024, 1024), (4096, 2048, 512), (300, 264, 72),
                                   (1, 8, 8), (129, 257 * 8, 1032), (8192, 8192, 1024)])
def test_gemm_shapes(M, N, K):
    torch.manual_seed(M + N + K)
    x = (torch.randn(M, K, device=DEV) * 0.5).to(torch.bfloat16)
    w = (torch.randn(N, K, device=DEV) * 0.5).to(torch.bfloat16)
    y = native.gemm_bf16(x, w)
    torch.cuda.synchronize()
    _check(y, x, w)


def test_gemm_bias_fp32_out():
    M, N, K = 512, 768, 320
    x = torch.randn(M, K, device=DEV).to(torch.bfloat16)
    w = torch.randn(N, K, device=DEV).to(torch.bfloat16)
    b = torch.randn(N, device=DEV)
    y = native.gemm_bf16(x, w, b, out_dtype=torch.float32)
    exp = x.float() @ w.float().t() + b
    assert torch.allclose(y, exp, atol=1e-2, rtol=1e-3)


def test_gemm_structured_operands():
    """Catches descriptor / swizzle mistakes that random data can hide: W = identity-like."""
    M, N, K = 256, 256, 256
    x = torch.arange(M * K, device=DEV, dtype=torch.float32).reshape(M, K).remainder(17).to(torch.bfloat16)
    w = torch.eye(N, K, device=DEV).to(torch.bfloat16)
    y = native.gemm_bf16(x, w, out_dtype=torch.float32)
    assert torch.equal(y, x.float())


def test_gemm_repeatable():
    x = torch.randn(1024, 512, device=DEV).to(torch.bfloat16)
    w = torch.randn(1024, 512, device=DEV).to(torch.bfloat16)
    a = native.gemm_bf16(x, w)
    b = native.gemm_bf16(x, w)
    assert torch.equal(a, b)


@pytest.mark.parametrize("world", [2, 4])
def test_gemm_allgather_emulated(world):
    M, K = 640, 512
    N = 256 * world * 2
    torch.manual_seed(world)
    fabs = Fabric.local_group(world, DEV, heap_bytes=N * K * 2 + (4 << 20))
    bufs = [f.alloc(N * K * 2, align=4096) for f in fabs]
    w_full = (torch.randn(N, K, device=DEV) * 0.5).to(torch.bfloat16)
    rows = N // world
    for r, buf in enumerate(bufs):  # each replica holds only its own row slice
        t = buf.view(torch.bfloat16).view(N, K)
        t.zero_()
        t[r * rows:(r + 1) * rows].copy_(w_full[r * rows:(r + 1) * rows])
    x = (torch.randn(M, K, device=DEV) * 0.5).to(torch.bfloat16)
    bias = torch.randn(N, device=DEV)
    lib = _lib.load()
    for r, f in enumerate(fabs):
        a = _lib.PsbGemmAG()
        a.fabric = f.c_fabric()
        for q in range(world):
            a.w[q] = bufs[q].peer_ptr(q)
        y = torch.empty(M, N, dtype=torch.bfloat16, device=DEV)
        a.x, a.bias, a.y = x.data_ptr(), bias.data_ptr(), y.data_ptr()
        a.M, a.N, a.K = M, N, K
        a.out_dtype = _lib.BF16
        a.bucket, a.epoch, a.skip_wait = 0, None, 1
        _lib.check(lib.psb_gemm_allgather_bf16(C.byref(a), torch.cuda.current_stream().cuda_stream))
        torch.cuda.synchronize()
        _check(y, x, w_full, bias)
        # the gather persisted the full weight into this rank's replica
        assert torch.equal(bufs[r].view(torch.bfloat16).view(N, K), w_full)
    for f in fabs:
        f.close()
