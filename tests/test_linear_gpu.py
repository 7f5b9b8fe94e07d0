"""Training GEMMs on the CTA-pair tcgen05 kernel: MN-major operands (dX = dY·W, dW = dYᵀ·X read
in place, no transposed copies), the fused GELU / dGELU epilogues, the autograd link built
from them and the flagship MLP running on it.  Oracle: plain PyTorch in fp32."""
import pytest
import torch
import torch.nn.functional as F

from parameter_server_distributed_b200.models import MLP, MLPConfig
from parameter_server_distributed_b200.ops import linear, native

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda:0")


def _rand(*shape, scale=0.5):
    return (torch.randn(*shape, device=DEV) * scale).to(torch.bfloat16)


def _close(got, exp, tol=1.5e-2):
    err = (got.float() - exp.float()).abs().max().item()
    scale = exp.float().abs().max().item() + 1e-6
    assert err / scale < tol, f"max err {err} (scale {scale})"


SHAPES = [(256, 256, 64), (256, 256, 256), (512, 768, 320), (304, 264, 72), (8, 8, 8),
          (136, 2056, 1032), (4096, 2048, 512), (1024, 4096, 4096)]


@pytest.mark.parametrize("M,N,K", SHAPES)
def test_gemm_b_mn_major(M, N, K):
    """dX-shaped: A [M,K] K-major, B stored [K,N]."""
    torch.manual_seed(M + N + K)
    a, b = _rand(M, K), _rand(K, N)
    y = native.gemm(a, b, trans_b=True)
    _close(y, a.float() @ b.float())


@pytest.mark.parametrize("M,N,K", SHAPES)
def test_gemm_both_mn_major(M, N, K):
    """dW-shaped: A stored [K,M], B stored [K,N]."""
    torch.manual_seed(M + N + K + 1)
    a, b = _rand(K, M), _rand(K, N)
    y = native.gemm(a, b, trans_a=True, trans_b=True)
    _close(y, a.float().t() @ b.float())


@pytest.mark.parametrize("M,N,K", [(256, 256, 128), (520, 264, 200)])
def test_gemm_a_mn_major_only(M, N, K):
    a, b = _rand(K, M), _rand(N, K)
    y = native.gemm(a, b, trans_a=True, out_dtype=torch.float32)
    _close(y, a.float().t() @ b.float().t())


def test_gemm_mn_major_structured():
    """Identity operands expose any swizzle / descriptor mistake exactly."""
    M = N = K = 256
    x = torch.arange(M * K, device=DEV, dtype=torch.float32).reshape(M, K).remainder(17).to(torch.bfloat16)
    eye = torch.eye(K, N, device=DEV).to(torch.bfloat16)
    assert torch.equal(native.gemm(x, eye, trans_b=True, out_dtype=torch.float32), x.float())
    xt = x.t().contiguous()  # stored [K, M]
    assert torch.equal(native.gemm(xt, eye, trans_a=True, trans_b=True, out_dtype=torch.float32),
                       x.float())
    # larger than one tile in every dimension, still exact (small integers)
    M, N, K = 768, 512, 384
    a = torch.randint(-3, 4, (K, M), device=DEV).to(torch.bfloat16)
    b = torch.randint(-3, 4, (K, N), device=DEV).to(torch.bfloat16)
    y = native.gemm(a, b, trans_a=True, trans_b=True, out_dtype=torch.float32)
    assert torch.equal(y, a.float().t() @ b.float())


@pytest.mark.parametrize("M,N,K", [(256, 256, 128), (300, 264, 72), (2048, 1024, 512)])
def test_gelu_epilogue(M, N, K):
    x, w = _rand(M, K), _rand(N, K, scale=0.1)
    bias = torch.randn(N, device=DEV)
    z = torch.empty(M, N, device=DEV, dtype=torch.bfloat16)
    act = native.gemm(x, w, bias=bias, epilogue="gelu", aux=z)
    z_exp = x.float() @ w.float().t() + bias
    _close(z, z_exp)
    # the activation is computed from the ROUNDED pre-activation
    _close(act, F.gelu(z.float(), approximate="tanh"), tol=1e-2)
    act2 = native.gemm(x, w, bias=bias, epilogue="gelu")  # without storing Z
    assert torch.equal(act, act2)


@pytest.mark.parametrize("M,N,K", [(256, 256, 128), (300, 264, 72), (2048, 1024, 512)])
def test_dgelu_epilogue(M, N, K):
    gy, w = _rand(M, K), _rand(K, N, scale=0.1)
    z = _rand(M, N, scale=1.5)
    got = native.gemm(gy, w, trans_b=True, epilogue="dgelu", aux=z)
    exp = (gy.float() @ w.float()) * linear._dgelu(z)
    _close(got, exp)


def test_dgelu_matches_autograd():
    z = torch.linspace(-6, 6, 4001, device=DEV, dtype=torch.float32, requires_grad=True)
    F.gelu(z, approximate="tanh").sum().backward()
    assert torch.allclose(linear._dgelu(z.detach()), z.grad, atol=1e-6)


@pytest.mark.parametrize("batch,dims", [(256, (256, 512, 256)), (200, (136, 264, 520, 72))])
def test_link_chain_backward(batch, dims):
    """A chain of links against the same chain in fp32 autograd."""
    torch.manual_seed(7)
    ws = [(_rand(dims[i + 1], dims[i], scale=dims[i] ** -0.5)).requires_grad_() for i in range(len(dims) - 1)]
    bs = [(_rand(dims[i + 1], scale=0.1)).requires_grad_() for i in range(len(dims) - 1)]
    x = _rand(batch, dims[0]).requires_grad_()
    z, act = x, None
    n = len(ws)
    for i in range(n):
        z, act = linear.link(z, act, ws[i], bs[i], gelu_in=i > 0, gelu_out=i + 1 < n)
    tgt = _rand(batch, dims[-1])
    F.mse_loss(z.float(), tgt.float()).backward()

    wr = [w.detach().float().requires_grad_() for w in ws]
    br = [b.detach().float().requires_grad_() for b in bs]
    xr = x.detach().float().requires_grad_()
    zr = xr
    for i in range(n):
        zr, _ = linear.link_reference(zr, wr[i], br[i], gelu_in=i > 0, gelu_out=False)
    F.mse_loss(zr, tgt.float()).backward()
    _close(z, zr, tol=3e-2)
    _close(x.grad, xr.grad, tol=5e-2)
    for w, r in zip(ws, wr):
        _close(w.grad, r.grad, tol=5e-2)
    for b, r in zip(bs, br):
        _close(b.grad, r.grad, tol=5e-2)


def test_colsum_scratch_reuse_across_shapes():
    """One scratch serves every shape on a stream: counters must not alias partial sums."""
    for M, N in [(777, 2056), (200, 264), (4096, 1024), (200, 72), (200, 520), (777, 2056)]:
        x = _rand(M, N, scale=1.0)
        got = native.colsum(x, out_dtype=torch.float32)
        exp = x.float().sum(0)
        assert (got - exp).abs().max().item() <= 1e-4 * exp.abs().max().item() + 1e-6, (M, N)


@pytest.mark.parametrize("M,N", [(4096, 8192), (4096, 1024), (100, 264), (1, 8), (777, 2056)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_colsum(M, N, dtype):
    x = _rand(M, N, scale=1.0)
    for _ in range(2):  # the arrival counters must be left clean
        got = native.colsum(x, out_dtype=dtype)
        exp = x.float().sum(0)
        assert got.dtype == dtype
        tol = 1e-2 if dtype == torch.bfloat16 else 1e-5
        assert (got.float() - exp).abs().max().item() <= tol * exp.abs().max().item() + 1e-6
    assert torch.equal(native.colsum(x, out_dtype=dtype), got)  # fixed summation order


@pytest.mark.parametrize("shape", [(4096, 1024), (33, 24), (7,), (300, 1000)])
def test_fused_mse_loss(shape):
    from parameter_server_distributed_b200.models import mse_loss
    a = _rand(*shape, scale=1.0).requires_grad_()
    b = _rand(*shape, scale=1.0)
    loss = mse_loss(a, b)
    (loss * 3.0).backward()
    ar = a.detach().float().requires_grad_()
    ref = F.mse_loss(ar, b.float())
    (ref * 3.0).backward()
    assert loss.dtype == torch.float32
    assert abs(loss.item() - ref.item()) <= 1e-5 * abs(ref.item())
    _close(a.grad, ar.grad, tol=1e-2)
    assert mse_loss(a, b).item() == loss.item()  # scratch left clean, fixed order


def test_gemm_bf16_bias():
    x, w = _rand(256, 128), _rand(264, 128)
    bias = _rand(264, scale=2.0)
    y = native.gemm(x, w, bias=bias, out_dtype=torch.float32)
    _close(y, x.float() @ w.float().t() + bias.float(), tol=1e-5)


def test_mlp_uses_own_gemm_and_matches_library_path():
    from parameter_server_distributed_b200 import _lib
    torch.manual_seed(3)
    cfg = MLPConfig(d_in=512, hidden=(1024, 1024), d_out=256)
    model = MLP(cfg).to(DEV, torch.bfloat16)
    x, y = _rand(384, 512), _rand(384, 256)
    n0 = _lib.launch_count()
    out = model(x)
    F.mse_loss(out.float(), y.float()).backward()
    # 3 forward, 3 dW, 2 dX (the input needs no grad), 3 bias-gradient column sums
    assert _lib.launch_count() - n0 == 3 + 3 + 2 + 3
    grads = [p.grad.clone() for p in model.parameters()]
    model.zero_grad(set_to_none=True)
    linear.ENABLED = False
    try:
        ref = model(x)
        F.mse_loss(ref.float(), y.float()).backward()
    finally:
        linear.ENABLED = True
    _close(out, ref, tol=3e-2)
    for g, p in zip(grads, model.parameters()):
        _close(g, p.grad, tol=5e-2)


def test_gpt2_blocks_on_own_gemms_match_library_path():
    import dataclasses
    from parameter_server_distributed_b200.models import GPT2, GPT2Config, lm_loss
    cfg = dataclasses.replace(GPT2Config.tiny(), n_embd=128, n_head=4)
    torch.manual_seed(0)
    lib = GPT2(cfg).to(DEV, torch.bfloat16)
    torch.manual_seed(0)
    own = GPT2(dataclasses.replace(cfg, own_gemms=True)).to(DEV, torch.bfloat16)
    idx = torch.randint(0, cfg.vocab_size, (4, 65), device=DEV)
    x, y = idx[:, :-1].contiguous(), idx[:, 1:].contiguous()
    lo, ll = own(x), lib(x)
    _close(lo, ll, tol=3e-2)
    lm_loss(lo, y).backward()
    lm_loss(ll, y).backward()
    for (n, a), b in zip(own.named_parameters(), lib.parameters()):
        assert a.grad is not None, n
        err = (a.grad.float() - b.grad.float()).abs().max().item()
        scale = b.grad.float().abs().max().item() + 1e-6
        assert err / scale < 8e-2, (n, err, scale)


@pytest.mark.parametrize("lo,hi", [(0, 256), (256, 768), (512, 520)])
def test_gemm_column_block_operands(lo, hi):
    """dW of a row block of a weight: dY[:, lo:hi]^T · X with both operands used in place
    (row stride > row length), and the K-major forms of the same."""
    torch.manual_seed(lo + hi)
    batch, n_out, k_in = 384, 768, 264
    dy, x = _rand(batch, n_out), _rand(batch, k_in)
    got = native.gemm(dy[:, lo:hi], x, trans_a=True, trans_b=True)
    _close(got, dy[:, lo:hi].float().t() @ x.float())
    wide = _rand(320, 1024)                       # K-major operands cut out of a wider matrix
    a, b = wide[:, 128:392], _rand(136, 1024)[:, 256:520]
    _close(native.gemm(a, b), a.float() @ b.float().t())


def test_row_split_weight_trains_like_the_whole_one():
    """Engine split_mb: the first layer's dW is produced in row blocks, each into its own
    gradient bucket and pushed at once — same numbers as the unsplit run."""
    import parameter_server_distributed_b200 as psd
    from parameter_server_distributed_b200.models import mse_loss
    ctx = psd.init(heap_bytes=64 << 20)
    cfg = MLPConfig(d_in=512, hidden=(1024, 256), d_out=128)
    finals = []
    try:
        for split in (0.0, 0.25):
            torch.manual_seed(5)
            model = MLP(cfg).to(DEV, torch.bfloat16)
            tr = psd.Trainer(model, mse_loss, ctx, psd.OptimConfig(kind="momentum", lr=1e-2),
                             bucket_mb=1.0, split_mb=split)
            n_pieces = len(tr.ps.plan.slots["layers.0.weight"].pieces)
            assert n_pieces == (4 if split else 0), n_pieces
            gen = torch.Generator().manual_seed(9)
            for _ in range(3):
                x = torch.randn(256, 512, generator=gen).to(DEV, torch.bfloat16)
                y = torch.randn(256, 128, generator=gen).to(DEV, torch.bfloat16)
                tr.step_device(x, y)
            torch.cuda.synchronize()
            finals.append([p.detach().float().clone() for p in model.parameters()])
            tr.ps.close()
        for a, b in zip(*finals):
            assert torch.equal(a, b)
    finally:
        ctx.close()
