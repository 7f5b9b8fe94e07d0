"""T4: every elementwise / optimizer / norm kernel against a plain PyTorch fp32 reference."""
import pytest
import torch

from parameter_server_distributed_b200 import OptimConfig, _lib
from parameter_server_distributed_b200.ops import native, reference

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


@pytest.mark.parametrize("n", [1, 7, 8, 1000, 1 << 20, (1 << 20) + 3])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_fill(n, dtype):
    t = torch.zeros(n + 16, dtype=dtype, device=DEV)
    native.fill_(t[5:5 + n], 0.01)  # unaligned start exercises head/tail handling
    torch.cuda.synchronize()
    exp = torch.zeros(n + 16, dtype=dtype, device=DEV)
    exp[5:5 + n] = 0.01
    assert torch.equal(t, exp)


@pytest.mark.parametrize("n", [8, 1000, (1 << 20) + 5])
def test_cast_scale(n):
    src = torch.randn(n, device=DEV)
    dst = torch.empty(n, dtype=torch.bfloat16, device=DEV)
    native.cast_scale(src, dst, 0.5)
    assert torch.equal(dst, (src * 0.5).to(torch.bfloat16))
    back = torch.empty(n, device=DEV)
    native.cast_scale(dst, back, 2.0)
    assert torch.equal(back, dst.float() * 2.0)


@pytest.mark.parametrize("sdt,ddt", [(torch.bfloat16, torch.bfloat16), (torch.float32, torch.bfloat16),
                                     (torch.float32, torch.float32), (torch.bfloat16, torch.float32)])
def test_gather_to_bucket(sdt, ddt):
    sizes = [1000, 8, 77, 4096 * 33, 1] + [64] * 120  # > MAX_GATHER tensors => several launches
    srcs = [torch.randn(s, device=DEV).to(sdt) for s in sizes]
    offs, cur = [], 0
    for s in sizes:
        offs.append(cur)
        cur += (s + 127) // 128 * 128
    bucket = torch.zeros(cur, dtype=ddt, device=DEV)
    native.gather_to_bucket(srcs, bucket, offs, scale=0.25)
    for s, o, t in zip(sizes, offs, srcs):
        assert torch.equal(bucket[o:o + s], (t.float() * 0.25).to(ddt))


@pytest.mark.parametrize("kind", ["sgd", "momentum", "adam", "adamw"])
@pytest.mark.parametrize("gdt", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("n", [8, 4096 + 3, 1 << 20])
def test_optim_step(kind, gdt, n):
    torch.manual_seed(0)
    cfg = OptimConfig(kind=kind, lr=0.01, weight_decay=0.1, grad_scale=0.5)
    master = torch.randn(n, device=DEV)
    m = torch.zeros(n, device=DEV)
    v = torch.zeros(n, device=DEV)
    out = torch.empty(n, dtype=torch.bfloat16, device=DEV)
    step = torch.zeros(1, dtype=torch.int32, device=DEV)
    clip = torch.full((1,), 0.7, device=DEV)
    rm, rmm, rv = master.clone(), m.clone(), v.clone()
    for it in range(3):
        g = torch.randn(n, device=DEV).to(gdt)
        native.optim_step(cfg, master, g, m, v, out, clip, step)
        native.counter_add_(step)
        reference.optim_step(cfg, rm, g, rmm, rv, clip_coef=0.7, step=it)
    torch.cuda.synchronize()
    # Adam divides by sqrt(v): where the effective gradient cancels to ~0 the quotient is
    # ill-conditioned, so allow a vanishing fraction of outliers bounded by one lr-sized step.
    bad = ~torch.isclose(master, rm, atol=1e-5, rtol=1e-5)
    assert bad.float().mean().item() < 1e-4
    assert (master - rm).abs().max().item() < (2 * cfg.lr if cfg.needs_v() else 1e-5)
    if cfg.needs_m():
        assert torch.allclose(m, rmm, atol=1e-5, rtol=1e-4)
    if cfg.needs_v():
        assert torch.allclose(v, rv, atol=1e-6, rtol=1e-4)
    assert torch.equal(out, master.to(torch.bfloat16))
    assert step.item() == 3


def test_optim_compat_matches_reference_server():
    """lr = 1 SGD on the mean gradient is the reference's `p -= g` (golden trace App. C)."""
    cfg = OptimConfig(kind="sgd", lr=1.0, grad_scale=0.5)
    p = torch.full((104,), 0.01, device=DEV)
    g = torch.full((104,), 0.02, device=DEV)  # sum of two workers' 0.01
    for _ in range(4):
        native.optim_step(cfg, p, g)
    assert torch.allclose(p, torch.full_like(p, -0.03), atol=1e-7)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("n", [1, 1000, (1 << 22) + 7])
def test_sumsq_and_clip(dtype, n):
    x = torch.randn(n, device=DEV).to(dtype)
    got = native.sumsq(x)
    exp = reference.sumsq(x)
    assert torch.allclose(got, exp, rtol=1e-4)
    coef, norm = native.clip_coef(got, 1.0)
    assert abs(norm.item() - exp.sqrt().item()) <= 1e-4 * exp.sqrt().item() + 1e-6
    assert abs(coef.item() - reference.clip_coef(exp.item(), 1.0)) < 1e-4


def test_errors_are_loud():
    with pytest.raises(TypeError):
        native.fill_(torch.zeros(4), 1.0)  # CPU tensor: no silent fallback
    m = torch.zeros(16, device=DEV)
    with pytest.raises(_lib.PsbError):
        native.optim_step(OptimConfig(kind="adam"), m, m.clone())  # missing m / v


def test_launch_counter_moves():
    before = _lib.launch_count()
    native.fill_(torch.empty(64, device=DEV), 1.0)
    assert _lib.launch_count() == before + 1
