"""Model families of the BASELINE.json configs through the public Trainer API on one GPU:
loss goes down, the bf16 replica is exactly the rounded fp32 master shard, channels-last
convolution weights keep their layout, the async engine gives the same result as the sync one
when there is a single worker."""
import pytest
import torch

import parameter_server_distributed_b200 as psd
from parameter_server_distributed_b200.models import (GPT2, GPT2Config, MLP, MLPConfig, ce_loss,
                                                      lm_loss, mse_loss, resnet_tiny)

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = psd.init(heap_bytes=96 << 20)
    yield c
    c.close()


def _check_replica(tr):
    masters = tr.ps.gather_state("master")
    for n, p in tr.model.named_parameters():
        assert torch.equal(p.detach(), masters[n].to(torch.bfloat16)), n
        assert torch.isfinite(p).all()


def test_gpt2_tiny_adamw_clip(ctx):
    torch.manual_seed(0)
    model = GPT2(GPT2Config.tiny()).to(ctx.device, torch.bfloat16)
    tr = psd.Trainer(model, lm_loss, ctx, psd.OptimConfig(kind="adamw", lr=3e-3, weight_decay=0.01),
                     bucket_mb=0.1, clip_norm=1.0)
    ids = torch.randint(0, 512, (8, 33))
    x, y = ids[:, :-1].contiguous().pin_memory(), ids[:, 1:].contiguous().pin_memory()
    losses = [tr.step(x, y).item() for _ in range(12)]
    assert losses[-1] < losses[0] - 0.5, losses
    assert tr.ps.grad_norm().item() > 0
    _check_replica(tr)
    tr.ps.close()


def test_resnet_tiny_momentum_channels_last(ctx):
    torch.manual_seed(1)
    model = resnet_tiny().to(ctx.device, torch.bfloat16).to(memory_format=torch.channels_last)
    tr = psd.Trainer(model, ce_loss, ctx, psd.OptimConfig(kind="momentum", lr=0.05), bucket_mb=0.05)
    assert len(tr.ps._nhwc) > 0
    x = torch.randn(16, 3, 64, 64).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    y = torch.randint(0, 10, (16,))
    xs, ys = x.pin_memory(), y.pin_memory()
    losses = [tr.step(xs, ys).item() for _ in range(10)]
    assert losses[-1] < losses[0], losses
    w = model.layer1[0].conv2.weight
    assert w.is_contiguous(memory_format=torch.channels_last)
    _check_replica(tr)
    tr.ps.close()


def test_async_single_worker_equals_sync(ctx):
    cfg = MLPConfig(d_in=256, hidden=(512,), d_out=64)
    outs = []
    for mode in ("sync", "ssp"):
        torch.manual_seed(2)
        model = MLP(cfg).to(ctx.device, torch.bfloat16)
        tr = psd.Trainer(model, mse_loss, ctx, psd.OptimConfig(kind="momentum", lr=0.02),
                         mode=mode, staleness=1, bucket_mb=0.2)
        g = torch.Generator().manual_seed(3)
        for _ in range(5):
            x = torch.randn(64, 256, generator=g).to(torch.bfloat16).pin_memory()
            y = torch.randn(64, 64, generator=g).to(torch.bfloat16).pin_memory()
            tr.step(x, y).item()
        if mode != "sync":
            tr.ps.drain()
        outs.append({n: p.detach().clone() for n, p in model.named_parameters()})
        tr.ps.close()
    for n in outs[0]:
        assert torch.equal(outs[0][n], outs[1][n]), n


@pytest.mark.gpu
def test_second_engine_on_reused_heap_ranges_sees_clean_padding(ctx):
    """Closing an engine returns its symmetric buffers to the heap's free list and the next
    engine gets the same ranges back: the padding between tensors (reduced and normed like real
    elements) must not carry the first engine's values."""
    import copy
    dev = ctx.device
    if True:
        for rnd in range(3):
            torch.manual_seed(rnd)
            cfg = MLPConfig(d_in=264, hidden=(520,), d_out=72)  # odd sizes: padded slots
            model = MLP(cfg).to(dev, torch.bfloat16)
            oracle = copy.deepcopy(model).float()
            tr = psd.Trainer(model, mse_loss, ctx, psd.OptimConfig(kind="sgd", lr=1e-3),
                             bucket_mb=0.25, clip_norm=1e6)
            x = torch.randn(64, 264, device=dev).to(torch.bfloat16)
            y = torch.randn(64, 72, device=dev).to(torch.bfloat16)
            tr.step_device(x, y)
            torch.nn.functional.mse_loss(oracle(x.float()), y.float()).backward()
            exp = torch.sqrt(sum((p.grad ** 2).sum() for p in oracle.parameters())).item()
            got = tr.ps.grad_norm().item()
            assert abs(got - exp) < 3e-2 * exp, (rnd, got, exp)
            tr.ps.close()


@pytest.mark.gpu
def test_prefetched_batches_train_like_batches_passed_to_step(ctx):
    """Trainer.prefetch(next batch) + step() is the same training as step(batch)."""
    dev = ctx.device
    cfg = MLPConfig(d_in=256, hidden=(512,), d_out=128)
    gen = torch.Generator().manual_seed(3)
    batches = [(torch.randn(64, 256, generator=gen).to(torch.bfloat16).pin_memory(),
                torch.randn(64, 128, generator=gen).to(torch.bfloat16).pin_memory()) for _ in range(5)]
    runs = []
    for use_prefetch in (False, True):
        torch.manual_seed(11)
        model = MLP(cfg).to(dev, torch.bfloat16)
        tr = psd.Trainer(model, mse_loss, ctx, psd.OptimConfig(kind="sgd", lr=0.05), bucket_mb=0.25)
        losses = []
        if use_prefetch:
            tr.prefetch(*batches[0])
            for i in range(len(batches)):
                h = tr.step()
                if i + 1 < len(batches):
                    tr.prefetch(*batches[i + 1])
                losses.append(h.item())
            with pytest.raises(ValueError):
                tr.step()  # nothing prefetched, nothing passed
        else:
            losses = [tr.step(*b).item() for b in batches]
        runs.append(losses)
        tr.ps.close()
    assert runs[0] == runs[1], runs
