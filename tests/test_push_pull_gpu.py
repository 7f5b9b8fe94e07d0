"""T5 on one GPU: the fused push (K1) and pull (K2) with N emulated ranks whose heaps all
live on cuda:0.  The ranks' kernels are launched one after the other with NO_WAIT (spinning
kernels must never be serialised on one GPU); the cross-rank protocol itself is exercised
by tests/test_multigpu.py on >= 2 GPUs.  Oracle: fixed-order fp32 sum + PyTorch optimizer
(reference math: /root/reference/src/parameter_server.cpp:49-60, 83-90)."""
import pytest
import torch

from parameter_server_distributed_b200 import OptimConfig, _lib
from parameter_server_distributed_b200.ops import reference
from parameter_server_distributed_b200.parallel.buckets import plan_buckets
from parameter_server_distributed_b200.parallel.device_backend import DeviceBackend
from parameter_server_distributed_b200.parallel.fabric import Fabric

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda:0")


def _setup(world, kind, clip=False, broadcast=True, shapes=None, tma=False):
    shapes = shapes or [("w0", (257, 130)), ("b0", (257,)), ("w1", (1000, 257))]
    plan = plan_buckets(shapes, world, bucket_elems=200_000)
    fabs = Fabric.local_group(world, DEV, heap_bytes=plan.total_elems * 4 + (8 << 20))
    cfg = OptimConfig(kind=kind, lr=0.05 if kind in ("sgd", "momentum") else 0.002,
                      weight_decay=0.01, grad_scale=1.0 / world)
    bes = [DeviceBackend(f, plan, cfg, broadcast=broadcast, clip=clip, push_tma=tma) for f in fabs]
    return plan, fabs, cfg, bes


@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("kind", ["sgd", "adam"])
@pytest.mark.parametrize("tma", [False, True], ids=["ldg", "tma"])
def test_fused_push_broadcast(world, kind, tma):
    torch.manual_seed(world)
    plan, fabs, cfg, bes = _setup(world, kind, tma=tma)
    total, shard = plan.total_elems, plan.shard_elems
    init = torch.randn(total, device=DEV).to(torch.bfloat16)
    ref_master = init.float().clone()
    ref_m, ref_v = torch.zeros(total, device=DEV), torch.zeros(total, device=DEV)
    for r, be in enumerate(bes):
        be.param_flat.copy_(init)
        for b, bk in enumerate(plan.buckets):
            s = bk.slice_elems
            be.shard_slice(be.master, b).copy_(init[bk.base + r * s: bk.base + (r + 1) * s].float())
    for it in range(3):
        grads = [torch.randn(total, device=DEV).to(torch.bfloat16) for _ in range(world)]
        for be, g in zip(bes, grads):
            be.grad_flat.copy_(g)
        for b in range(len(plan.buckets)):
            for be in bes:
                be.push(b, no_wait=True)
        for be in bes:
            be.advance_step()
        torch.cuda.synchronize()
        mean = reference.reduce_mean(grads)
        rcfg = OptimConfig(**{**cfg.__dict__, "grad_scale": 1.0})
        reference.optim_step(rcfg, ref_master, mean, ref_m, ref_v, step=it)
        exp_bf16 = ref_master.to(torch.bfloat16)
        for r, be in enumerate(bes):
            # every replica holds every owner's new slice
            assert torch.allclose(be.param_flat.float(), exp_bf16.float(), atol=1e-2, rtol=1e-2)
            for b, bk in enumerate(plan.buckets):
                s = bk.slice_elems
                exp = ref_master[bk.base + r * s: bk.base + (r + 1) * s]
                got = be.shard_slice(be.master, b)
                # Adam's m / sqrt(v) is ill-conditioned where the gradient cancels to ~0:
                # tolerate a vanishing fraction of lr-sized outliers
                bad = ~torch.isclose(got, exp, atol=2e-5, rtol=1e-5)
                assert bad.float().mean().item() < 1e-3
                assert (got - exp).abs().max().item() < (
                    2e-5 if kind == "sgd" else 2 * cfg.lr * (it + 1))
            assert int(be.epoch[0].item()) == it + 1 and int(be.done.sum().item()) == 0
            ver = fabs[r].pad.view(torch.int32)[_lib.PAD_VERSION_OFF:_lib.PAD_VERSION_OFF + world]
            assert ver.tolist() == [it + 1] * world
    for f in fabs:
        f.close()


@pytest.mark.parametrize("world", [2, 4])
def test_pull_mode_then_allgather(world):
    torch.manual_seed(7)
    plan, fabs, cfg, bes = _setup(world, "sgd", broadcast=False)
    total = plan.total_elems
    for r, be in enumerate(bes):
        be.param_flat.zero_()
    grads = [torch.randn(total, device=DEV).to(torch.bfloat16) for _ in range(world)]
    for be, g in zip(bes, grads):
        be.grad_flat.copy_(g)
    for b in range(len(plan.buckets)):
        for be in bes:
            be.push(b, no_wait=True)
    torch.cuda.synchronize()
    exp = (-cfg.lr * reference.reduce_mean(grads)).to(torch.bfloat16)
    for r, be in enumerate(bes):  # only the owner's slice is fresh before the pull
        for bk in plan.buckets:
            s = bk.slice_elems
            own = slice(bk.base + r * s, bk.base + (r + 1) * s)
            assert torch.allclose(be.param_flat[own].float(), exp[own].float(), atol=1e-2, rtol=1e-2)
    for b in range(len(plan.buckets)):
        for be in bes:
            be.pull(b)  # version flags are already published => the in-kernel wait passes
    torch.cuda.synchronize()
    for be in bes:
        assert torch.allclose(be.param_flat.float(), exp.float(), atol=1e-2, rtol=1e-2)
    for f in fabs:
        f.close()


def test_clip_path_reduce_then_update():
    world = 4
    torch.manual_seed(3)
    plan, fabs, cfg, bes = _setup(world, "momentum", clip=True)
    total = plan.total_elems
    grads = [torch.randn(total, device=DEV).to(torch.bfloat16) for _ in range(world)]
    for be, g in zip(bes, grads):
        be.grad_flat.copy_(g)
    nb = len(plan.buckets)
    for b in range(nb):
        for be in bes:
            be.push(b, reduce_only=True, no_wait=True)
    torch.cuda.synchronize()
    mean = reference.reduce_mean(grads)
    # per-shard Σg² adds up to the global one; emulate the cross-shard sum on the host
    tot = sum(be.sumsq_acc.item() for be in bes)
    assert abs(tot - (mean ** 2).sum().item()) / tot < 1e-4
    coef = reference.clip_coef(tot, 1.0)
    for be in bes:
        be.clip_coef.fill_(coef)
        be.sumsq_acc.zero_()
    for b in range(nb):
        for be in bes:
            be.push(b, from_reduced=True, clip=True)
    torch.cuda.synchronize()
    ref_master = torch.zeros(total, device=DEV)
    ref_m = torch.zeros(total, device=DEV)
    rcfg = OptimConfig(**{**cfg.__dict__, "grad_scale": 1.0})
    reference.optim_step(rcfg, ref_master, mean, ref_m, None, clip_coef=coef)
    for be in bes:
        assert torch.allclose(be.param_flat.float(), ref_master.to(torch.bfloat16).float(),
                              atol=1e-2, rtol=1e-2)
    for f in fabs:
        f.close()


def test_barrier_and_wait_single_rank():
    fab = Fabric(0, 1, DEV, 1 << 20)
    fab.barrier()
    fab.barrier()
    torch.cuda.synchronize()
    assert fab._barrier_seq[0].item() == 2
    fab.close()
