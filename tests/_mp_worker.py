"""Multi-GPU worker script (one rank per GPU, launched by torch.distributed.run from
test_multigpu.py).  Checks the real cross-rank protocol of K1/K2/K2f over NVLink peer memory
against PyTorch fp32 math, with the C++ coordinator as the only control plane (no
torch.distributed process group is created).  Prints 'MP_OK <rank>' on success."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch

import parameter_server_distributed_b200 as psd
from parameter_server_distributed_b200 import OptimConfig
from parameter_server_distributed_b200.models import MLP, MLPConfig, mse_loss
from parameter_server_distributed_b200.ops import reference


def _log(msg):
    """Progress trace per rank (post-mortem of a hang): gpurun_out/mp_rank<r>.log if possible."""
    d = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(d):
        with open(os.path.join(d, f"mp_rank{os.environ.get('RANK', '0')}.log"), "a") as f:
            f.write(msg + "\n")


def check_training(ctx, pull, kind, clip, tma=False, split_mb=None, nvls=None):
    """Every iteration is checked as ONE step from the engine's actual state: N workers through
    the fused push/pull == fp32 PyTorch update with the mean of the workers' gradients."""
    torch.manual_seed(0)
    cfg = MLPConfig(d_in=512, hidden=(2048, 768), d_out=256)
    model = MLP(cfg).to(ctx.device, torch.bfloat16)
    optim = OptimConfig(kind=kind, lr=1e-2 if kind == "sgd" else 1e-3)
    tr = psd.Trainer(model, mse_loss, ctx, optim, pull=pull, bucket_mb=1.0, clip_norm=clip,
                     push_tma=tma, split_mb=split_mb, push_nvls=nvls)
    if nvls:
        assert tr.ps._variant(0).get("nvls"), "the NVLS push was not selected"
    if split_mb:
        assert len(tr.ps.plan.slots["layers.0.weight"].pieces) > 1, "first layer was not split"
    omodel = MLP(cfg).to(ctx.device, torch.bfloat16)
    gen = torch.Generator(device="cpu").manual_seed(5)
    for it in range(4):
        masters = tr.ps.gather_state("master")
        m = tr.ps.gather_state("m") or {n: torch.zeros_like(p) for n, p in masters.items()}
        v = tr.ps.gather_state("v") or {n: torch.zeros_like(p) for n, p in masters.items()}
        omodel.load_state_dict({k: p.detach().clone() for k, p in model.state_dict().items()})
        xs = torch.randn(ctx.world, 256, 512, generator=gen).to(torch.bfloat16)
        ys = torch.randn(ctx.world, 256, 256, generator=gen).to(torch.bfloat16)
        tr.step(xs[ctx.rank].pin_memory(), ys[ctx.rank].pin_memory()).item()
        grads = {n: torch.zeros_like(p) for n, p in masters.items()}
        for w in range(ctx.world):
            omodel.zero_grad()
            mse_loss(omodel(xs[w].to(ctx.device)), ys[w].to(ctx.device)).backward()
            for n, p in omodel.named_parameters():
                grads[n] += p.grad.float()  # bf16 grads summed in fp32, like K1
        tot = sum(((g / ctx.world) ** 2).sum() for g in grads.values()).item()
        coef = reference.clip_coef(tot, clip) if clip else 1.0
        rcfg = OptimConfig(**{**optim.__dict__, "grad_scale": 1.0 / ctx.world})
        for n in masters:
            reference.optim_step(rcfg, masters[n].view(-1), grads[n].view(-1), m[n].view(-1),
                                 v[n].view(-1), clip_coef=coef, step=it)
        if pull in ("gather", "hybrid"):
            # the first-layer weight is gathered lazily by the fused GEMM of the next forward
            with torch.no_grad():
                model(xs[ctx.rank].to(ctx.device))
        torch.cuda.synchronize()
        new_masters = tr.ps.gather_state("master")
        for n, p in model.named_parameters():
            exp = masters[n].to(torch.bfloat16).float()
            diff = (p.float() - exp).abs()
            tol = 1e-2 * (exp.abs().max().item() + 1e-3)
            if kind in ("adam", "adamw"):
                # sign(g) decides an lr-sized move: bf16 noise on near-zero gradients flips it
                tol += 2.2 * optim.lr
            assert diff.max().item() <= tol, (pull, kind, it, n, diff.max().item(), tol)
            assert torch.equal(p.detach(), new_masters[n].to(torch.bfloat16)), \
                ("replica != master", pull, kind, it, n)
        if clip:
            got = tr.ps.grad_norm().item()
            assert abs(got - tot ** 0.5) < 2e-2 * tot ** 0.5, (it, got, tot ** 0.5)
    # all ranks hold identical replicas
    flat = tr.ps.backend.param_flat.clone()
    ref0 = ctx.fabric.all_gather(flat)[0]
    assert torch.equal(flat, ref0), "replicas diverged"
    tr.ps.close()


def check_fused_gemm(ctx):
    import ctypes as C
    from parameter_server_distributed_b200 import _lib
    world, rank = ctx.world, ctx.rank
    M, K = 1024, 1024
    N = 256 * world * 2
    buf = ctx.fabric.alloc(N * K * 2, align=4096)
    torch.manual_seed(11)
    w_full = (torch.randn(N, K, device=ctx.device) * 0.5).to(torch.bfloat16)
    rows = N // world
    t = buf.view(torch.bfloat16).view(N, K)
    t.zero_()
    t[rank * rows:(rank + 1) * rows].copy_(w_full[rank * rows:(rank + 1) * rows])
    torch.cuda.synchronize()
    ctx.barrier()
    x = (torch.randn(M, K, device=ctx.device) * 0.5).to(torch.bfloat16)
    y = torch.empty(M, N, dtype=torch.bfloat16, device=ctx.device)
    a = _lib.PsbGemmAG()
    a.fabric = ctx.fabric.c_fabric()
    for q in range(world):
        a.w[q] = buf.peer_ptr(q)
    a.x, a.bias, a.y = x.data_ptr(), None, y.data_ptr()
    a.M, a.N, a.K = M, N, K
    a.out_dtype, a.bucket, a.epoch, a.skip_wait = _lib.BF16, 0, None, 1
    _lib.check(_lib.load().psb_gemm_allgather_bf16(C.byref(a),
                                                   torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    exp = x.float() @ w_full.float().t()
    err = (y.float() - exp).abs().max().item() / exp.abs().max().item()
    assert err < 1.5e-2, err
    assert torch.equal(t, w_full), "gathered weight differs"
    ctx.barrier()
    buf.free()


class _Vec(torch.nn.Module):
    def __init__(self, n):
        super().__init__()
        self.a = torch.nn.Parameter(torch.zeros(n))
        self.b = torch.nn.Parameter(torch.zeros(n // 3))


def check_async(ctx, staleness):
    """Async / SSP on versioned shard buffers: every gradient is applied exactly once, the
    staleness bound holds at every step, replicas agree after drain().  Rank 1 is a straggler."""
    from parameter_server_distributed_b200 import _lib
    from parameter_server_distributed_b200.parallel.async_engine import AsyncParameterServer
    torch.manual_seed(3)
    n, T, lr = 300_000, 12, 0.5
    model = _Vec(n).to(ctx.device, torch.bfloat16)
    ps = AsyncParameterServer(list(model.named_parameters()), OptimConfig(kind="sgd", lr=lr),
                              ctx.fabric, bucket_mb=0.3, staleness=staleness)
    assert len(ps.plan.buckets) >= 2
    pad = ctx.fabric.pad.view(torch.int32)
    total = {k: torch.zeros_like(p, dtype=torch.float32) for k, p in model.named_parameters()}
    worst = 0
    for c in range(1, T + 1):
        if ctx.rank == 1 and c % 3 == 0:
            torch.cuda._sleep(int(2e8))  # ~100 ms straggle on the device
        loss = 0
        for k, p in model.named_parameters():
            # gradient of worker w at clock c: a constant pattern that is exact in bf16
            coef = torch.full_like(p, float((ctx.rank + 1) * (c % 4 + 1)) / 64.0)
            loss = loss + (p * coef).sum()
        loss.backward()
        ps.step()
        _log(f"async s={staleness} clock {c} launched")
        torch.cuda.synchronize()
        _log(f"async s={staleness} clock {c} done")
        clocks = pad[_lib.PAD_CLOCK_OFF:_lib.PAD_CLOCK_OFF + ctx.world]
        lag = ps.clock - int(clocks.min().item())
        worst = max(worst, lag)
        if staleness is not None:
            # clock c started only after everybody had published clock c - 1 - staleness
            assert lag <= staleness + 1, (ctx.rank, c, lag, staleness)
        assert lag <= ps.ring
    ps.drain()
    _log(f"async s={staleness} drained")
    assert int(ps.applied_clocks().min().item()) == T and int(ps.applied_clocks().max().item()) == T
    for k, p in model.named_parameters():
        s_all = sum((w + 1) * (c % 4 + 1) for w in range(ctx.world) for c in range(1, T + 1)) / 64.0
        exp = torch.full_like(p, -lr * s_all / ctx.world, dtype=torch.float32)
        got = ps.gather_state("master")[k]
        assert torch.allclose(got, exp, rtol=1e-5, atol=1e-5), (k, got.flatten()[:3], exp.flatten()[:3])
        assert torch.equal(p.detach(), got.to(torch.bfloat16)), f"replica of {k} is not the master"
    flat = ps.backend.param_flat.clone()
    ref0 = ctx.fabric.all_gather(flat)[0]
    assert torch.equal(flat, ref0), "async replicas diverged after drain"
    ps.close()
    return worst


def check_tracks_oracle(ctx):
    """A run whose loss MOVES (about 5x down): 200 steps of a teacher-student regression through the
    sharded parameter server must follow a single-process fp32 run on the global batch.  At the
    noise floor of random targets a wrong reduction hides; here a missing or doubled worker
    changes the trajectory by far more than the tolerance."""
    torch.manual_seed(21)
    cfg = MLPConfig(d_in=256, hidden=(512,), d_out=256)
    teacher = MLP(cfg).to(ctx.device, torch.float32)
    model = MLP(cfg).to(ctx.device, torch.bfloat16)
    oracle = MLP(cfg).to(ctx.device, torch.float32)
    oracle.load_state_dict({k: v.float() for k, v in model.state_dict().items()})
    opt = torch.optim.SGD(oracle.parameters(), lr=4.0)
    tr = psd.Trainer(model, mse_loss, ctx, OptimConfig(kind="sgd", lr=4.0), bucket_mb=0.25)
    gen = torch.Generator(device="cpu").manual_seed(77)
    B = 128
    first = last = olast = None
    for it in range(200):
        x = torch.randn(ctx.world, B, cfg.d_in, generator=gen)
        with torch.no_grad():
            y = teacher(x.to(ctx.device).view(-1, cfg.d_in)).view(ctx.world, B, cfg.d_out)
        loss = tr.step_device(x[ctx.rank].to(ctx.device, torch.bfloat16),
                              y[ctx.rank].to(torch.bfloat16))
        opt.zero_grad()
        ol = torch.nn.functional.mse_loss(oracle(x.to(ctx.device).view(-1, cfg.d_in)),
                                          y.view(-1, cfg.d_out))
        ol.backward()
        opt.step()
        if it == 0:
            first = float(ol.item())
        last, olast = float(loss.item()), float(ol.item())
    mean_last = sum(ctx.all_gather_object(last, "oracle/loss")) / ctx.world
    assert olast < 0.5 * first, ("the oracle itself did not learn", first, olast)
    assert abs(mean_last - olast) < 0.15 * olast + 1e-4, \
        ("loss left the fp32 oracle's trajectory", mean_last, olast, first)
    worst = 0.0
    for (n, p), (_, q) in zip(model.named_parameters(), oracle.named_parameters()):
        rel = (p.float() - q).norm().item() / (q.norm().item() + 1e-12)
        worst = max(worst, rel)
    assert worst < 0.05, ("weights left the fp32 oracle's trajectory", worst)
    tr.ps.close()
    return first, olast, mean_last, worst


def check_control_plane(ctx):
    """The coordinator is the membership / shard-assignment authority of the GPU plane."""
    assert ctx.coord is not None, "the GPU plane did not come up through the coordinator"
    assert not torch.distributed.is_initialized(), "a torch.distributed group exists"
    assert ctx.coord.list_workers() == list(range(ctx.world)), ctx.coord.list_workers()
    shards, total = ctx.coord.shard_map()
    assert [s[0] for s in shards] == list(range(ctx.world)) and total == ctx.world, (shards, total)
    assert sorted(s[2] for s in shards) == list(range(ctx.world)), shards  # one GPU each
    got = ctx.all_gather_object({"rank": ctx.rank}, "cp/obj")
    assert [g["rank"] for g in got] == list(range(ctx.world))
    assert ctx.reduce_max(float(ctx.rank), "cp/max") == ctx.world - 1


def main():
    ctx = psd.init(heap_bytes=256 << 20)
    assert ctx.world >= 2
    check_control_plane(ctx)
    _log("control plane ok")
    check_fused_gemm(ctx)
    _log("fused gemm ok")
    for pull, kind, clip, tma in [("broadcast", "sgd", None, False), ("broadcast", "adamw", 0.5, False),
                                  ("gather", "momentum", None, False), ("gather", "adam", None, False),
                                  ("broadcast", "sgd", None, True), ("gather", "adamw", None, True),
                                  ("hybrid", "momentum", None, True), ("hybrid", "adamw", 1.0, False)]:
        check_training(ctx, pull, kind, clip, tma)
        ctx.barrier()
        _log(f"training {pull} {kind} tma={tma} ok")
    # NVLS: the switch reduces (multimem.ld_reduce) and replicates (multimem.st); tolerance test
    if ctx.fabric.mc_base:
        for kind, split in (("sgd", None), ("adamw", 0.5), ("momentum", None)):
            check_training(ctx, "broadcast", kind, None, None, split_mb=split, nvls=True)
            ctx.barrier()
            _log(f"training nvls {kind} ok")
    if ctx.rank == 0:
        print(f"fabric: {ctx.fabric.kind} heap, multicast "
              f"{'bound (NVLS push tested)' if ctx.fabric.mc_base else 'not available'}", flush=True)
    # the first layer's weight in row blocks, one bucket each (dW block by block, early pushes)
    for kind, tma in (("sgd", None), ("adamw", None), ("momentum", False)):
        check_training(ctx, "broadcast", kind, None, tma, split_mb=0.5)
        ctx.barrier()
        _log(f"training split {kind} ok")
    for st in (None, 1, 0):
        worst = check_async(ctx, st)
        ctx.barrier()
        if ctx.rank == 0:
            print(f"async staleness={st}: worst observed lag {worst}", flush=True)
    first, olast, mine, wrel = check_tracks_oracle(ctx)
    ctx.barrier()
    if ctx.rank == 0:
        print(f"oracle tracking: loss {first:.4f} -> oracle {olast:.5f} / ours {mine:.5f}, "
              f"worst relative weight distance {wrel:.4f}", flush=True)
    print(f"MP_OK {ctx.rank}", flush=True)
    ctx.close()


if __name__ == "__main__":
    try:
        main()
    except BaseException as e:  # one compact line per rank: torchrun's own traceback is noise
        import traceback
        tb = traceback.extract_tb(e.__traceback__)[-1]
        print(f"MP_FAIL rank={os.environ.get('RANK')} {type(e).__name__}: {str(e)[:400]} "
              f"({os.path.basename(tb.filename)}:{tb.lineno})", flush=True)
        os._exit(1)
