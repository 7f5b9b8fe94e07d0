"""Engine on the host backend: equals torch.optim single-process; gloo world_size=2 equals
data-parallel mean-gradient training (the reference's BSP semantics,
/root/reference/src/parameter_server.cpp:49-60,83-90, with lr/momentum/Adam added)."""
import copy
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from parameter_server_distributed_b200 import OptimConfig
from parameter_server_distributed_b200.models import MLP, MLPConfig, mse_loss
from parameter_server_distributed_b200.parallel.engine import ShardedParameterServer

CFG = MLPConfig(d_in=32, hidden=(64, 48), d_out=8)


def _torch_opt(kind, params):
    return {"sgd": lambda: torch.optim.SGD(params, lr=0.01, weight_decay=0.01),
            "momentum": lambda: torch.optim.SGD(params, lr=0.01, momentum=0.9, weight_decay=0.01),
            "adam": lambda: torch.optim.Adam(params, lr=0.01, weight_decay=0.01),
            "adamw": lambda: torch.optim.AdamW(params, lr=0.01, weight_decay=0.01)}[kind]()


@pytest.mark.parametrize("kind", ["sgd", "momentum", "adam", "adamw"])
def test_single_process_matches_torch(kind):
    torch.manual_seed(0)
    m = MLP(CFG)
    ref = copy.deepcopy(m)
    ps = ShardedParameterServer(list(m.named_parameters()),
                                OptimConfig(kind=kind, lr=0.01, weight_decay=0.01), bucket_mb=0.005)
    opt = _torch_opt(kind, list(ref.parameters()))
    assert len(ps.plan.buckets) > 1
    for _ in range(4):
        x, y = torch.randn(16, 32), torch.randn(16, 8)
        mse_loss(m(x), y).backward()
        ps.step()
        opt.zero_grad()
        mse_loss(ref(x), y).backward()
        opt.step()
    for a, b in zip(m.parameters(), ref.parameters()):
        assert torch.allclose(a, b, atol=2e-6, rtol=1e-5)


def test_row_split_tail_tensor_matches_torch():
    """The tensors whose gradients come last are cut into row blocks with a bucket each
    (engine: split_mb); gradients delivered whole by autograd are handed out block by block."""
    torch.manual_seed(3)
    cfg = MLPConfig(d_in=128, hidden=(256, 64), d_out=8)
    m = MLP(cfg)
    ref = copy.deepcopy(m)
    ps = ShardedParameterServer(list(m.named_parameters()), OptimConfig(kind="adamw", lr=0.01),
                                bucket_mb=0.0625, split_mb=0.02)
    whole = ps.plan.slots["layers.0.weight"]
    assert len(whole.pieces) == 4, [len(s.pieces) for s in ps.plan.slots.values()]
    assert not ps.plan.slots["layers.1.weight"].pieces  # only the tail of backward is split
    opt = torch.optim.AdamW(ref.parameters(), lr=0.01, weight_decay=0.0)
    for _ in range(4):
        x, y = torch.randn(16, 128), torch.randn(16, 8)
        mse_loss(m(x), y).backward()
        ps.step()
        opt.zero_grad()
        mse_loss(ref(x), y).backward()
        opt.step()
    for a, b in zip(m.parameters(), ref.parameters()):
        assert torch.allclose(a, b, atol=2e-6, rtol=1e-5)
    got = ps.gather_state("master")
    for (n, p) in ref.named_parameters():
        assert torch.allclose(got[n], p, atol=2e-6, rtol=1e-5), n
    ps.scatter_state({n: torch.zeros_like(p) for n, p in got.items()})
    assert all(float(p.abs().max()) == 0.0 for p in m.parameters())


def test_clip_matches_torch():
    torch.manual_seed(1)
    m = MLP(CFG)
    ref = copy.deepcopy(m)
    ps = ShardedParameterServer(list(m.named_parameters()), OptimConfig(kind="sgd", lr=0.1),
                                bucket_mb=0.005, clip_norm=0.05)
    opt = torch.optim.SGD(ref.parameters(), lr=0.1)
    for _ in range(3):
        x, y = torch.randn(16, 32), torch.randn(16, 8)
        mse_loss(m(x), y).backward()
        ps.step()
        opt.zero_grad()
        mse_loss(ref(x), y).backward()
        n = torch.nn.utils.clip_grad_norm_(ref.parameters(), 0.05)
        opt.step()
        assert abs(ps.grad_norm().item() - n.item()) < 1e-5
    for a, b in zip(m.parameters(), ref.parameters()):
        assert torch.allclose(a, b, atol=2e-6, rtol=1e-5)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, kind, broadcast, q, alt=False, split_mb=None, big=False):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(0)
    cfg = MLPConfig(d_in=256, hidden=(512,), d_out=64) if big else CFG
    m = MLP(cfg)
    ref = copy.deepcopy(m)
    ps = ShardedParameterServer(list(m.named_parameters()), OptimConfig(kind=kind, lr=0.01),
                                world=world, rank=rank, bucket_mb=0.3 if big else 0.005,
                                broadcast=broadcast,
                                alt_params=alt, **({} if split_mb is None else {"split_mb": split_mb}))
    if split_mb:
        pieces = sum(len(getattr(s_, "pieces", None) or []) for s_ in ps.plan.slots.values())
        assert pieces >= 2, "the row-split path was not taken"
    fwd = _SinkMLP(m) if alt else m  # with two replicas: sink layers that commit before dX
    opt = _torch_opt(kind, list(ref.parameters()))
    for g in opt.param_groups:
        g["weight_decay"] = 0.0
    gen = torch.Generator().manual_seed(123)
    for _ in range(3):
        xs = torch.randn(world, 8, cfg.d_in, generator=gen)
        ys = torch.randn(world, 8, cfg.d_out, generator=gen)
        mse_loss(fwd(xs[rank]), ys[rank]).backward()
        ps.step()
        # oracle: mean over workers of the per-worker gradients
        opt.zero_grad()
        for w in range(world):
            (mse_loss(ref(xs[w]), ys[w]) / world).backward()
        opt.step()
    err = max((a - b).abs().max().item() for a, b in zip(m.parameters(), ref.parameters()))
    q.put((rank, err))
    dist.destroy_process_group()


@pytest.mark.parametrize("kind,broadcast,alt", [("sgd", True, False), ("adam", True, False),
                                                ("momentum", False, False), ("sgd", True, True),
                                                ("adam", True, True)])
def test_gloo_two_workers_match_mean_gradient(kind, broadcast, alt):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, kind, broadcast, q, alt))
             for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, err in res:
        assert err < 5e-6, (rank, err)


def test_gloo_four_workers_with_row_split_buckets():
    """world = 4 with the first-produced tensors cut into row blocks (what the GPU engine does
    from 4 ranks up so that the last bucket's push overlaps its own dW GEMM): same result as the
    mean-gradient oracle."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 4, port, "adam", True, q, False, 0.125, True))
             for r in range(4)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, err in res:
        assert err < 5e-6, (rank, err)


def test_async_engine_host_emulation_matches_sequential_application():
    """Host emulation of the async mode: each worker's gradient is applied on its own, scaled by
    1/N — for plain SGD that equals the synchronous mean update."""
    from parameter_server_distributed_b200.parallel.async_engine import AsyncParameterServer
    torch.manual_seed(2)
    m = MLP(CFG)
    ref = copy.deepcopy(m)
    ps = AsyncParameterServer(list(m.named_parameters()), OptimConfig(kind="sgd", lr=0.05),
                              bucket_mb=0.005, staleness=1)
    assert ps.ring == 2 and ps.staleness == 1
    opt = torch.optim.SGD(ref.parameters(), lr=0.05)
    for _ in range(3):
        x, y = torch.randn(16, 32), torch.randn(16, 8)
        mse_loss(m(x), y).backward()
        ps.step()
        opt.zero_grad()
        mse_loss(ref(x), y).backward()
        opt.step()
    assert ps.clock == 3
    for a, b in zip(m.parameters(), ref.parameters()):
        assert torch.allclose(a, b, atol=1e-6)
    ps.drain()  # no-op on the host


def test_engine_metrics_and_prometheus_text():
    torch.manual_seed(4)
    m = MLP(CFG)
    ps = ShardedParameterServer(list(m.named_parameters()), OptimConfig(kind="sgd", lr=0.1),
                                bucket_mb=0.005)
    for _ in range(2):
        mse_loss(m(torch.randn(4, 32)), torch.randn(4, 8)).backward()
        ps.step()
    snap = ps.metrics.snapshot()
    assert snap["steps"] == 2 and snap["pushes"] == 2 * len(ps.plan.buckets)
    assert snap["nvlink_in_bytes"] == 0  # one rank: nothing crosses a link
    text = ps.metrics.prometheus()
    assert "# TYPE psb_steps counter" in text and "psb_pushes " in text


class _SinkLinear(torch.autograd.Function):
    """A layer that uses the engine's gradient sinks the way ops/linear.py does on the GPU:
    parameter gradients are written into the bucket and committed after dX is computed."""

    @staticmethod
    def forward(ctx, x, w, b):
        from parameter_server_distributed_b200.ops.linear import grad_to_sink  # noqa: F401
        ctx.save_for_backward(x, w)
        ctx.set_materialize_grads(False)
        from parameter_server_distributed_b200.ops.linear import take_sink
        ctx.sinks = (take_sink(w, ctx.needs_input_grad[1]), take_sink(b, ctx.needs_input_grad[2]))
        return x @ w.t() + b

    @staticmethod
    def backward(ctx, gy):
        from parameter_server_distributed_b200.ops.linear import grad_to_sink
        x, w = ctx.saved_tensors
        sw, sb = ctx.sinks
        live = [s for s in (sw, sb) if s is not None and s.exclusive]
        early = bool(live) and all(s.early for s in live)
        if not early:
            gx = gy @ w  # before the commits: a commit may start the in-place update of w
        gb = grad_to_sink(sb, lambda out: torch.sum(gy, 0, out=out))
        gw = grad_to_sink(sw, lambda out: torch.mm(gy.t(), x, out=out))
        if early:
            gx = gy @ w  # the push went to the other replica (or only reduced): w is intact
        return gx, gw, gb


class _SinkMLP(torch.nn.Module):
    def __init__(self, base: MLP):
        super().__init__()
        self.layers = base.layers

    def forward(self, x):
        for i, layer in enumerate(self.layers):
            x = _SinkLinear.apply(x, layer.weight, layer.bias)
            if i + 1 < len(self.layers):
                x = torch.nn.functional.gelu(x, approximate="tanh")
        return x


def test_gradient_sinks_push_once_per_bucket_and_match_torch():
    torch.manual_seed(5)
    m = MLP(CFG)
    ref = copy.deepcopy(m)
    ps = ShardedParameterServer(list(m.named_parameters()), OptimConfig(kind="sgd", lr=0.05),
                                bucket_mb=0.005)
    assert all(hasattr(p, "_psb_sink") for p in m.parameters())
    sm = _SinkMLP(m)
    opt = torch.optim.SGD(ref.parameters(), lr=0.05)
    nb = len(ps.plan.buckets)
    for it in range(3):
        x, y = torch.randn(16, 32), torch.randn(16, 8)
        mse_loss(sm(x), y).backward()
        assert all(p.grad is None for p in m.parameters())  # delivered through the sinks
        ps.step()
        assert ps.metrics.counters["pushes"] == nb * (it + 1)
        opt.zero_grad()
        mse_loss(ref(x), y).backward()
        opt.step()
    for a, b in zip(m.parameters(), ref.parameters()):
        assert torch.allclose(a, b, atol=2e-6, rtol=1e-5)
    ps.close()
    assert not any(hasattr(p, "_psb_sink") for p in m.parameters())


def test_gradient_sinks_with_tied_weights_fall_back_to_autograd():
    """A parameter used by two layer calls in one step needs its gradients added: the sink is
    not exclusive then and the ordinary hook path (autograd accumulation) delivers it once."""
    torch.manual_seed(6)
    lin = torch.nn.Linear(16, 16)
    ref = copy.deepcopy(lin)
    ps = ShardedParameterServer(list(lin.named_parameters()), OptimConfig(kind="sgd", lr=0.1),
                                bucket_mb=0.005)
    opt = torch.optim.SGD(ref.parameters(), lr=0.1)
    for it in range(2):
        x = torch.randn(8, 16)
        y = _SinkLinear.apply(torch.tanh(_SinkLinear.apply(x, lin.weight, lin.bias)),
                              lin.weight, lin.bias)
        y.square().mean().backward()
        ps.step()
        assert ps.metrics.counters["pushes"] == len(ps.plan.buckets) * (it + 1)
        opt.zero_grad()
        ref(torch.tanh(ref(x))).square().mean().backward()
        opt.step()
    for a, b in zip(lin.parameters(), ref.parameters()):
        assert torch.allclose(a, b, atol=2e-6, rtol=1e-5)


@pytest.mark.parametrize("clip", [None, 0.05])
def test_alternating_replicas_allow_commit_before_dx(clip):
    """With two parameter replicas (or in clip mode, where a flush only reduces) the sinks are
    `early`: the layer commits dW before it computes dX from the same weights.  The host
    backend applies a push synchronously, so any aliasing of the two replicas would corrupt dX
    and show up against torch."""
    torch.manual_seed(8)
    m = MLP(CFG)
    ref = copy.deepcopy(m)
    ps = ShardedParameterServer(list(m.named_parameters()), OptimConfig(kind="sgd", lr=0.05),
                                bucket_mb=0.005, alt_params=clip is None, clip_norm=clip)
    assert all(p._psb_sink.early for p in m.parameters())
    sm = _SinkMLP(m)
    opt = torch.optim.SGD(ref.parameters(), lr=0.05)
    ptrs = []
    for it in range(4):
        x, y = torch.randn(16, 32), torch.randn(16, 8)
        mse_loss(sm(x), y).backward()
        ps.step()
        ptrs.append(m.layers[1].weight.data_ptr())
        opt.zero_grad()
        mse_loss(ref(x), y).backward()
        if clip is not None:
            torch.nn.utils.clip_grad_norm_(ref.parameters(), clip)
        opt.step()
    for a, b in zip(m.parameters(), ref.parameters()):
        assert torch.allclose(a, b, atol=2e-6, rtol=1e-5)
    if clip is None:  # the model reads the two replicas in turn
        assert ptrs[0] == ptrs[2] and ptrs[1] == ptrs[3] and ptrs[0] != ptrs[1]
        # state written back (checkpoint restore) lands in the replica the model reads
        state = ps.gather_state("master")
        ps.scatter_state({k: v * 0 + 1.5 for k, v in state.items()}, "master")
        assert all(torch.all(p == 1.5) for p in m.parameters())
