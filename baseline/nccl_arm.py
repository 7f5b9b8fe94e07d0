"""`bench.py --impl nccl`: the bar BASELINE.json sets — "a path that only calls NCCL for push/pull
is the baseline, not the product".

The same workload as the product arm, built the way a competent user builds a sharded parameter
server out of library calls only: ``torch.nn`` layers (cuBLAS GEMMs, eager elementwise kernels),
parameters and gradients in flat bf16 buckets, and per bucket, on a communication stream that
overlaps the rest of backward,

    ncclReduceScatter(AVG)  ->  fused torch.optim step on the rank's fp32 master shard
                            ->  bf16 cast  ->  ncclAllGather into every replica

i.e. ZeRO-1 expressed as push (reduce-scatter + server update) and pull (all-gather).  None of
this repo's kernels, engine or fabric is on this path; it prints the same JSON line as the
product arm with ``"impl": "nccl"``.  (The reference's own NCCL usage — one blocking
``ncclAllReduce`` per tensor on the null stream with host staging, src/worker.cpp:408-448 — is
the `--impl reference` arm; this arm is the strong baseline SURVEY §6 calls B1.)
"""
from __future__ import annotations

import json
import os
import sys
import time


class NcclShardedPS:
    """Flat-bucket ZeRO-1 over NCCL collectives."""

    def __init__(self, model, lr: float, kind: str, bucket_mb: float, device, overlap: bool = True):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.device = device
        self.overlap = overlap
        params = [p for p in model.parameters() if p.requires_grad]
        order = params[::-1]  # backward produces the last layers' gradients first
        limit = int(bucket_mb * (1 << 20) / 2)
        quantum = self.world * 512
        self.buckets = []  # (params, offsets, numel_padded)
        cur, fill = [], 0
        for p in order:
            if cur and fill + p.numel() > limit:
                self.buckets.append((cur, -(-fill // quantum) * quantum))
                cur, fill = [], 0
            cur.append((p, fill))
            fill += -(-p.numel() // 128) * 128
        if cur:
            self.buckets.append((cur, -(-fill // quantum) * quantum))
        total = sum(n for _, n in self.buckets)
        self.flat_p = torch.zeros(total, dtype=torch.bfloat16, device=device)
        self.flat_g = torch.zeros(total, dtype=torch.bfloat16, device=device)
        self.masters, self.shard_g, self.bucket_of = [], [], {}
        base = 0
        with torch.no_grad():
            for b, (plist, n) in enumerate(self.buckets):
                for p, off in plist:
                    view = self.flat_p[base + off: base + off + p.numel()].view(p.shape)
                    view.copy_(p)
                    p.data = view
                    self.bucket_of[p] = (b, base + off)
                s = n // self.world
                m = self.flat_p[base + self.rank * s: base + (self.rank + 1) * s].float()
                self.masters.append(torch.nn.Parameter(m))
                self.shard_g.append(torch.empty(s, dtype=torch.bfloat16, device=device))
                self.buckets[b] = (plist, n, base)
                base += n
        if kind == "sgd":
            self.opts = [torch.optim.SGD([m], lr=lr, foreach=True) for m in self.masters]
        else:
            self.opts = [torch.optim.AdamW([m], lr=lr, weight_decay=0.0, fused=True) for m in self.masters]
        self.pending = [0] * len(self.buckets)
        self.comm = torch.cuda.Stream(device=device, priority=-1)
        self.events = [torch.cuda.Event() for _ in self.buckets]
        for p in params:
            p.register_post_accumulate_grad_hook(self._hook)

    def _hook(self, p):
        b, off = self.bucket_of[p]
        self.flat_g[off: off + p.numel()].view(p.shape).copy_(p.grad)
        p.grad = None
        self.pending[b] += 1
        if self.pending[b] == len(self.buckets[b][0]):
            self._push_pull(b)

    def _push_pull(self, b):
        torch, dist = self.torch, self.dist
        _, n, base = self.buckets[b]
        s = n // self.world
        cur = torch.cuda.current_stream(self.device)
        stream = self.comm if self.overlap else cur
        if self.overlap:
            self.events[b].record(cur)
            self.comm.wait_event(self.events[b])
        with torch.cuda.stream(stream):
            g = self.flat_g[base: base + n]
            if self.world > 1:
                dist.reduce_scatter_tensor(self.shard_g[b], g, op=dist.ReduceOp.AVG)
                sg = self.shard_g[b]
            else:
                sg = g
            self.masters[b].grad = sg.float()
            self.opts[b].step()
            mine = self.flat_p[base + self.rank * s: base + (self.rank + 1) * s]
            mine.copy_(self.masters[b].detach())
            if self.world > 1:
                dist.all_gather_into_tensor(self.flat_p[base: base + n], mine)

    def step(self):
        for b in range(len(self.buckets)):
            if self.pending[b] and self.pending[b] != len(self.buckets[b][0]):
                self._push_pull(b)  # parameters that took no part: push what arrived
            self.pending[b] = 0
        if self.overlap:
            self.torch.cuda.current_stream(self.device).wait_stream(self.comm)


def run_nccl(args) -> int:
    import torch
    import torch.distributed as dist

    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from bench import Workload
    from parameter_server_distributed_b200.utils import ClockSampler

    if not torch.cuda.is_available():
        print(json.dumps({"impl": "nccl", "error": "no CUDA device"}))
        return 1
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    # same host placement as the other arm: pinned batches next to the GPU
    from parameter_server_distributed_b200.utils.numa import bind_to_gpu_node
    numa = bind_to_gpu_node(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=device)
    # the library path end to end: no tcgen05 links, no fused loss kernels of this repo
    os.environ["PSB_TCGEN05_LINEAR"] = "0"
    from parameter_server_distributed_b200.ops import linear as _linear
    _linear.ENABLED = False
    wl = Workload(args.config, args.batch_per_gpu)
    torch.manual_seed(1234)
    model = wl.build().to(device, torch.bfloat16)
    if wl.memory_format is not None:
        model = model.to(memory_format=wl.memory_format)
    lr = 1e-3 if args.optimizer != "sgd" else 1e-2
    ps = NcclShardedPS(model, lr, args.optimizer, args.bucket_mb, device, overlap=not args.no_overlap)

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def barrier():
        torch.cuda.synchronize(device)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(device)

    gen = torch.Generator().manual_seed(100 + rank)
    n_host = 4
    batches = [wl.make(gen) for _ in range(n_host)]
    host = [(b[0].pin_memory(), b[1].pin_memory()) for b in batches]
    dev = [(x.to(device), y.to(device)) for x, y in host]

    def step_device(x, y):
        loss = wl.loss(model(x), y)
        loss.backward()
        ps.step()
        return loss.detach()

    for i in range(args.warmup):
        step_device(*dev[i % n_host])
    barrier()
    with ClockSampler(local) as clocks:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(args.steps):
            loss = step_device(*dev[i % n_host])
        e1.record()
        barrier()
        dev_ms = e0.elapsed_time(e1)
    dev_ms = max_over_ranks(dev_ms)
    final_loss = float(loss.item())

    # push + pull alone (gradients already in the buckets)
    barrier()
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    p0.record()
    for _ in range(args.steps):
        for b in range(len(ps.buckets)):
            ps._push_pull(b)
        ps.step()
    p1.record()
    barrier()
    pp_ms = max_over_ranks(p0.elapsed_time(p1)) / args.steps

    # end to end: pinned host batch -> device, loss -> host every step
    copy_stream = torch.cuda.Stream(device=device)
    bufs = [[torch.empty_like(t, device=device) for t in host[0]] for _ in range(2)]
    loss_host = [torch.zeros((), dtype=torch.float32).pin_memory() for _ in range(4)]

    def step_e2e(i, x, y):
        slot = i & 1
        cur = torch.cuda.current_stream(device)
        copy_stream.wait_stream(cur)
        with torch.cuda.stream(copy_stream):
            bufs[slot][0].copy_(x, non_blocking=True)
            bufs[slot][1].copy_(y, non_blocking=True)
        cur.wait_stream(copy_stream)
        loss = wl.loss(model(bufs[slot][0]), bufs[slot][1])
        h = loss_host[i % 4]
        h.copy_(loss.detach().float(), non_blocking=True)
        ev = torch.cuda.Event()
        ev.record(cur)
        loss.backward()
        ps.step()
        return h, ev

    for i in range(args.warmup):
        h, ev = step_e2e(i, *host[i % n_host])
        ev.synchronize()
    barrier()
    t0 = time.perf_counter()
    s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s0.record()
    prev = None
    for i in range(args.steps):
        cur = step_e2e(i, *host[i % n_host])
        if prev is not None:
            prev[1].synchronize()
            float(prev[0].item())
        prev = cur
    prev[1].synchronize()
    e2e_loss = float(prev[0].item())
    s1.record()
    torch.cuda.synchronize(device)  # same clock as the other arm: local completion, then max over ranks
    wall_ms = (time.perf_counter() - t0) * 1e3
    barrier()
    e2e_ms = max_over_ranks(max(s0.elapsed_time(s1), wall_ms))

    gb = wl.batch * world
    if rank == 0:
        h2d = sum(t.numel() * t.element_size() for t in host[0])
        print(json.dumps({
            "metric": "samples/sec (training through parameter-server push/pull)",
            "value": gb * args.steps / (dev_ms / 1e3), "unit": "samples/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": dev_ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
            "data": "synthetic (random batches, random-init weights)", "impl": "nccl",
            "config": {"model": f"{wl.label} ({wl.nparams / 1e6:.1f}M params)", "global_batch": gb,
                       "seq_len": wl.seq_len, "parallelism": f"dp{world} / {world} PS shards (ZeRO-1)",
                       "optimizer": args.optimizer, "bucket_mb": args.bucket_mb,
                       "layer_gemms": "library (cuBLAS)", "numa_bind": numa,
                       "push_pull": "ncclReduceScatter + torch.optim on the fp32 shard + ncclAllGather, "
                                    "bucketed, overlapped with backward",
                       "l2": "working set exceeds the 126 MB L2; inputs rotate over 4 batches"},
            "push_pull_ms_per_step": pp_ms,
            "e2e": {"value": gb * args.steps / (e2e_ms / 1e3), "unit": "samples/s",
                    "ms_per_step": e2e_ms / args.steps, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": 4},
            "gpu_launches": 0, "clocks": clocks.summary(),
            "final_loss": final_loss, "e2e_final_loss": e2e_loss}), flush=True)
    barrier()
    if world > 1:
        dist.destroy_process_group()
    return 0
