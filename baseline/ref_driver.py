"""Reference arm: train the same synthetic MLP through the UNMODIFIED reference code.

What a user of araju6/parameter-server-distributed gets on this box: a host-resident
``ParameterServerCore`` (fp32 ``std::vector`` tensors, ``p -= mean(g)``) and, when more than
one GPU is visible, the worker's ``NCCLManager`` aggregation of every gradient tensor through
host staging.  The reference ships no model, so the forward/backward is a stock
``torch.nn`` MLP of the benchmark's shape (none of this repo's models, kernels or engine is
on this path).  Per iteration, as in ``Worker::run_iteration`` (src/worker.cpp:331-406):

    pull (serve_parameters -> host -> device) -> forward/backward -> gradients to host ->
    [N > 1: aggregate_gradients_multi_gpu pattern over NCCLManager] -> push (receive_gradients,
    which aggregates and applies because total_workers = 1)

The reference drives all GPUs from ONE process and feeds every GPU the same gradient
(src/worker.cpp:420-425), so extra GPUs add communication but no samples: global batch =
per-GPU batch.  Under torchrun only rank 0 works; the other ranks exit at once.
"""
from __future__ import annotations

import ctypes as C
import json
import os
import subprocess
import sys
import time

import numpy as np


def _arrays(names, shapes, bufs):
    n = len(names)
    c_names = (C.c_char_p * n)(*[s.encode() for s in names])
    ndims = (C.c_int * n)(*[len(s) for s in shapes])
    flat = [d for s in shapes for d in s]
    dims = (C.c_int * len(flat))(*flat)
    ptrs = (C.c_void_p * n)(*[b.ctypes.data for b in bufs])
    sizes = (C.c_longlong * n)(*[b.size for b in bufs])
    return n, c_names, ndims, dims, ptrs, sizes


def _probe_ungrouped(lib_path: str, n_gpus: int) -> bool:
    """Does the reference's un-grouped all-reduce loop complete on this NCCL?  """
    code = (
        "import ctypes as C, numpy as np, sys\n"
        f"lib = C.CDLL({lib_path!r})\n"
        "lib.ref_nccl_create.restype = C.c_void_p\n"
        "m = lib.ref_nccl_create()\n"
        f"assert lib.ref_nccl_init(C.c_void_p(m), {n_gpus}) == 1\n"
        "x = np.ones(1024, dtype=np.float32)\n"
        "rc = lib.ref_nccl_aggregate(C.c_void_p(m), C.c_void_p(x.ctypes.data), C.c_longlong(1024), 0)\n"
        "sys.exit(0 if rc == 1 and abs(float(x[0]) - 1.0) < 1e-6 else 3)\n")
    try:
        r = subprocess.run([sys.executable, "-c", code], timeout=60, capture_output=True)
        return r.returncode == 0
    except subprocess.TimeoutExpired:
        return False


def run(args, lib_path: str) -> int:
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0  # single-process multi-GPU design: rank 0 drives every GPU
    import torch
    if not torch.cuda.is_available():
        print(json.dumps({"impl": "reference", "unavailable": "no CUDA device"}))
        return 0
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from bench import model_config  # shapes only
    if args.config not in ("mlp", "mlp-small"):
        print(json.dumps({"impl": "reference", "unavailable":
                          f"reference arm implements the flagship MLP config only, not {args.config}"}))
        return 0
    cfg = model_config(args.config)
    n_gpus = min(args.gpus, torch.cuda.device_count())

    use_group = 0
    if n_gpus > 1:
        forced = os.environ.get("REF_NCCL_GROUP")
        if forced is not None:
            use_group = int(forced)
        elif not _probe_ungrouped(lib_path, n_gpus):
            use_group = 1

    lib = C.CDLL(lib_path)
    lib.ref_ps_create.restype = C.c_void_p
    lib.ref_nccl_create.restype = C.c_void_p
    ps = C.c_void_p(lib.ref_ps_create(1))
    mgr = None
    if n_gpus > 1:
        mgr = C.c_void_p(lib.ref_nccl_create())
        if lib.ref_nccl_init(mgr, n_gpus) != 1:
            print(json.dumps({"impl": "reference",
                              "unavailable": "NCCLManager::initialize failed"}))
            return 0

    dev = torch.device("cuda:0")
    torch.manual_seed(1234)
    dims = cfg.dims
    layers = []
    for i in range(len(dims) - 1):
        layers.append(torch.nn.Linear(dims[i], dims[i + 1], bias=cfg.bias))
        if i + 2 < len(dims):
            layers.append(torch.nn.GELU(approximate="tanh"))
    model = torch.nn.Sequential(*layers).to(dev, torch.bfloat16)
    params = list(model.named_parameters())
    names = [n for n, _ in params]
    shapes = [list(p.shape) for _, p in params]
    host_p = [np.ascontiguousarray(p.detach().float().cpu().numpy()).reshape(-1) for _, p in params]
    host_g = [np.zeros_like(b) for b in host_p]
    lib.ref_ps_init(ps, *_arrays(names, shapes, host_p))
    push_args = _arrays(names, shapes, host_g)
    pull_ptrs = (C.c_void_p * len(names))(*[b.ctypes.data for b in host_p])
    pull_sizes = (C.c_longlong * len(names))(*[b.size for b in host_p])

    B = args.batch_per_gpu or 4096
    gen = torch.Generator().manual_seed(100)
    hx = [torch.randn(B, cfg.d_in, generator=gen).to(torch.bfloat16).pin_memory() for _ in range(4)]
    hy = [torch.randn(B, cfg.d_out, generator=gen).to(torch.bfloat16).pin_memory() for _ in range(4)]

    def step(it: int) -> float:
        # pull: ParameterServerCore::serve_parameters -> host -> device
        lib.ref_ps_pull(ps, it, len(names), pull_ptrs, pull_sizes)
        with torch.no_grad():
            for (_, p), b in zip(params, host_p):
                p.copy_(torch.from_numpy(b).view(p.shape))
        x = hx[it % 4].to(dev, non_blocking=True)
        y = hy[it % 4].to(dev, non_blocking=True)
        model.zero_grad(set_to_none=True)
        loss = torch.nn.functional.mse_loss(model(x).float(), y.float())
        loss.backward()
        for (_, p), g in zip(params, host_g):
            g[:] = p.grad.detach().float().cpu().numpy().reshape(-1)
        if mgr is not None:  # Worker::aggregate_gradients_multi_gpu, per tensor
            for g in host_g:
                rc = lib.ref_nccl_aggregate(mgr, C.c_void_p(g.ctypes.data), C.c_longlong(g.size),
                                            use_group)
                if rc != 1:
                    raise RuntimeError(f"reference NCCL aggregation failed (rc={rc})")
        # push: ParameterServerCore::receive_gradients (aggregates + applies, 1 worker)
        lib.ref_ps_push(ps, 0, it, *push_args)
        return float(loss.item())

    it = 0
    for _ in range(args.warmup):
        step(it)
        it += 1
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        loss = step(it)
        it += 1
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    nparams = sum(b.size for b in host_p)
    out = {
        "metric": "samples/sec (training through parameter-server push/pull)",
        "value": B * args.steps / dt, "unit": "samples/s", "n_gpus": n_gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16 compute, fp32 parameter server",
        "data": "synthetic (random batches, random-init weights)", "impl": "reference",
        "config": {"model": f"MLP {'-'.join(map(str, dims))} ({nparams / 1e6:.1f}M params)",
                   "global_batch": B, "seq_len": 1,
                   "parallelism": f"1 worker process driving {n_gpus} GPU(s), 1 host PS",
                   "nccl_group_fix": bool(use_group),
                   "note": "reference replicates one gradient to every GPU (worker.cpp:420-425): "
                           "extra GPUs add no samples"},
        "e2e": {"value": B * args.steps / dt, "unit": "samples/s",
                "h2d_bytes_per_step": nparams * 2 + hx[0].numel() * 2 + hy[0].numel() * 2,
                "d2h_bytes_per_step": nparams * 4 + 4},
        "final_loss": loss,
    }
    print(json.dumps(out), flush=True)
    return 0
