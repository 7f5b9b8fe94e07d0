"""One rank of the fault-injection test (tests/test_elastic_gpu.py).  Started directly (not by
torchrun, which would tear the whole group down when one worker dies) with RANK / LOCAL_RANK /
WORLD_SIZE / PSB_COORDINATOR in the environment, the way scripts/deploy_box.sh MODE=gpu does.

phase "train":  train with periodic checkpoints; rank DIE_RANK kills itself at step DIE_STEP;
                the survivors must notice (PeerLost), leave cleanly and print ELASTIC_LEFT.
phase "resume": a smaller world loads the checkpoint, checks it and trains on.
"""
import os
import signal
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch

import parameter_server_distributed_b200 as psd
from parameter_server_distributed_b200.models import MLP, MLPConfig, mse_loss
from parameter_server_distributed_b200.utils.checkpoint import read_reference_file
from parameter_server_distributed_b200.utils.elastic import PeerLost


def main():
    phase, ckpt = sys.argv[1], sys.argv[2]
    die_rank, die_step = int(os.environ.get("DIE_RANK", "-1")), int(os.environ.get("DIE_STEP", "0"))
    ctx = psd.init(heap_bytes=128 << 20)
    torch.manual_seed(0)
    cfg = MLPConfig(d_in=256, hidden=(512, 512), d_out=128)
    model = MLP(cfg).to(ctx.device, torch.bfloat16)
    tr = psd.Trainer(model, mse_loss, ctx, psd.OptimConfig(kind="adamw", lr=1e-3), bucket_mb=0.5,
                     checkpoint_every=4 if phase == "train" else 0,
                     checkpoint_path=ckpt if phase == "train" else None)
    gen = torch.Generator().manual_seed(10 + ctx.rank)
    if phase == "resume":
        epoch, it = tr.load_checkpoint(ckpt)
        _, _, tensors = read_reference_file(ckpt)
        got = tr.ps.gather_state("master")
        for name, shape, data in tensors:
            exp = torch.from_numpy(data).view(*shape).to(ctx.device)
            assert torch.equal(got[name], exp), f"{name}: shard state differs from the file"
            p = dict(model.named_parameters())[name]
            assert torch.equal(p.detach(), exp.to(torch.bfloat16)), f"{name}: replica differs"
        assert tr.ps.gather_state("m"), "optimizer moments were not restored"
        losses = []
        for _ in range(5):
            x = torch.randn(128, cfg.d_in, generator=gen).to(torch.bfloat16).pin_memory()
            y = torch.randn(128, cfg.d_out, generator=gen).to(torch.bfloat16).pin_memory()
            losses.append(tr.step(x, y).item())
        assert all(l == l and l < 100 for l in losses), losses
        print(f"ELASTIC_RESUMED rank={ctx.rank} world={ctx.world} epoch={epoch} iteration={it} "
              f"loss={losses[-1]:.4f}", flush=True)
        ctx.close()
        return 0
    step = 0
    try:
        while step < 10_000:
            if ctx.rank == die_rank and step == die_step:
                os.kill(os.getpid(), signal.SIGKILL)  # no goodbye: heartbeats just stop
            x = torch.randn(128, cfg.d_in, generator=gen).to(torch.bfloat16).pin_memory()
            y = torch.randn(128, cfg.d_out, generator=gen).to(torch.bfloat16).pin_memory()
            tr.step(x, y).item()
            step += 1
        print(f"ELASTIC_NEVER_NOTICED rank={ctx.rank}", flush=True)
        return 3
    except PeerLost as e:
        alive = ctx.coord.list_workers()
        # the CUDA context is still usable: kernels gave up waiting instead of trapping
        torch.cuda.synchronize()
        probe = float((torch.ones(8, device=ctx.device) * 2).sum().item())
        print(f"ELASTIC_LEFT rank={ctx.rank} step={step} alive={alive} probe={probe} :: {e}",
              flush=True)
        ctx.close()
        return 0


if __name__ == "__main__":
    sys.exit(main())
