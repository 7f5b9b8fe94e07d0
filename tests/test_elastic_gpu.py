"""T7 on the GPU plane: a rank dies mid-run -> the coordinator evicts it -> the survivors'
kernels stop waiting for it (abort word), Trainer.step raises PeerLost, they leave cleanly with
their CUDA contexts intact, and a smaller world resumes from the last periodic checkpoint.

The reference has no counterpart: its survivors poll for the dead worker's iteration and give
up on every step for ever after (/root/reference/src/worker.cpp:372-401)."""
import os
import re
import subprocess
import sys

import pytest
import torch

from parameter_server_distributed_b200.control import client as cc

HERE = os.path.dirname(os.path.abspath(__file__))
WORKER = os.path.join(HERE, "_elastic_worker.py")


def _launch(world, addr, phase, ckpt, extra_env=None):
    procs = []
    for r in range(world):
        env = dict(os.environ, RANK=str(r), LOCAL_RANK=str(r), WORLD_SIZE=str(world),
                   PSB_COORDINATOR=addr, PSB_HEARTBEAT_MS="100", PSB_SPIN_TIMEOUT_S="60",
                   # peer-load kernels: the survivors keep a mapping of the dead rank's segment, so
                   # their loads stay legal.  A multicast team does not survive the death of a
                   # member (its binding goes away under running multimem kernels and the survivors
                   # get a CUDA error instead of PeerLost) — jobs that want the clean exit run so.
                   PSB_NVLS="0")
        env.update(extra_env or {})
        procs.append(subprocess.Popen([sys.executable, WORKER, phase, ckpt], env=env,
                                      stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = []
    for p in procs:
        try:
            out, _ = p.communicate(timeout=240)
        except subprocess.TimeoutExpired:
            p.kill()
            out, _ = p.communicate()
            out += "\n<timeout>"
        outs.append((p.returncode, out))
    return outs


@pytest.mark.gpu
@pytest.mark.multigpu
def test_rank_dies_survivors_leave_cleanly_and_a_smaller_world_resumes(tmp_path):
    n = min(torch.cuda.device_count(), 4)
    ckpt = str(tmp_path / "elastic.ckpt")
    fast = {"PSB_STALE_SCAN_MS": "100", "PSB_STALE_TIMEOUT_S": "1"}
    proc, addr = cc.spawn_coordinator(env=fast, log_path=str(tmp_path / "co.log"))
    try:
        die = n - 1
        outs = _launch(n, addr, "train", ckpt, {"DIE_RANK": str(die), "DIE_STEP": "14"})
        assert outs[die][0] == -9, outs[die]
        for r in range(n - 1):
            rc, out = outs[r]
            m = re.search(r"ELASTIC_LEFT rank=%d step=(\d+) alive=\[([\d, ]*)\] probe=16.0" % r, out)
            assert rc == 0 and m, (r, rc, out[-1500:])
            assert int(m.group(1)) >= 14 and str(die) not in m.group(2).split(", ")
        assert os.path.exists(ckpt) and os.path.exists(ckpt + ".psb")
    finally:
        cc._stop(proc)
    # the job comes back with the ranks that are left
    proc, addr = cc.spawn_coordinator(env=fast, log_path=str(tmp_path / "co2.log"))
    try:
        outs = _launch(n - 1, addr, "resume", ckpt)
        for r, (rc, out) in enumerate(outs):
            assert rc == 0 and f"ELASTIC_RESUMED rank={r} world={n - 1} epoch=3" in out, \
                (r, rc, out[-1500:])
    finally:
        cc._stop(proc)
