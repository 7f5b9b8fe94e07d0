"""T5 on real peers: launches tests/_mp_worker.py with one rank per GPU."""
import os
import subprocess
import sys

import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.gpu
@pytest.mark.multigpu
def test_cross_rank_push_pull_and_fused_gemm():
    n = min(torch.cuda.device_count(), 8)
    n = 8 if n >= 8 else 4 if n >= 4 else 2
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
           "--master-addr", "127.0.0.1", "--master-port", "29613",
           os.path.join(HERE, "_mp_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=240)
    ok = r.stdout.count("MP_OK")
    fails = [l for l in r.stdout.splitlines() if "MP_FAIL" in l]
    assert r.returncode == 0 and ok == n, f"rc={r.returncode} ok={ok}\n" + "\n".join(fails[:8]) + \
        f"\n{r.stderr[-1500:]}"
