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


def _run_sparse(n, env=None):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
           "--master-addr", "127.0.0.1", "--master-port", "29617",
           os.path.join(HERE, "_mp_sparse_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=240, env=env)
    ok = r.stdout.count("MP_OK")
    fails = [l for l in r.stdout.splitlines() if "MP_FAIL" in l]
    assert r.returncode == 0 and ok == n, f"rc={r.returncode} ok={ok}\n" + "\n".join(fails[:8]) + \
        f"\n{r.stderr[-1500:]}"


@pytest.mark.gpu
@pytest.mark.multigpu
def test_sparse_embedding_push_pull_across_real_peers():
    """K5 over NVLink: every rank stages its rows, the owners read them from the peers' heaps,
    every rank pulls every owner's rows (tests/_mp_sparse_worker.py)."""
    n = min(torch.cuda.device_count(), 8)
    _run_sparse(8 if n >= 8 else 4 if n >= 4 else 2)


def test_sparse_worker_script_on_three_cpu_ranks():
    """The same script over gloo (host path): its batches, oracle and assertions are exercised
    here, so that the multi-GPU run tests the kernels and not the script."""
    _run_sparse(3, env=dict(os.environ, CUDA_VISIBLE_DEVICES=""))
