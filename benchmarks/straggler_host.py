#!/usr/bin/env python
"""Straggler benchmark of the host plane: sync (BSP) vs SSP vs async on the C++ binaries.

The reference is BSP only — every iteration waits for the slowest worker
(/root/reference/src/parameter_server.cpp:34-38 completes an iteration when
``gradients.size() == total_workers``; /root/reference/src/worker.cpp:366-395 polls until it
does).  ``parameter_server --mode=ssp|async`` removes that barrier; this script shows the case
those modes exist for.  One coordinator, one parameter server and W ``worker_main`` processes
run on this machine; a worker's compute step is simulated (``--compute-ms``), so the numbers
are about the barrier, not about the model.

Scenarios
  persistent   worker W-1 is 10 % slower than the others on every iteration
  transient    every worker's iteration takes 5x as long with probability 0.1
               (the same per-worker sequence in every mode)

Reported per mode: wall time until the last worker finished K iterations, the sum of the
workers' own rates (iterations/s each worker sustained while it ran), and both relative to sync.

    python benchmarks/straggler_host.py [--workers 4] [--iters 400] [--compute-ms 5] [--out FILE]
"""
import argparse
import json
import os
import re
import socket
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "build", "bin")


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def wait_port(port, timeout=10.0):
    end = time.time() + timeout
    while time.time() < end:
        try:
            socket.create_connection(("127.0.0.1", port), timeout=0.2).close()
            return True
        except OSError:
            time.sleep(0.02)
    return False


def run(mode, workers, iters, compute_ms, scenario, staleness, tmp, ps_flags=()):
    ps_port, co_port = free_port(), free_port()
    logs = []

    def spawn(argv, name):
        f = open(os.path.join(tmp, f"{mode}_{scenario}_{name}.log"), "w")
        logs.append(f)
        return subprocess.Popen(argv, stdout=f, stderr=subprocess.STDOUT, cwd=tmp)

    co = spawn([f"{BIN}/coordinator", f"127.0.0.1:{co_port}", f"127.0.0.1:{ps_port}"], "co")
    ps = spawn([f"{BIN}/parameter_server", f"127.0.0.1:{ps_port}", str(workers), "0",
                f"--mode={mode}", f"--staleness={staleness}", *ps_flags], "ps")
    procs = []
    try:
        assert wait_port(ps_port) and wait_port(co_port), "servers did not come up"
        t0 = time.time()
        for w in range(workers):
            argv = [f"{BIN}/worker_main", f"127.0.0.1:{co_port}", str(w), str(iters),
                    f"--mode={mode}", "--quiet", "--sync-timeout-ms=60000", "--seed=7"]
            if scenario == "persistent":
                ms = compute_ms * (1.10 if w == workers - 1 else 1.0)
                argv += [f"--compute-ms={ms}"]
            else:
                argv += [f"--compute-ms={compute_ms}", "--straggle-prob=0.1", "--straggle-factor=5"]
            procs.append(spawn(argv, f"w{w}"))
        for p in procs:
            p.wait(timeout=600)
        wall = time.time() - t0
    finally:
        for p in procs + [ps, co]:
            if p.poll() is None:
                p.terminate()
        for p in procs + [ps, co]:
            try:
                p.wait(timeout=5)
            except subprocess.TimeoutExpired:
                p.kill()
        for f in logs:
            f.close()
    per = []
    for w in range(workers):
        text = open(os.path.join(tmp, f"{mode}_{scenario}_w{w}.log")).read()
        m = re.search(r"summary iterations=(\d+) done=(\d+) elapsed_ms=([0-9.e+]+)", text)
        assert m, text[-400:]
        assert int(m.group(2)) == iters, f"worker {w}: {m.group(2)} of {iters} iterations completed"
        per.append(float(m.group(3)))
    return {"scenario": scenario, "mode": mode, "workers": workers, "iters": iters,
            "compute_ms": compute_ms, "staleness": staleness if mode == "ssp" else None,
            "job_ms": max(per), "worker_ms": [round(x, 1) for x in per],
            "sum_rate_it_s": round(sum(iters / (x / 1e3) for x in per), 1),
            "wall_s_incl_startup": round(wall, 2)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workers", type=int, default=4)
    ap.add_argument("--iters", type=int, default=400)
    ap.add_argument("--compute-ms", type=float, default=5.0)
    ap.add_argument("--staleness", type=int, default=4)
    ap.add_argument("--scenarios", default="persistent,transient")
    ap.add_argument("--modes", default="sync,ssp,async")
    ap.add_argument("--ps-flags", default="", help="extra parameter_server flags, comma separated")
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    if not os.path.exists(os.path.join(BIN, "worker_main")):
        sys.exit("build the binaries first: make -j8 binaries")
    rows = []
    with tempfile.TemporaryDirectory() as tmp:
        for scenario in a.scenarios.split(","):
            base = None
            for mode in a.modes.split(","):
                r = run(mode, a.workers, a.iters, a.compute_ms, scenario, a.staleness, tmp,
                        tuple(f for f in a.ps_flags.split(",") if f))
                if base is None:
                    base = r
                r["job_speedup_vs_sync"] = round(base["job_ms"] / r["job_ms"], 3)
                r["rate_vs_sync"] = round(r["sum_rate_it_s"] / base["sum_rate_it_s"], 3)
                rows.append(r)
                print(json.dumps(r), flush=True)
    if a.out:
        with open(a.out, "w") as f:
            for r in rows:
                f.write(json.dumps(r) + "\n")


if __name__ == "__main__":
    main()
