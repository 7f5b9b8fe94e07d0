"""T3 / T6 / T7 on the CPU binaries: the test_local.sh flow with assertions (BASELINE.json
config 1), server-side optimizers and async / SSP modes through the unchanged wire schema,
periodic checkpoints, resume from a checkpoint, and a missing worker."""
import os
import random
import re
import subprocess
import time

import grpc
import numpy as np
import pytest
import torch

from parameter_server_distributed_b200.control import protos
from tests._native import BIN, ROOT, Proc, ensure_binaries, free_port, wait_port

ensure_binaries()
M = protos.msg
T = lambda name, data: M("parameter_server.Tensor")(name=name, shape=[len(data)], data=data)


def _stub(port):
    return protos.Stub(grpc.insecure_channel(f"127.0.0.1:{port}"), "parameter_server.ParameterServer")


def test_selftest_binary():
    r = subprocess.run([f"{BIN}/psb_selftest"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and "psb_selftest ok" in r.stdout, r.stderr


def test_test_local_script(tmp_path):
    env = dict(os.environ, COORD_PORT=str(free_port()), PS_PORT=str(free_port()),
               LOG_DIR=str(tmp_path))
    r = subprocess.run(["bash", os.path.join(ROOT, "scripts", "test_local.sh")], env=env,
                       capture_output=True, text=True, timeout=180)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "PASS: test_local (10 x done=true)" in r.stdout
    assert "final weight = -0.0300, checkpoint = 458 bytes" in r.stdout


@pytest.mark.parametrize("kind", ["momentum", "adam", "adamw"])
def test_server_side_optimizers_match_torch(tmp_path, kind):
    port = free_port()
    ps = Proc([f"{BIN}/parameter_server", f"127.0.0.1:{port}", "2", "0", f"--optimizer={kind}",
               "--lr=0.05", "--weight-decay=0.01"], str(tmp_path), str(tmp_path / "ps.log"))
    try:
        assert wait_port(port)
        stub = _stub(port)
        rng = np.random.default_rng(0)
        w0 = rng.standard_normal(64).astype(np.float32)
        # bootstrap: the first aggregate becomes the model
        for wid in (0, 1):
            stub.ReceiveGradients(M("parameter_server.GradientUpdate")(
                worker_id=wid, iteration=0, gradients=[T("w", w0.tolist())]), timeout=5)
        p = torch.nn.Parameter(torch.tensor(w0))
        opt = {"momentum": lambda: torch.optim.SGD([p], lr=0.05, momentum=0.9, weight_decay=0.01),
               "adam": lambda: torch.optim.Adam([p], lr=0.05, weight_decay=0.01),
               "adamw": lambda: torch.optim.AdamW([p], lr=0.05, weight_decay=0.01)}[kind]()
        for it in range(1, 5):
            g0 = rng.standard_normal(64).astype(np.float32)
            g1 = rng.standard_normal(64).astype(np.float32)
            stub.ReceiveGradients(M("parameter_server.GradientUpdate")(
                worker_id=0, iteration=it, gradients=[T("w", g0.tolist())]), timeout=5)
            r = stub.ReceiveGradients(M("parameter_server.GradientUpdate")(
                worker_id=1, iteration=it, gradients=[T("w", g1.tolist())]), timeout=5)
            assert r.aggregation_complete
            p.grad = torch.tensor((g0 + g1) / np.float32(2))
            opt.step()
        got = np.array(stub.ServeParameters(M("parameter_server.PullRequest")(), timeout=5)
                       .parameters[0].data, dtype=np.float32)
        assert np.allclose(got, p.detach().numpy(), atol=2e-5, rtol=1e-4)
        # optimizer state survives a checkpoint through the sidecar file
        assert stub.SaveCheckpoint(M("parameter_server.SaveCheckpointRequest")(
            epoch=1, path="opt.ckpt"), timeout=5).success
        assert (tmp_path / "opt.ckpt.psb").exists()
        assert stub.LoadCheckpoint(M("parameter_server.LoadCheckpointRequest")(path="opt.ckpt"),
                                   timeout=5).success
    finally:
        ps.stop()


def test_async_mode_applies_each_push_and_ssp_bounds_staleness(tmp_path):
    port = free_port()
    ps = Proc([f"{BIN}/parameter_server", f"127.0.0.1:{port}", "2", "0", "--mode=ssp",
               "--staleness=1"], str(tmp_path), str(tmp_path / "ps.log"))
    try:
        assert wait_port(port)
        stub = _stub(port)
        push = lambda wid, it, v: stub.ReceiveGradients(M("parameter_server.GradientUpdate")(
            worker_id=wid, iteration=it, gradients=[T("w", [v] * 8)]), timeout=5)
        pull = lambda wid, it: stub.ServeParameters(M("parameter_server.PullRequest")(
            worker_id=wid, iteration=it), timeout=5)
        assert push(0, 0, 1.0).aggregation_complete       # bootstrap: the push becomes the model
        w = np.array(pull(0, 1).parameters[0].data)
        assert np.allclose(w, 1.0)                         # unscaled, as in sync mode (D8)
        assert push(0, 1, 0.2).aggregation_complete        # applied at once: no barrier
        assert np.allclose(np.array(pull(0, 1).parameters[0].data), 1.0 - 0.1)  # share = 1/2
        push(0, 1, 0.2)                                    # a re-sent clock is not applied twice
        assert np.allclose(np.array(pull(0, 1).parameters[0].data), 1.0 - 0.1)
        # worker 1 never pushed: it has completed 0 clocks, so clock 2 of worker 0 is 2 ahead
        assert pull(0, 1).ready            # 1 - 0 <= 1
        assert not pull(0, 2).ready        # 2 - 0 >  1
        push(1, 0, 0.0)
        assert pull(0, 2).ready            # slowest finished clock 0 -> 2 - 1 <= 1
        assert not pull(0, 3).ready
    finally:
        ps.stop()


def test_periodic_checkpoint_and_resume(tmp_path):
    ps_port, co_port = free_port(), free_port()
    env = {"PSB_CHECKPOINT_SCAN_MS": "100"}
    ps = Proc([f"{BIN}/parameter_server", f"127.0.0.1:{ps_port}", "1", "2"], str(tmp_path),
              str(tmp_path / "ps.log"), env=env)
    co = Proc([f"{BIN}/coordinator", f"127.0.0.1:{co_port}", f"127.0.0.1:{ps_port}"],
              str(tmp_path), str(tmp_path / "co.log"))
    try:
        assert wait_port(ps_port) and wait_port(co_port)
        w = subprocess.run([f"{BIN}/worker_main", f"127.0.0.1:{co_port}", "0", "5"],
                           capture_output=True, text=True, timeout=60, cwd=tmp_path)
        assert w.stdout.count("done=true") == 5
        time.sleep(0.5)
        out = ps.output()
        assert "periodic checkpointing every 2 iterations" in out
        assert "saved checkpoint: checkpoint_epoch_2.ckpt (epoch 2)" in out  # iteration 4 / 2
        ps.stop()
        # a fresh server resumes from the file when a worker passes it as argv[6]
        ps = Proc([f"{BIN}/parameter_server", f"127.0.0.1:{ps_port}", "1", "0"], str(tmp_path),
                  str(tmp_path / "ps2.log"))
        assert wait_port(ps_port)
        w = subprocess.run([f"{BIN}/worker_main", f"127.0.0.1:{co_port}", "0", "1", "", "0",
                            "checkpoint_epoch_2.ckpt"], capture_output=True, text=True, timeout=60,
                           cwd=tmp_path)
        assert "worker 0 loaded checkpoint from checkpoint_epoch_2.ckpt (epoch 2)" in w.stdout
        got = _stub(ps_port).ServeParameters(M("parameter_server.PullRequest")(), timeout=5)
        # -0.03 after 5 iterations, minus one more 0.01 step
        assert np.allclose(np.array(got.parameters[0].data), -0.04, atol=1e-6)
    finally:
        ps.stop()
        co.stop()


def test_missing_worker_times_out_with_done_false(tmp_path):
    ps_port, co_port = free_port(), free_port()
    ps = Proc([f"{BIN}/parameter_server", f"127.0.0.1:{ps_port}", "2", "0"], str(tmp_path),
              str(tmp_path / "ps.log"))
    co = Proc([f"{BIN}/coordinator", f"127.0.0.1:{co_port}", f"127.0.0.1:{ps_port}"],
              str(tmp_path), str(tmp_path / "co.log"))
    try:
        assert wait_port(ps_port) and wait_port(co_port)
        t0 = time.time()
        w = subprocess.run([f"{BIN}/worker_main", f"127.0.0.1:{co_port}", "0", "1",
                            "--sync-timeout-ms=400"], capture_output=True, text=True, timeout=60)
        assert "worker 0 iter 0 done=false" in w.stdout and w.returncode == 0
        assert time.time() - t0 < 10
        # no coordinator at all -> exit code 1 and the reference's stderr line
        w = subprocess.run([f"{BIN}/worker_main", f"127.0.0.1:{free_port()}", "5", "1"],
                           capture_output=True, text=True, timeout=60)
        assert w.returncode == 1 and "worker 5 failed to initialize" in w.stderr
    finally:
        ps.stop()
        co.stop()


def test_example_script_trains_and_checkpoints(tmp_path):
    """examples/train_mlp.py: the public API end to end on the host backend."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    ckpt = tmp_path / "mlp.ckpt"
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    r = subprocess.run([sys.executable, os.path.join(root, "examples", "train_mlp.py"), "--steps", "12",
                        "--checkpoint", str(ckpt)], capture_output=True, text=True, timeout=300,
                       env=env)
    assert r.returncode == 0, r.stdout[-400:] + r.stderr[-400:]
    assert "checkpoint" in r.stdout and ckpt.exists() and ckpt.stat().st_size > 1000
    # the same script in bounded-staleness mode (host emulation of the async engine)
    r = subprocess.run([sys.executable, os.path.join(root, "examples", "train_mlp.py"), "--steps", "8",
                        "--mode", "ssp"], capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0, r.stdout[-400:] + r.stderr[-400:]
    first, last = re.search(r"loss ([0-9.]+) -> ([0-9.]+)", r.stdout).groups()
    assert float(last) < float(first)


def test_python_worker_example_against_cpp_servers(tmp_path):
    """examples/python_worker.py: a grpcio worker drives the C++ coordinator and parameter
    server through the reference's wire schema."""
    import sys
    bin_dir, root = BIN, ROOT
    cport, pport = free_port(), free_port()
    ps = Proc([os.path.join(bin_dir, "parameter_server"), f"0.0.0.0:{pport}", "1", "0",
               "--optimizer=sgd", "--lr=0.1"], str(tmp_path), str(tmp_path / "ps.log"))
    co = Proc([os.path.join(bin_dir, "coordinator"), f"0.0.0.0:{cport}", f"127.0.0.1:{pport}"],
              str(tmp_path), str(tmp_path / "co.log"))
    try:
        assert wait_port(pport) and wait_port(cport)
        r = subprocess.run([sys.executable, os.path.join(root, "examples", "python_worker.py"),
                            "--coordinator", f"127.0.0.1:{cport}", "--iterations", "6"],
                           capture_output=True, text=True, timeout=120)
        assert r.returncode == 0, r.stdout[-400:] + r.stderr[-400:]
        assert r.stdout.count("done=true") == 6
    finally:
        ps.stop()
        co.stop()


def test_eviction_script(tmp_path):
    """scripts/test_eviction.sh: kill 1 of 3 workers, the others finish (shell-level T7)."""
    r = subprocess.run(["bash", os.path.join(ROOT, "scripts", "test_eviction.sh")],
                       capture_output=True, text=True, timeout=300,
                       env=dict(os.environ, ITERS="600", COORD_PORT=str(free_port()),
                                PS_PORT=str(free_port())))
    assert r.returncode == 0 and "PASS:" in r.stdout, r.stdout[-800:] + r.stderr[-400:]


def test_straggler_benchmark_modes_agree_on_the_work_done(tmp_path):
    """benchmarks/straggler_host.py: every worker completes all of its iterations in sync, SSP
    and async mode with simulated compute time and transient stragglers (timing is reported, not
    asserted — this box's scheduler is part of it)."""
    import json
    import subprocess
    import sys
    out = tmp_path / "s.jsonl"
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "benchmarks", "straggler_host.py"),
                        "--workers", "3", "--iters", "40", "--compute-ms", "2", "--scenarios",
                        "transient", "--out", str(out)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-500:] + r.stderr[-500:]
    rows = [json.loads(l) for l in open(out)]
    assert [x["mode"] for x in rows] == ["sync", "ssp", "async"]
    for x in rows:
        assert len(x["worker_ms"]) == 3 and x["job_ms"] >= 40 * 2


def test_multi_megabyte_model_through_cpp_worker_and_server(tmp_path):
    """6 MB of parameters resumed from a checkpoint, two C++ workers x 3 iterations: every pull is
    a response sent from the server's shared per-version encoding in several DATA frames, every
    push is received straight into the stream body and moved into the staging map.  The weights
    must come out as init - 3 x 0.01 exactly as with the 100-float model."""
    from parameter_server_distributed_b200.utils.checkpoint import write_reference_file
    rng = np.random.default_rng(5)
    init = [("big.weight", (1024, 1024), rng.standard_normal(1024 * 1024).astype(np.float32)),
            ("big.bias", (1024,), rng.standard_normal(1024).astype(np.float32)),
            ("tail.weight", (512, 1024), rng.standard_normal(512 * 1024).astype(np.float32))]
    write_reference_file(str(tmp_path / "init.ckpt"), 0, 0, init)
    ps_port, co_port = free_port(), free_port()
    co = Proc([f"{BIN}/coordinator", f"127.0.0.1:{co_port}", f"127.0.0.1:{ps_port}"],
              str(tmp_path), str(tmp_path / "co.log"))
    ps = Proc([f"{BIN}/parameter_server", f"127.0.0.1:{ps_port}", "2", "0", "--resume=init.ckpt"],
              str(tmp_path), str(tmp_path / "ps.log"))
    workers = []
    try:
        assert wait_port(ps_port) and wait_port(co_port)
        for wid in range(2):
            workers.append(Proc([f"{BIN}/worker_main", f"127.0.0.1:{co_port}", str(wid), "3",
                                 "--trace"], str(tmp_path), str(tmp_path / f"w{wid}.log")))
        for w in workers:
            assert w.p.wait(timeout=120) == 0
        for w in workers:
            out = w.output()
            assert out.count("done=true") >= 3 and "pull_ms=" in out, out[-300:]
        stub = protos.Stub(grpc.insecure_channel(
            f"127.0.0.1:{ps_port}", options=[("grpc.max_receive_message_length", -1)]),
            "parameter_server.ParameterServer")
        r = stub.ServeParameters(protos.msg("parameter_server.PullRequest")(worker_id=0, iteration=9),
                                 timeout=30)
        assert [t.name for t in r.parameters] == [n for n, _, _ in init]
        for t, (_, shape, a) in zip(r.parameters, init):
            assert tuple(t.shape) == shape
            exp = a
            for _ in range(3):
                exp = exp - np.float32(0.01)
            np.testing.assert_array_equal(np.asarray(t.data, dtype=np.float32), exp)
    finally:
        for w in workers:
            w.stop()
        ps.stop()
        co.stop()


def test_cpp_workers_over_two_parameter_server_shards(tmp_path):
    """worker_main --shards=2: three tensors resumed from per-shard checkpoints (a, c on shard 0;
    b on shard 1), two C++ workers x 4 iterations; every tensor ends at init - 4 x 0.01 and the
    workers report done=true throughout (each iteration needs both shards' barriers)."""
    from parameter_server_distributed_b200.utils.checkpoint import write_reference_file
    rng = np.random.default_rng(11)
    tensors = [("a", (64, 32)), ("b", (300,)), ("c", (8, 8, 4))]
    init = {n: rng.standard_normal(int(np.prod(s))).astype(np.float32) for n, s in tensors}
    co_port = free_port()
    co = Proc([f"{BIN}/coordinator", f"127.0.0.1:{co_port}", "127.0.0.1:1"], str(tmp_path),
              str(tmp_path / "co.log"))
    procs, ports = [], []
    try:
        assert wait_port(co_port)
        for sid in range(2):
            d = tmp_path / f"s{sid}"
            d.mkdir()
            mine = [(n, s, init[n]) for k, (n, s) in enumerate(tensors) if k % 2 == sid]
            write_reference_file(str(d / "init.ckpt"), 0, 0, mine)
            port = free_port()
            ports.append(port)
            procs.append(Proc([f"{BIN}/parameter_server", f"127.0.0.1:{port}", "2", "0",
                               "--resume=init.ckpt", f"--coordinator=127.0.0.1:{co_port}",
                               f"--shard-id={sid}", "--membership-poll-ms=100"], str(d),
                              str(d / "ps.log")))
        for p in ports:
            assert wait_port(p)
        workers = [Proc([f"{BIN}/worker_main", f"127.0.0.1:{co_port}", str(w), "4", "--shards=2",
                         "--heartbeat-ms=200"], str(tmp_path), str(tmp_path / f"w{w}.log"))
                   for w in range(2)]
        procs += workers
        for w in workers:
            assert w.p.wait(timeout=60) == 0
            assert w.output().count("done=true") == 4, w.output()[-300:]
        got = {}
        for port in ports:
            stub = protos.Stub(grpc.insecure_channel(f"127.0.0.1:{port}"),
                               "parameter_server.ParameterServer")
            r = stub.ServeParameters(protos.msg("parameter_server.PullRequest")(worker_id=0, iteration=9),
                                     timeout=10)
            for t in r.parameters:
                got[t.name] = (tuple(t.shape), np.asarray(t.data, dtype=np.float32))
        assert sorted(got) == ["a", "b", "c"]
        for n, s in tensors:
            exp = init[n]
            for _ in range(4):
                exp = exp - np.float32(0.01)
            assert got[n][0] == s
            np.testing.assert_array_equal(got[n][1], exp)
    finally:
        for p in procs:
            p.stop()
        co.stop()


def _wait_for(pred, timeout):
    end = time.time() + timeout
    while time.time() < end:
        if pred():
            return True
        time.sleep(0.1)
    return pred()


@pytest.mark.parametrize("shards", [1, 2])
def test_deploy_box_and_scale_workers_scripts(tmp_path, shards):
    """S5 / S6 on one box: deploy_box.sh (host mode; PS_SHARDS=2 splits the model over two
    parameter servers), then scale_workers.sh up 3 — checkpoint, server restart with the new
    worker count, the first new worker reloads the checkpoint — and stop."""
    run = tmp_path / "run"
    co_port = free_port()
    ps_port = free_port()
    env = dict(os.environ, RUN_DIR=str(run), COORDINATOR_PORT=str(co_port), PS_PORT=str(ps_port),
               WORKER_COUNT="2", ITERATIONS="30", CHECKPOINT_INTERVAL="0", PS_SHARDS=str(shards),
               WORKER_FLAGS="--sync-timeout-ms=60000")  # a slow box must not turn into done=false
    r = subprocess.run(["bash", os.path.join(ROOT, "scripts", "deploy_box.sh")], env=env,
                       capture_output=True, text=True, timeout=300)
    try:
        assert r.returncode == 0, r.stdout[-500:] + r.stderr[-500:]

        def done(wid, n):
            p = run / f"worker_{wid}.log"
            return p.exists() and p.read_text().count("done=true") >= n

        assert _wait_for(lambda: done(0, 30) and done(1, 30), 120), (run / "worker_0.log").read_text()[-300:]
        if shards == 2:
            assert "registered as shard 1" in (run / "parameter_server_1.log").read_text()
            return
        env["ITERATIONS"] = "10"
        r = subprocess.run(["bash", os.path.join(ROOT, "scripts", "scale_workers.sh"), "up", "3"],
                           env=env, capture_output=True, text=True, timeout=120)
        assert r.returncode == 0 and "scaled to 3 workers" in r.stdout, r.stdout + r.stderr
        assert (run / "scale.ckpt").exists()
        assert _wait_for(lambda: all(done(w, 10) for w in range(3)), 120), "\n".join(
            (run / f"worker_{w}.log").read_text()[-200:] for w in range(3))
        assert "loaded checkpoint from scale.ckpt" in (run / "worker_0.log").read_text()
        st = subprocess.run(["bash", os.path.join(ROOT, "scripts", "scale_workers.sh"), "status"],
                            env=env, capture_output=True, text=True, timeout=30)
        assert "parameter_server: running" in st.stdout and "coordinator: running" in st.stdout
    finally:
        subprocess.run(["bash", os.path.join(ROOT, "scripts", "scale_workers.sh"), "stop"], env=env,
                       capture_output=True, text=True, timeout=60)


def test_binaries_explain_themselves():
    """--help prints the positional contract of the reference binaries plus the extension flags
    (the reference has no --help: SURVEY §5.6) and exits 0 without starting anything."""
    for exe, words in (("coordinator", ["usage: coordinator", "PSB_STALE_TIMEOUT_S"]),
                       ("parameter_server", ["total_workers=2", "--resume=PATH|latest", "--shard-id"]),
                       ("worker_main", ["iterations=1", "--shards=N", "--trace"])):
        r = subprocess.run([f"{BIN}/{exe}", "--help"], capture_output=True, text=True, timeout=10)
        assert r.returncode == 0
        for w in words:
            assert w in r.stdout, (exe, w)
    r = subprocess.run([f"{BIN}/parameter_server", "127.0.0.1:1", "2", "0", "--no-such-flag=1"],
                       capture_output=True, text=True, timeout=10)
    assert r.returncode == 2 and "unknown flag" in r.stderr


def test_parameter_server_stats_line(tmp_path):
    """--stats-every-s: one JSON line of counters per period (the reference prints nothing after
    its start-up lines)."""
    import json
    ps_port, co_port = free_port(), free_port()
    ps = Proc([f"{BIN}/parameter_server", f"127.0.0.1:{ps_port}", "1", "0", "--stats-every-s=1"],
              str(tmp_path), str(tmp_path / "ps.log"))
    co = Proc([f"{BIN}/coordinator", f"127.0.0.1:{co_port}", f"127.0.0.1:{ps_port}"],
              str(tmp_path), str(tmp_path / "co.log"))
    try:
        assert wait_port(ps_port) and wait_port(co_port)
        w = subprocess.run([f"{BIN}/worker_main", f"127.0.0.1:{co_port}", "0", "50", "--quiet"],
                           capture_output=True, text=True, timeout=60, cwd=tmp_path)
        assert "done=50" in w.stdout
        def last():
            lines = [l for l in ps.output().splitlines() if l.startswith("stats ")]
            return json.loads(lines[-1][len("stats "):]) if lines else None

        assert _wait_for(lambda: (last() or {}).get("pushes") == 50, 8), ps.output()[-300:]
        st = last()
        assert st["pushes"] == 50 and st["pulls"] == 50 and st["iteration"] == 49
        assert st["bytes_in"] > 50 * 400 and st["total_workers"] == 1
    finally:
        ps.stop()
        co.stop()


def test_async_pushes_of_a_multi_megabyte_model_are_applied_exactly_once(tmp_path):
    """--mode=async, four C++ workers x 15 iterations on a 4 MB model, all at once: every push is
    applied exactly once with share 1/4 (final = init - 60 x 0.0025) while pulls of the shared
    per-version encoding race with the updates."""
    from parameter_server_distributed_b200.utils.checkpoint import write_reference_file
    rng = np.random.default_rng(6)
    init = rng.standard_normal(1 << 20).astype(np.float32)
    write_reference_file(str(tmp_path / "init.ckpt"), 0, 0, [("w", (1024, 1024), init)])
    ps_port, co_port = free_port(), free_port()
    co = Proc([f"{BIN}/coordinator", f"127.0.0.1:{co_port}", f"127.0.0.1:{ps_port}"],
              str(tmp_path), str(tmp_path / "co.log"))
    ps = Proc([f"{BIN}/parameter_server", f"127.0.0.1:{ps_port}", "4", "0", "--mode=async",
               "--resume=init.ckpt"], str(tmp_path), str(tmp_path / "ps.log"))
    workers = []
    try:
        assert wait_port(ps_port) and wait_port(co_port)
        workers = [Proc([f"{BIN}/worker_main", f"127.0.0.1:{co_port}", str(w), "15", "--mode=async",
                         "--quiet"], str(tmp_path), str(tmp_path / f"w{w}.log")) for w in range(4)]
        for w in workers:
            assert w.p.wait(timeout=120) == 0
            assert "done=15" in w.output(), w.output()[-200:]
        stub = protos.Stub(grpc.insecure_channel(
            f"127.0.0.1:{ps_port}", options=[("grpc.max_receive_message_length", -1)]),
            "parameter_server.ParameterServer")
        r = stub.ServeParameters(protos.msg("parameter_server.PullRequest")(worker_id=0, iteration=99),
                                 timeout=30)
        got = np.asarray(r.parameters[0].data, dtype=np.float32)
        np.testing.assert_allclose(got, init - np.float32(60 * 0.0025), atol=2e-5)
    finally:
        for w in workers:
            w.stop()
        ps.stop()
        co.stop()
