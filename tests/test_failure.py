"""T7: a worker dies -> the coordinator evicts it -> the parameter server's sync
barrier shrinks and the live workers carry on; plus the native coordinator client that gives the
GPU ranks the same membership view (barriers, blob exchange, peer-loss watch).

The reference stalls for ever in this situation (D9: /root/reference/src/coordinator.cpp:52-67
evicts, src/parameter_main.cpp:13-15 keeps total_workers fixed, nobody tells anybody)."""
import ctypes as C
import multiprocessing as mp
import os
import re
import signal
import struct
import subprocess
import time

import pytest

from parameter_server_distributed_b200.control import client as cc
from tests._native import BIN, Proc, ensure_binaries, free_port, wait_port

ensure_binaries()

FAST = {"PSB_STALE_SCAN_MS": "100", "PSB_STALE_TIMEOUT_S": "1"}


def _done_lines(text):
    return [(int(m.group(1)), m.group(2) == "true")
            for m in re.finditer(r"iter (\d+) done=(true|false)", text)]


def test_dead_worker_no_longer_stalls_the_parameter_server(tmp_path):
    """Kill 1 of 3 C++ workers mid-run: the other two reach done=true on every iteration."""
    ps_port, co_port = free_port(), free_port()
    co = Proc([f"{BIN}/coordinator", f"127.0.0.1:{co_port}", f"127.0.0.1:{ps_port}"],
              str(tmp_path), str(tmp_path / "co.log"), env=FAST)
    ps = Proc([f"{BIN}/parameter_server", f"127.0.0.1:{ps_port}", "3", "0",
               f"--coordinator=127.0.0.1:{co_port}", "--membership-poll-ms=100"],
              str(tmp_path), str(tmp_path / "ps.log"))
    workers = []
    try:
        assert wait_port(ps_port) and wait_port(co_port)
        iters = 1500
        for wid in range(3):
            n = iters if wid < 2 else 10_000_000  # worker 2 never finishes on its own
            workers.append(Proc([f"{BIN}/worker_main", f"127.0.0.1:{co_port}", str(wid), str(n),
                                 "--heartbeat-ms=100", "--sync-timeout-ms=8000"], str(tmp_path),
                                str(tmp_path / f"w{wid}.log")))
        # let the job get going, then kill worker 2 without ceremony
        end = time.time() + 20
        while time.time() < end and len(_done_lines(workers[0].output())) < 20:
            time.sleep(0.02)
        assert len(_done_lines(workers[0].output())) >= 20, workers[0].output()[-300:]
        workers[2].p.send_signal(signal.SIGKILL)
        killed_at = len(_done_lines(workers[0].output()))
        for w in workers[:2]:
            w.p.wait(timeout=120)
        for wid in (0, 1):
            lines = _done_lines(workers[wid].output())
            assert len(lines) == iters, (wid, len(lines))
            assert all(ok for _, ok in lines), [i for i, ok in lines if not ok][:5]
        assert killed_at < iters, "worker 2 was killed after the others had finished"
        out = ps.output()
        assert "following worker membership of coordinator" in out
        assert re.search(r"membership: total_workers 3 -> 2 \(2 live\)", out), out[-400:]
        assert "evicted stale worker 2" in co.output()
    finally:
        for w in workers:
            w.stop()
        ps.stop()
        co.stop()


def test_worker_that_returns_grows_the_barrier_back(tmp_path):
    """ParameterServerCore::set_total_workers through the service: 2 configured, one evicted,
    then it registers again — the count follows (never above the configured one)."""
    ps_port, co_port = free_port(), free_port()
    co = Proc([f"{BIN}/coordinator", f"127.0.0.1:{co_port}", f"127.0.0.1:{ps_port}"],
              str(tmp_path), str(tmp_path / "co.log"), env=FAST)
    ps = Proc([f"{BIN}/parameter_server", f"127.0.0.1:{ps_port}", "2", "0",
               f"--coordinator=127.0.0.1:{co_port}", "--membership-poll-ms=50"],
              str(tmp_path), str(tmp_path / "ps.log"))
    try:
        assert wait_port(ps_port) and wait_port(co_port)
        a = cc.CoordinatorClient(f"127.0.0.1:{co_port}", 0, 2, heartbeat_ms=100)
        b = cc.CoordinatorClient(f"127.0.0.1:{co_port}", 1, 2, heartbeat_ms=100)
        assert a.ps_address == f"127.0.0.1:{ps_port}"
        time.sleep(0.4)
        b.close()  # heartbeats of worker 1 stop -> eviction after ~1 s
        end = time.time() + 10
        while time.time() < end and "total_workers 2 -> 1" not in ps.output():
            time.sleep(0.05)
        assert "membership: total_workers 2 -> 1 (1 live)" in ps.output()
        b = cc.CoordinatorClient(f"127.0.0.1:{co_port}", 1, 2, heartbeat_ms=100)
        end = time.time() + 10
        while time.time() < end and "total_workers 1 -> 2" not in ps.output():
            time.sleep(0.05)
        assert "membership: total_workers 1 -> 2 (2 live)" in ps.output()
        a.close()
        b.close()
    finally:
        ps.stop()
        co.stop()


# ------------------------------------------------------------------ native coordinator client
def _rank(rank, world, addr, q, die):
    try:
        c = cc.CoordinatorClient(addr, rank, world, heartbeat_ms=100)
        c.register_shard(rank, f"hbm://test/gpu{rank}", rank)
        c.barrier("up")
        got = c.allgather("handles", bytes([rank]) * 64)
        objs = c.allgather_object("objs", {"rank": rank, "pad": "x" * (rank * 100)})
        mx = c.reduce_max("mx", float(rank))
        for _ in range(3):  # the same key again: rounds must not mix
            again = c.allgather("handles", bytes([rank + 10]) * 64)
        shards, total = c.shard_map()
        word = C.c_uint32(0)
        c.barrier("registered")
        c.watch(world, C.addressof(word))
        if rank == die:
            os._exit(0)  # no close(): heartbeats simply stop
        t0 = time.time()
        while not c.peer_lost and time.time() - t0 < 15:
            time.sleep(0.02)
        q.put((rank, [g[0] for g in got], [o["rank"] for o in objs], mx, [a[0] for a in again],
               shards, total, c.peer_lost, word.value, c.list_workers()))
        c.close()
    except BaseException as e:  # noqa: BLE001
        q.put((rank, "error", repr(e)))


def test_ranks_bootstrap_and_watch_each_other_through_the_coordinator(tmp_path):
    """What psd.init() does on a GPU box, minus the GPUs: register, shard map, handle exchange,
    barrier, small all-gathers — then one rank dies and the others are told (abort word set)."""
    proc, addr = cc.spawn_coordinator(env=FAST, log_path=str(tmp_path / "co.log"))
    try:
        world, die = 3, 2
        ctx = mp.get_context("spawn")
        q = ctx.Queue()
        ps = [ctx.Process(target=_rank, args=(r, world, addr, q, die)) for r in range(world)]
        for p in ps:
            p.start()
        res = {}
        for _ in range(world - 1):
            r = q.get(timeout=60)
            assert r[1] != "error", r
            res[r[0]] = r
        for p in ps:
            p.join(timeout=10)
        for rank in (0, 1):
            _, got, objs, mx, again, shards, total, lost, word, alive = res[rank]
            assert got == [0, 1, 2] and objs == [0, 1, 2] and mx == 2.0
            assert again == [10, 11, 12]
            assert [(s[0], s[2]) for s in shards] == [(0, 0), (1, 1), (2, 2)] and total == 3
            assert shards[1][1] == "hbm://test/gpu1"
            assert lost and word == 1, "the survivor was not told about the lost peer"
            assert alive == [0, 1]
    finally:
        cc._stop(proc)


def test_ensure_coordinator_rendezvous(tmp_path, monkeypatch):
    """The torchrun case: nobody started a coordinator; rank 0 launches the binary and the
    address reaches the other ranks through the rendezvous file."""
    monkeypatch.setenv("PSB_RENDEZVOUS_DIR", str(tmp_path))
    monkeypatch.setenv("MASTER_PORT", "31999")
    monkeypatch.delenv("PSB_COORDINATOR", raising=False)
    c0 = cc.ensure_coordinator(0, 2, heartbeat_ms=200)
    c1 = cc.ensure_coordinator(1, 2, heartbeat_ms=200)
    assert c0.address == c1.address
    assert c0._proc is not None and c1._proc is None
    assert c0.list_workers() == [0, 1]
    c1.close()
    proc = c0._proc
    c0.close()
    assert not os.listdir(tmp_path), os.listdir(tmp_path)
    assert proc.poll() is not None, "the coordinator launched by rank 0 outlived it"


def test_parameter_server_restart_resumes_from_its_latest_checkpoint(tmp_path):
    """The other half of T7: the parameter server dies (SIGKILL) mid-run and a supervisor starts
    it again with --resume=latest.  The model continues from the newest periodic checkpoint and
    the workers — which only saw some RPCs fail — carry on against the new process.  The
    reference's restarted server is empty until somebody passes CHECKPOINT_PATH to a worker
    (/root/reference/scripts/scale_workers.sh:137-144, scripts/start_worker.sh:22-23)."""
    from parameter_server_distributed_b200.utils.checkpoint import read_reference_file
    ps_port, co_port = free_port(), free_port()
    env = {"PSB_CHECKPOINT_SCAN_MS": "50"}
    argv = [f"{BIN}/parameter_server", f"127.0.0.1:{ps_port}", "2", "20", "--resume=latest"]
    co = Proc([f"{BIN}/coordinator", f"127.0.0.1:{co_port}", f"127.0.0.1:{ps_port}"],
              str(tmp_path), str(tmp_path / "co.log"))
    ps = Proc(argv, str(tmp_path), str(tmp_path / "ps1.log"), env=env)
    ps2 = None
    workers = []
    try:
        assert wait_port(ps_port) and wait_port(co_port)
        assert "resume: no checkpoint_epoch_*.ckpt here, starting empty" in ps.output()
        iters = 1200
        for wid in range(2):
            workers.append(Proc([f"{BIN}/worker_main", f"127.0.0.1:{co_port}", str(wid), str(iters),
                                 "--heartbeat-ms=200", "--sync-timeout-ms=3000", "--compute-ms=1"],
                                str(tmp_path), str(tmp_path / f"w{wid}.log")))
        def saved_epochs():  # the checkpoint thread may skip epochs when iterations fly by
            return [int(e) for e in re.findall(r"saved checkpoint: checkpoint_epoch_(\d+)\.ckpt",
                                               ps.output())]

        end = time.time() + 30
        while time.time() < end and not any(e >= 5 for e in saved_epochs()):
            time.sleep(0.02)
        assert any(e >= 5 for e in saved_epochs()), ps.output()[-300:]
        assert max(saved_epochs()) < 40, "the run was nearly over before the kill"
        ps.p.send_signal(signal.SIGKILL)
        ps.p.wait(timeout=5)
        ps.f.close()
        time.sleep(0.3)  # the workers meet a dead port for a moment
        ps2 = Proc(argv, str(tmp_path), str(tmp_path / "ps2.log"), env=env)
        assert wait_port(ps_port)
        m = re.search(r"resumed from (checkpoint_epoch_(\d+)\.ckpt) \(epoch (\d+), iteration (\d+)\)",
                      ps2.output())
        end = time.time() + 5
        while m is None and time.time() < end:
            time.sleep(0.05)
            m = re.search(r"resumed from (checkpoint_epoch_(\d+)\.ckpt) \(epoch (\d+), iteration (\d+)\)",
                          ps2.output())
        assert m, ps2.output()[-300:]
        assert int(m.group(2)) >= 5 and int(m.group(4)) >= 100
        for w in workers:
            w.p.wait(timeout=180)
        for wid in range(2):
            lines = _done_lines(workers[wid].output())
            assert len(lines) == iters
            lost = [i for i, ok in lines if not ok]
            assert len(lost) <= 20, f"worker {wid} lost {len(lost)} iterations: {lost[:8]}"
            assert all(ok for i, ok in lines if i > (lost[-1] if lost else 0))
        # the model went on from the checkpoint: weight = 0.01 - 0.01 * (#aggregations - 1), so
        # a server that restarted empty would end near -0.01 * (iterations after the restart)
        def newest_checkpoint():
            files = sorted(tmp_path.glob("checkpoint_epoch_*.ckpt"),
                           key=lambda p: int(p.stem.split("_")[-1]))
            return read_reference_file(str(files[-1]))

        epoch, iteration, tensors = newest_checkpoint()
        end = time.time() + 10  # the checkpoint thread scans every 50 ms; give it a moment
        while iteration < iters - 25 and time.time() < end:
            time.sleep(0.1)
            epoch, iteration, tensors = newest_checkpoint()
        w = float(tensors[0][2].reshape(-1)[0])
        assert iteration >= iters - 25
        assert w < -0.01 * (iters - 60), (w, iteration)
    finally:
        for w in workers:
            w.stop()
        ps.stop() if ps.p.poll() is None else None
        if ps2:
            ps2.stop()
        co.stop()


def test_coordinator_restart_rebuilds_its_soft_state(tmp_path):
    """The coordinator keeps nothing on disk: after a crash and restart the workers' next
    heartbeat is refused (unknown id), they register again, and a sharded parameter server
    announces itself again — registry and shard map are back without touching the job."""
    ps_port, co_port = free_port(), free_port()
    coargs = [f"{BIN}/coordinator", f"127.0.0.1:{co_port}", f"127.0.0.1:{ps_port}"]
    co = Proc(coargs, str(tmp_path), str(tmp_path / "co1.log"), env=FAST)
    ps = Proc([f"{BIN}/parameter_server", f"127.0.0.1:{ps_port}", "2", "0",
               f"--coordinator=127.0.0.1:{co_port}", "--shard-id=0", "--membership-poll-ms=100"],
              str(tmp_path), str(tmp_path / "ps.log"))
    co2 = None
    workers = []
    try:
        assert wait_port(ps_port) and wait_port(co_port)
        for wid in range(2):
            workers.append(Proc([f"{BIN}/worker_main", f"127.0.0.1:{co_port}", str(wid), "100000",
                                 "--heartbeat-ms=100", "--compute-ms=2", "--quiet"], str(tmp_path),
                                str(tmp_path / f"w{wid}.log")))
        probe = cc.CoordinatorClient(f"127.0.0.1:{co_port}", 99, 3, heartbeat_ms=100)
        end = time.time() + 10
        while time.time() < end and probe.list_workers() != [0, 1, 99]:
            time.sleep(0.05)
        assert probe.list_workers() == [0, 1, 99]
        assert [s[0] for s in probe.shard_map()[0]] == [0]
        co.p.send_signal(signal.SIGKILL)
        co.p.wait(timeout=5)
        time.sleep(0.3)
        co2 = Proc(coargs, str(tmp_path), str(tmp_path / "co2.log"), env=FAST)
        assert wait_port(co_port)
        fresh = cc.CoordinatorClient(f"127.0.0.1:{co_port}", 98, 4, heartbeat_ms=100)
        end = time.time() + 10
        while time.time() < end and not (set(fresh.list_workers()) >= {0, 1, 99}
                                         and fresh.shard_map()[0]):
            time.sleep(0.05)
        assert set(fresh.list_workers()) >= {0, 1, 99}, fresh.list_workers()
        shards, _ = fresh.shard_map()
        assert [(s[0], s[1]) for s in shards] == [(0, f"127.0.0.1:{ps_port}")]
        assert probe.evictions() >= 1  # it found itself unknown and registered again
        for w in workers:
            assert w.p.poll() is None, "a worker died with the coordinator"
        probe.close()
        fresh.close()
    finally:
        for w in workers:
            w.stop()
        ps.stop()
        co.stop() if co.p.poll() is None else None
        if co2:
            co2.stop()


def _init_rank(rank, world, port, rdv, q):
    try:
        os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world),
                          MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), PSB_RENDEZVOUS_DIR=rdv,
                          PSB_HEARTBEAT_MS="100")
        os.environ.pop("PSB_COORDINATOR", None)
        import parameter_server_distributed_b200 as psd
        ctx = psd.init(control="coordinator")  # no GPU here: control plane only, fabric is None
        assert ctx.coord is not None and ctx.world == world and ctx.fabric is None
        ctx.barrier("up")
        objs = ctx.all_gather_object({"rank": rank, "blob": "x" * (50 * rank)}, key="t")
        mx = [ctx.reduce_max(float(rank * (i + 1)), key="m") for i in range(3)]
        workers = ctx.coord.list_workers()
        ctx.barrier("down")
        ctx.close()
        q.put((rank, [o["rank"] for o in objs], mx, workers))
    except BaseException as e:  # noqa: BLE001
        q.put((rank, "error", repr(e)))


def test_psd_init_bootstraps_ranks_through_a_coordinator_it_launches(tmp_path):
    """The torchrun path of the GPU benchmark, minus the GPUs: nobody started a coordinator,
    rank 0's psd.init() launches the binary, the others find it through the rendezvous file, and
    the host collectives of Context (barrier, all_gather_object, reduce_max) run through it."""
    world, port = 4, free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_init_rank, args=(r, world, port, str(tmp_path), q))
          for r in range(world)]
    for p in ps:
        p.start()
    res = {}
    for _ in range(world):
        r = q.get(timeout=120)
        assert r[1] != "error", r
        res[r[0]] = r
    for p in ps:
        p.join(timeout=20)
        assert p.exitcode == 0
    for rank in range(world):
        _, ranks, mx, workers = res[rank]
        assert ranks == list(range(world))
        assert mx == [float((world - 1) * (i + 1)) for i in range(3)]
        assert workers == list(range(world))
    assert not os.listdir(tmp_path), os.listdir(tmp_path)  # the rendezvous file is gone


def _fd_rank(rank, world, addr, q):
    try:
        from parameter_server_distributed_b200.parallel import fabric as fb
        c = cc.CoordinatorClient(addr, rank, world, heartbeat_ms=100)
        ctl = fb._Control(lambda key, obj: c.allgather_object(key, obj), lambda name: c.barrier(name))
        mine = []
        for k in range(2):  # two descriptors per rank, like heap segment + signal pad
            fd = os.memfd_create(f"psb_test_{rank}_{k}")
            os.write(fd, f"rank {rank} segment {k}".encode())
            mine.append(fd)
        fds = fb._exchange_fds(ctl, rank, world, mine)
        seen = []
        for r in range(world):
            row = []
            for fd in fds[r]:
                row.append(os.pread(fd, 64, 0).decode())  # the offset is shared between ranks
            seen.append(row)
        c.barrier("done")
        c.close()
        q.put((rank, seen))
    except BaseException as e:  # noqa: BLE001
        q.put((rank, "error", repr(e)))


def test_file_descriptors_travel_between_ranks(tmp_path):
    """The VMM heap's bootstrap step without a GPU: every rank's POSIX descriptors (here memfds)
    reach every peer over Unix sockets whose paths were exchanged through the coordinator
    (parallel/fabric.py: _exchange_fds, the route of cuMemExportToShareableHandle fds)."""
    proc, addr = cc.spawn_coordinator(env=FAST, log_path=str(tmp_path / "co.log"))
    try:
        world = 4
        ctx = mp.get_context("spawn")
        q = ctx.Queue()
        ps = [ctx.Process(target=_fd_rank, args=(r, world, addr, q)) for r in range(world)]
        for p in ps:
            p.start()
        for _ in range(world):
            r = q.get(timeout=120)
            assert r[1] != "error", r
            assert r[1] == [[f"rank {o} segment {k}" for k in range(2)] for o in range(world)]
        for p in ps:
            p.join(timeout=20)
    finally:
        cc._stop(proc)


def test_restarted_worker_joins_at_the_jobs_iteration(tmp_path):
    """A worker dies, the barrier shrinks, a supervisor starts the worker again with --join: it
    asks the server where the job is, continues there, the barrier grows back, and all three
    finish.  (Started at iteration 0 — what the reference does, scripts/scale_workers.sh:113 —
    it would push iterations nobody waits for any more.)"""
    ps_port, co_port = free_port(), free_port()
    co = Proc([f"{BIN}/coordinator", f"127.0.0.1:{co_port}", f"127.0.0.1:{ps_port}"],
              str(tmp_path), str(tmp_path / "co.log"), env=FAST)
    ps = Proc([f"{BIN}/parameter_server", f"127.0.0.1:{ps_port}", "3", "0",
               f"--coordinator=127.0.0.1:{co_port}", "--membership-poll-ms=100"],
              str(tmp_path), str(tmp_path / "ps.log"))
    iters = 4000
    common = ["--heartbeat-ms=100", "--sync-timeout-ms=8000", "--compute-ms=1"]
    workers, again = [], None
    try:
        assert wait_port(ps_port) and wait_port(co_port)
        for wid in range(3):
            workers.append(Proc([f"{BIN}/worker_main", f"127.0.0.1:{co_port}", str(wid), str(iters),
                                 *common], str(tmp_path), str(tmp_path / f"w{wid}.log")))
        end = time.time() + 20
        while time.time() < end and len(_done_lines(workers[0].output())) < 50:
            time.sleep(0.02)
        workers[2].p.send_signal(signal.SIGKILL)
        end = time.time() + 10
        while time.time() < end and "total_workers 3 -> 2" not in ps.output():
            time.sleep(0.05)
        assert "membership: total_workers 3 -> 2" in ps.output()
        again = Proc([f"{BIN}/worker_main", f"127.0.0.1:{co_port}", "2", str(iters), "--join",
                      *common], str(tmp_path), str(tmp_path / "w2b.log"))
        for w in workers[:2] + [again]:
            assert w.p.wait(timeout=180) == 0
        m = re.search(r"worker 2 joining at iteration (\d+)", again.output())
        assert m and 50 <= int(m.group(1)) < iters, again.output()[:200]
        joined = int(m.group(1))
        lines = _done_lines(again.output())
        assert [i for i, _ in lines] == list(range(joined, iters))
        assert sum(ok for _, ok in lines) >= len(lines) - 2, [i for i, ok in lines if not ok][:5]
        for wid in (0, 1):
            lines = _done_lines(workers[wid].output())
            assert len(lines) == iters and all(ok for _, ok in lines)
        assert re.search(r"membership: total_workers 2 -> 3 \(3 live\)", ps.output())
        assert joined < iters - 100, "the job was over before the worker came back"
    finally:
        for w in workers + ([again] if again else []):
            w.stop()
        ps.stop()
        co.stop()


def test_eight_ranks_hammer_the_control_plane(tmp_path):
    """8 clients (threads, each with its own native handle and connections) run 150 rounds of
    barrier + all-gather + max-reduce + clock update against one coordinator: the control traffic
    of an 8-GPU job, compressed.  Every round must see every rank's value of THAT round."""
    import threading
    proc, addr = cc.spawn_coordinator(env=FAST, log_path=str(tmp_path / "co.log"))
    world, rounds = 8, 150
    errs, clients = [], []
    try:
        clients = [cc.CoordinatorClient(addr, r, world, heartbeat_ms=50) for r in range(world)]

        def work(r):
            c = clients[r]
            try:
                for i in range(rounds):
                    c.barrier("b")
                    got = c.allgather("v", struct.pack("<ii", r, i))
                    assert [struct.unpack("<ii", g) for g in got] == [(q, i) for q in range(world)]
                    assert c.reduce_max("m", float(r + i)) == float(world - 1 + i)
                    if i % 10 == 0:
                        assert c.update_clock(i) <= i
                        assert c.list_workers() == list(range(world))
            except BaseException as e:  # noqa: BLE001
                errs.append((r, repr(e)))

        ts = [threading.Thread(target=work, args=(r,)) for r in range(world)]
        t0 = time.time()
        for t in ts:
            t.start()
        for t in ts:
            t.join(timeout=120)
        assert not errs, errs[:3]
        assert all(not t.is_alive() for t in ts)
        assert time.time() - t0 < 60
    finally:
        for c in clients:
            c.close()
        cc._stop(proc)


def test_planned_stop_writes_a_last_checkpoint(tmp_path):
    """SIGTERM (systemctl stop, Ctrl-C): the parameter server saves where its periodic
    checkpoints go, exits 0, and --resume=latest continues from the very last update."""
    ps_port, co_port = free_port(), free_port()
    argv = [f"{BIN}/parameter_server", f"127.0.0.1:{ps_port}", "1", "1000", "--resume=latest"]
    co = Proc([f"{BIN}/coordinator", f"127.0.0.1:{co_port}", f"127.0.0.1:{ps_port}"],
              str(tmp_path), str(tmp_path / "co.log"))
    ps = Proc(argv, str(tmp_path), str(tmp_path / "ps1.log"))
    ps2 = None
    try:
        assert wait_port(ps_port) and wait_port(co_port)
        w = subprocess.run([f"{BIN}/worker_main", f"127.0.0.1:{co_port}", "0", "37", "--quiet"],
                           capture_output=True, text=True, timeout=60, cwd=tmp_path)
        assert "done=37" in w.stdout
        assert not list(tmp_path.glob("*.ckpt")), "no periodic checkpoint is due yet (interval 1000)"
        ps.p.send_signal(signal.SIGTERM)
        assert ps.p.wait(timeout=20) == 0
        out = ps.output()
        assert "signal 15: shutting down" in out
        assert "saved checkpoint: checkpoint_epoch_0.ckpt (epoch 0)" in out
        ps2 = Proc(argv, str(tmp_path), str(tmp_path / "ps2.log"))
        assert wait_port(ps_port)
        end = time.time() + 5
        while time.time() < end and "resumed from" not in ps2.output():
            time.sleep(0.05)
        assert "resumed from checkpoint_epoch_0.ckpt (epoch 0, iteration 36)" in ps2.output()
        w = subprocess.run([f"{BIN}/worker_main", f"127.0.0.1:{co_port}", "0", "40", "--join", "--quiet"],
                           capture_output=True, text=True, timeout=60, cwd=tmp_path)
        assert "joining at iteration 36" in w.stdout and "done=4" in w.stdout, w.stdout
    finally:
        ps.stop() if ps.p.poll() is None else ps.f.close()
        if ps2:
            ps2.stop()
        co.stop()


def test_c_example_of_the_control_abi(tmp_path):
    """examples/control_client.c: include/psb_control.h used from plain C (two processes)."""
    exe = str(tmp_path / "control_client")
    native = os.path.join(os.path.dirname(cc.LIB_PATH), "")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    subprocess.run(["gcc", "-std=c11", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(root, "include"),
                    os.path.join(root, "examples", "control_client.c"), "-L", native, "-lpsb_control",
                    f"-Wl,-rpath,{native}", "-o", exe], check=True)
    proc, addr = cc.spawn_coordinator(env=FAST, log_path=str(tmp_path / "co.log"))
    try:
        ps = [subprocess.Popen([exe, addr, str(r), "2"], stdout=subprocess.PIPE, text=True)
              for r in range(2)]
        outs = [p.communicate(timeout=60)[0] for p in ps]
        for r, (p, out) in enumerate(zip(ps, outs)):
            assert p.returncode == 0, out
            assert f"rank {r}: rank 0 says 'hello from 0'" in out
            assert f"rank {r}: rank 1 says 'hello from 1'" in out
            assert f"rank {r}: 2 worker(s) registered" in out
    finally:
        cc._stop(proc)
