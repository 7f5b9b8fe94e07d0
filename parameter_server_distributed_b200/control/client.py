"""Coordinator client of the Python workers (ctypes over ``libpsb_control.so``).

The ranks of a job bootstrap through the C++ coordinator exactly as the reference's workers do
(``Worker::initialize`` -> discovery -> ``RegisterWorker`` -> heartbeat thread,
/root/reference/src/worker.cpp:108-122, 141-186, 216-238) and use its extension service for
what the GPU data plane needs on the host side: shard registration, the exchange of
symmetric-heap handles, host barriers and small all-gathers.  ``torch.distributed`` is not
involved; the native library runs the heartbeats on its own thread (no GIL).

``ensure_coordinator`` covers the torchrun case, where nobody started a coordinator: rank 0
launches ``build/bin/coordinator`` and publishes its address in a rendezvous file on the box.
"""
from __future__ import annotations

import atexit
import ctypes as C
import json
import os
import pickle
import signal
import socket
import struct
import subprocess
import time
from pathlib import Path
from typing import List, Optional, Tuple

ROOT = Path(__file__).resolve().parents[2]
LIB_PATH = Path(__file__).resolve().parents[1] / "_native" / "libpsb_control.so"
COORDINATOR_BIN = Path(os.environ.get("PSB_COORDINATOR_BIN") or ROOT / "build" / "bin" / "coordinator")

_LIB = None


class ControlError(RuntimeError):
    pass


def available() -> bool:
    """Can this process use the coordinator path (library and binary are built)?"""
    return LIB_PATH.exists() and COORDINATOR_BIN.exists()


def load():
    global _LIB
    if _LIB is not None:
        return _LIB
    if not LIB_PATH.exists():
        from .. import _build
        _build.build_binaries()
    lib = C.CDLL(str(LIB_PATH))
    vp, i32, sz = C.c_void_p, C.c_int32, C.c_size_t
    sig = {
        "psbc_last_error": (C.c_char_p, []),
        "psbc_connect": (vp, [C.c_char_p, i32, C.c_char_p, i32, C.c_char_p, i32, i32]),
        "psbc_close": (None, [vp]),
        "psbc_ps_address": (C.c_char_p, [vp]),
        "psbc_set_status": (C.c_int, [vp, i32]),
        "psbc_list_workers": (C.c_int, [vp, C.POINTER(i32), i32, C.POINTER(i32)]),
        "psbc_live_workers": (C.c_int, [vp]),
        "psbc_evictions": (C.c_int, [vp]),
        "psbc_watch": (C.c_int, [vp, i32, vp]),
        "psbc_peer_lost": (C.c_int, [vp]),
        "psbc_register_shard": (C.c_int, [vp, i32, C.c_char_p, i32]),
        "psbc_shard_map": (C.c_int, [vp, C.POINTER(i32), C.POINTER(i32), i32, C.c_char_p, sz,
                                      C.POINTER(i32), C.POINTER(i32)]),
        "psbc_barrier": (C.c_int, [vp, C.c_char_p, i32, i32]),
        "psbc_allgather": (C.c_int, [vp, C.c_char_p, vp, sz, i32, vp, i32]),
        "psbc_put": (C.c_int, [vp, C.c_char_p, vp, sz]),
        "psbc_get": (C.c_int, [vp, C.c_char_p, i32, vp, sz, C.POINTER(sz)]),
        "psbc_update_clock": (C.c_int, [vp, i32, C.POINTER(i32)]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    _LIB = lib
    return lib


IDLE, TRAINING, CHECKPOINTING, ERROR = 0, 1, 2, 3


class CoordinatorClient:
    """One registered worker.  ``rank`` is the worker id; collectives run over ranks
    ``0..world-1``."""

    def __init__(self, address: str, rank: int, world: int, *, heartbeat_ms: int = 5000,
                 connect_timeout_s: float = 60.0, self_address: str = "", self_port: int = 0):
        self.lib = load()
        self.address, self.rank, self.world = address, rank, world
        self._h = self.lib.psbc_connect(address.encode(), rank, self_address.encode(), self_port,
                                        socket.gethostname().encode(), heartbeat_ms,
                                        int(connect_timeout_s * 1000))
        if not self._h:
            raise ControlError(self._err("psbc_connect"))
        self._proc: Optional[subprocess.Popen] = None  # coordinator this rank launched, if any
        self._rendezvous: Optional[str] = None

    # ------------------------------------------------------------------ helpers
    def _err(self, what: str) -> str:
        return f"{what}: {self.lib.psbc_last_error().decode(errors='replace')}"

    def _check(self, rc: int, what: str) -> None:
        if rc != 0:
            raise ControlError(f"{self._err(what)} (code {rc})")

    @property
    def ps_address(self) -> str:
        return self.lib.psbc_ps_address(self._h).decode()

    def set_status(self, status: int) -> None:
        self.lib.psbc_set_status(self._h, status)

    # ------------------------------------------------------------------ membership
    def list_workers(self) -> List[int]:
        ids = (C.c_int32 * 1024)()
        n = C.c_int32()
        self._check(self.lib.psbc_list_workers(self._h, ids, 1024, C.byref(n)), "ListWorkers")
        return sorted(ids[i] for i in range(min(n.value, 1024)))

    def live_workers(self) -> int:
        return int(self.lib.psbc_live_workers(self._h))

    def evictions(self) -> int:
        return int(self.lib.psbc_evictions(self._h))

    def watch(self, expected_world: int, abort_word_ptr: int = 0) -> None:
        """From now on a shrinking registry (a peer evicted) sets ``peer_lost`` and writes 1 to
        the 32-bit word at ``abort_word_ptr`` (host address; 0 = none).  ``expected_world=0``
        stops watching."""
        self._check(self.lib.psbc_watch(self._h, expected_world, abort_word_ptr or None), "watch")

    @property
    def peer_lost(self) -> bool:
        return bool(self.lib.psbc_peer_lost(self._h))

    # ------------------------------------------------------------------ shard map
    def register_shard(self, shard_id: int, address: str, device: int) -> None:
        self._check(self.lib.psbc_register_shard(self._h, shard_id, address.encode(), device),
                    "RegisterShard")

    def shard_map(self) -> Tuple[List[Tuple[int, str, int]], int]:
        """[(shard_id, address, device)], total_workers."""
        cap = 1024
        ids, devs = (C.c_int32 * cap)(), (C.c_int32 * cap)()
        buf = C.create_string_buffer(64 * cap)
        n, tw = C.c_int32(), C.c_int32()
        self._check(self.lib.psbc_shard_map(self._h, ids, devs, cap, buf, len(buf), C.byref(n),
                                            C.byref(tw)), "GetShardMap")
        addrs = buf.raw.split(b"\0")
        return [(ids[i], addrs[i].decode(), devs[i]) for i in range(min(n.value, cap))], tw.value

    # ------------------------------------------------------------------ host collectives
    def barrier(self, name: str = "host", timeout_s: float = 600.0, expected: Optional[int] = None):
        n = self.world if expected is None else expected
        self._check(self.lib.psbc_barrier(self._h, name.encode(), n, int(timeout_s * 1000)),
                    f"barrier({name})")

    def allgather(self, key: str, blob: bytes, timeout_s: float = 600.0) -> List[bytes]:
        """Fixed-size contribution per rank (all ranks pass the same length)."""
        if self.world == 1:
            return [bytes(blob)]
        n = len(blob)
        out = C.create_string_buffer(n * self.world)
        self._check(self.lib.psbc_allgather(self._h, key.encode(), blob, n, self.world, out,
                                            int(timeout_s * 1000)), f"allgather({key})")
        return [out.raw[r * n:(r + 1) * n] for r in range(self.world)]

    def allgather_object(self, key: str, obj) -> list:
        data = pickle.dumps(obj)
        sizes = [struct.unpack("<q", b)[0]
                 for b in self.allgather(key + "/n", struct.pack("<q", len(data)))]
        width = max(sizes)
        parts = self.allgather(key, data.ljust(width, b"\0"))
        return [pickle.loads(p[:s]) for p, s in zip(parts, sizes)]

    def reduce_max(self, key: str, x: float) -> float:
        return max(struct.unpack("<d", b)[0] for b in self.allgather(key, struct.pack("<d", x)))

    def update_clock(self, clock: int) -> int:
        lo = C.c_int32()
        self._check(self.lib.psbc_update_clock(self._h, clock, C.byref(lo)), "UpdateClock")
        return lo.value

    # ------------------------------------------------------------------ blobs
    def put(self, key: str, blob: bytes) -> None:
        buf = (C.c_char * max(1, len(blob))).from_buffer_copy(blob or b"\0")
        self._check(self.lib.psbc_put(self._h, key.encode(), buf, len(blob)), "PutHandle")

    def get(self, key: str, rank: int, cap: int = 1 << 16) -> Optional[bytes]:
        """Rank ``rank``'s blob of ``key`` (None when absent)."""
        out = (C.c_char * cap)()
        n = C.c_size_t()
        self._check(self.lib.psbc_get(self._h, key.encode(), rank, out, cap, C.byref(n)), "GetHandles")
        return bytes(out[:n.value]) if n.value else None

    # ------------------------------------------------------------------ teardown
    _BYE = "psb/bye"

    def close(self, linger_s: float = 10.0) -> None:
        """Leave the job.  Every rank says goodbye (a marker in the coordinator's blob store)
        before it disconnects; the rank that LAUNCHED the coordinator keeps it alive until all
        ``world`` ranks have said so (or ``linger_s`` has passed) — otherwise a peer that is
        still reading the answer of its last barrier would find the coordinator gone."""
        if self._h:
            try:
                self.put(self._BYE, b"1")
                if self._proc is not None and self.world > 1:
                    end = time.time() + linger_s
                    while time.time() < end and self._proc.poll() is None:
                        live = set(self.list_workers())  # an evicted (dead) rank is not waited for
                        if all(self.get(self._BYE, r) for r in range(self.world) if r in live):
                            break
                        time.sleep(0.02)
            except ControlError:
                pass  # coordinator already gone: nothing to wait for
            self.lib.psbc_close(self._h)
            self._h = None
        if self._proc is not None:
            _stop(self._proc)
            self._proc = None
        if self._rendezvous:
            try:
                os.unlink(self._rendezvous)
            except OSError:
                pass
            self._rendezvous = None


# ---------------------------------------------------------------------- launching a coordinator
def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _stop(proc: subprocess.Popen) -> None:
    if proc.poll() is None:
        proc.terminate()
        try:
            proc.wait(timeout=3)
        except subprocess.TimeoutExpired:
            proc.kill()


def _die_with_parent():
    """preexec hook: the coordinator must not outlive the rank that launched it."""
    try:
        C.CDLL("libc.so.6").prctl(1, signal.SIGTERM)  # PR_SET_PDEATHSIG
    except OSError:
        pass


def spawn_coordinator(host: str = "127.0.0.1", port: int = 0, ps_address: str = "localhost:50051",
                      log_path: Optional[str] = None,
                      env: Optional[dict] = None) -> Tuple[subprocess.Popen, str]:
    """Start ``build/bin/coordinator`` (same argv as the reference binary,
    src/coordinator_main.cpp:5-24) and wait until it accepts connections."""
    if not COORDINATOR_BIN.exists():
        from .. import _build
        _build.build_binaries()
    port = port or _free_port()
    addr = f"{host}:{port}"
    log = open(log_path, "ab") if log_path else subprocess.DEVNULL
    proc = subprocess.Popen([str(COORDINATOR_BIN), addr, ps_address], stdout=log,
                            stderr=subprocess.STDOUT, env={**os.environ, **(env or {})},
                            preexec_fn=_die_with_parent)
    deadline = time.time() + 10
    while time.time() < deadline:
        if proc.poll() is not None:
            raise ControlError(f"coordinator exited with code {proc.returncode} on {addr}")
        try:
            with socket.create_connection((host, port), timeout=0.2):
                return proc, addr
        except OSError:
            time.sleep(0.02)
    _stop(proc)
    raise ControlError(f"coordinator did not start listening on {addr}")


def _rendezvous_path() -> str:
    # ranks of one torchrun launch share the agent as parent and the master port
    tag = f"{os.getppid()}_{os.environ.get('MASTER_PORT', '0')}"
    return os.path.join(os.environ.get("PSB_RENDEZVOUS_DIR", "/tmp"), f"psb_coordinator_{tag}.json")


def ensure_coordinator(rank: int, world: int, *, heartbeat_ms: Optional[int] = None,
                       timeout_s: float = 60.0) -> CoordinatorClient:
    """Connect this rank to the job's coordinator.

    ``PSB_COORDINATOR=host:port`` names one that somebody started (scripts/deploy_box.sh,
    scripts/start_coordinator.sh).  Otherwise — the torchrun case — rank 0 launches the
    binary on a free port of this box and publishes the address in a rendezvous file that the
    other ranks poll."""
    hb = heartbeat_ms or int(os.environ.get("PSB_HEARTBEAT_MS", "5000"))
    addr = os.environ.get("PSB_COORDINATOR")
    proc, rdv = None, None
    if not addr:
        rdv = _rendezvous_path()
        if rank == 0:
            proc, addr = spawn_coordinator(log_path=os.environ.get("PSB_COORDINATOR_LOG"))
            tmp = rdv + f".{os.getpid()}.tmp"
            with open(tmp, "w") as f:
                json.dump({"address": addr, "pid": proc.pid, "started": time.time()}, f)
            os.replace(tmp, rdv)
            atexit.register(_stop, proc)
        else:
            deadline = time.time() + timeout_s
            while True:
                try:
                    with open(rdv) as f:
                        addr = json.load(f)["address"]
                    # a file left by an earlier launch names a coordinator that is gone: only
                    # an address that answers is accepted, else keep polling for rank 0's
                    host, _, port = addr.rpartition(":")
                    with socket.create_connection((host, int(port)), timeout=0.5):
                        break
                except (OSError, ValueError, KeyError):
                    if time.time() > deadline:
                        raise ControlError(f"no live coordinator address appeared in {rdv}")
                    time.sleep(0.05)
            rdv = None  # only rank 0 removes the file
    client = CoordinatorClient(addr, rank, world, heartbeat_ms=hb, connect_timeout_s=timeout_s)
    client._proc, client._rendezvous = proc, rdv
    return client
