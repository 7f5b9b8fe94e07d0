"""Symmetric HBM heap + signal pads across the GPUs of one NVSwitch box.

One process per GPU.  Every rank ``cudaMalloc``s one heap segment, exports a CUDA IPC
handle, the handles are exchanged through the control plane (``torch.distributed`` object
all-gather here; the coordinator's ``ExchangeHandles`` extension RPC for the C++ binaries)
and every rank maps every peer's segment.  Sub-allocations are made with the same bump
allocator on every rank, so a buffer has the same offset in every segment and
``peer_ptr(r) = base[r] + offset`` — the addresses the fused kernels load from / store to
over NVLink.

Replaces the reference's NCCL clique (src/nccl_manager.cpp:28-85: one process driving all
GPUs) and its per-iteration cudaMalloc/cudaFree of staging buffers (src/worker.cpp:420-443).
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Callable, List, Optional, Sequence

import torch

from .. import _lib
from .._lib import PsbFabric, check


class _RawCuda:
    """Adapter exposing a raw device allocation through ``__cuda_array_interface__``."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1",
                                         "data": (ptr, False), "version": 3}


def _wrap(ptr: int, nbytes: int, device: torch.device) -> torch.Tensor:
    return torch.as_tensor(_RawCuda(ptr, nbytes), device=device)


@dataclass
class SymBuffer:
    """A sub-allocation present at the same offset in every rank's heap."""
    fabric: "Fabric"
    offset: int
    nbytes: int

    @property
    def local(self) -> torch.Tensor:
        """uint8 view of the local copy."""
        return self.fabric.heap_u8[self.offset:self.offset + self.nbytes]

    def view(self, dtype: torch.dtype) -> torch.Tensor:
        return self.local.view(dtype)

    def peer_ptr(self, r: int) -> int:
        return self.fabric.heap_base[r] + self.offset

    def peer_ptrs(self) -> List[int]:
        return [self.peer_ptr(r) for r in range(self.fabric.world)]

    @property
    def mc_ptr(self) -> int:
        """Multicast address of the buffer (every rank's copy at once), 0 without NVLS."""
        return self.fabric.mc_base + self.offset if self.fabric.mc_base else 0

    def peer(self, r: int) -> torch.Tensor:
        """uint8 view of rank ``r``'s copy (reads / writes go over NVLink)."""
        if r == self.fabric.rank:
            return self.local
        return _wrap(self.peer_ptr(r), self.nbytes, self.fabric.device)

    def free(self) -> None:
        """Give the range back to the heap.  Collective in the same sense as ``alloc``: every
        rank frees the same buffers in the same order, and nothing may still be using it."""
        self.fabric._free(self)


class HeapAllocator:
    """Deterministic first-fit sub-allocator of one heap segment (offsets only, no memory).

    Every rank runs its own instance; ranks that issue the same sequence of ``alloc`` / ``free``
    calls get the same offsets, which is what makes the heap symmetric.  Freed ranges are merged
    with their neighbours and a free range at the end lowers the high-water mark again."""

    def __init__(self, capacity: int):
        self.capacity = int(capacity)
        self.cursor = 0
        self.holes: List[List[int]] = []  # free ranges [offset, nbytes] below the cursor, sorted

    def alloc(self, nbytes: int, align: int = 256) -> int:
        nbytes = int(nbytes)
        for i, (lo, size) in enumerate(self.holes):
            off = (lo + align - 1) // align * align
            if off + nbytes <= lo + size:
                rest = []
                if off > lo:
                    rest.append([lo, off - lo])
                if off + nbytes < lo + size:
                    rest.append([off + nbytes, lo + size - off - nbytes])
                self.holes[i:i + 1] = rest
                return off
        off = (self.cursor + align - 1) // align * align
        if off + nbytes > self.capacity:
            raise MemoryError(f"symmetric heap exhausted: need {off + nbytes} of {self.capacity} B")
        if off > self.cursor:
            self.holes.append([self.cursor, off - self.cursor])
        self.cursor = off + nbytes
        return off

    def free(self, offset: int, nbytes: int) -> None:
        if nbytes <= 0:
            return
        self.holes.append([int(offset), int(nbytes)])
        self.holes.sort()
        merged: List[List[int]] = []
        for lo, size in self.holes:
            if merged and merged[-1][0] + merged[-1][1] >= lo:
                merged[-1][1] = max(merged[-1][1], lo + size - merged[-1][0])
            else:
                merged.append([lo, size])
        if merged and merged[-1][0] + merged[-1][1] >= self.cursor:  # trailing hole: shrink
            self.cursor = merged.pop()[0]
        self.holes = merged

    def reset(self, keep: int) -> None:
        """Forget every allocation above ``keep`` bytes (benchmarks re-using one heap)."""
        self.cursor, self.holes = int(keep), []

    def largest_free(self) -> int:
        return max([self.capacity - self.cursor] + [size for _, size in self.holes])


class _Control:
    """What the fabric needs from the job's control plane while it comes up."""

    def __init__(self, gather_obj: Callable[[str, object], list], barrier: Callable[[str], None]):
        self.gather_obj, self.barrier = gather_obj, barrier


def _exchange_fds(ctl: _Control, rank: int, world: int, mine: List[int]) -> List[List[int]]:
    """POSIX file descriptors cannot ride in an RPC: every rank serves its own descriptors on a
    Unix-domain socket (SCM_RIGHTS), the socket paths travel through the control plane, and
    every rank fetches every peer's.  Returns ``fds[r]`` = rank r's descriptors as valid in
    THIS process (``fds[rank]`` is ``mine``)."""
    import os
    import socket
    import tempfile
    import threading
    path = os.path.join(tempfile.gettempdir(), f"psb_fd_{os.getpid()}_{rank}.sock")
    try:
        os.unlink(path)
    except OSError:
        pass
    srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    srv.bind(path)
    srv.listen(world)

    def serve():
        for _ in range(world - 1):
            conn, _ = srv.accept()
            with conn:
                socket.send_fds(conn, [len(mine).to_bytes(4, "little")], mine)

    th = threading.Thread(target=serve, daemon=True)
    th.start()
    paths = ctl.gather_obj("fabric/fd_sockets", path)
    out: List[List[int]] = [[] for _ in range(world)]
    out[rank] = list(mine)
    for r in range(world):
        if r == rank:
            continue
        with socket.socket(socket.AF_UNIX, socket.SOCK_STREAM) as c:
            c.connect(paths[r])
            _, fds, _, _ = socket.recv_fds(c, 4, 8)
            out[r] = list(fds)
    th.join(timeout=60)
    srv.close()
    ctl.barrier("fabric/fds_done")
    try:
        os.unlink(path)
    except OSError:
        pass
    return out


class Fabric:
    """Peer-mapped heap and signal pad of one rank.

    Two kinds of heap segment: ``vmm`` — ``cuMemCreate`` memory shared as POSIX file
    descriptors, which can also be bound to an NVLS multicast object (``mc_base``: one address
    for the same offset of every rank's segment; ``multimem.*`` kernels) — and the older
    ``cudaMalloc`` + CUDA IPC handles.  ``PSB_HEAP=ipc|vmm`` and ``PSB_NVLS=0`` override the
    automatic choice (VMM and multicast wherever every rank's device supports them)."""

    def __init__(self, rank: int, world: int, device: torch.device, heap_bytes: int,
                 exchange: Optional[Callable[[bytes], Sequence[bytes]]] = None,
                 _local_bases: Optional[List[int]] = None, control: Optional[_Control] = None,
                 vmm: Optional[bool] = None):
        import os
        if not 1 <= world <= _lib.MAX_PEERS:
            raise ValueError(f"world size {world} outside [1, {_lib.MAX_PEERS}]")
        self.rank, self.world, self.device = rank, world, torch.device(device)
        self.lib = _lib.load()
        self.heap_bytes = (int(heap_bytes) + self.pad_bytes() + (2 << 20) - 1) & ~((2 << 20) - 1)
        self._opened: List[int] = []
        self._owns = _local_bases is None
        self.mc_base = 0           # multicast mapping of the whole heap (0 = no NVLS)
        self._vmm: List[tuple] = []  # (ptr, handle) of every VMM mapping of this process
        self._mc_handle = 0
        self.kind = "ipc"
        with torch.cuda.device(self.device):
            if _local_bases is not None:
                # in-process emulation: all "ranks" live on this device (tests on one GPU)
                self.heap_base = list(_local_bases)
            else:
                use_vmm, use_mc = self._negotiate(control, vmm, os.environ)
                if use_vmm:
                    self._build_vmm(control, use_mc)
                else:
                    self._build_ipc(exchange if exchange is not None else (
                        (lambda h: control.gather_obj("fabric/heap", h)) if control else None))
        self.heap_u8 = _wrap(self.heap_base[rank], self.heap_bytes, self.device)
        self._heap = HeapAllocator(self.heap_bytes)
        self.pad = self.alloc(self.pad_bytes(), align=4096)
        self._barrier_seq = torch.zeros(64, dtype=torch.int32, device=self.device)
        self._cfab = PsbFabric()
        self._cfab.rank, self._cfab.world = rank, world
        for r in range(world):
            self._cfab.pad[r] = self.pad.peer_ptr(r)

    # ------------------------------------------------------------------ heap flavours
    def _negotiate(self, control, vmm, env) -> tuple:
        """(use_vmm, use_multicast): the same answer on every rank."""
        v, m = C.c_int(0), C.c_int(0)
        check(self.lib.psb_vmm_support(C.byref(v), C.byref(m)), "psb_vmm_support")
        want = env.get("PSB_HEAP", "").lower()
        ok_v = bool(v.value) and want != "ipc" and vmm is not False
        ok_m = ok_v and bool(m.value) and self.world > 1 and env.get("PSB_NVLS", "1") != "0"
        if self.world > 1:
            if control is None:
                return False, False  # a bare `exchange` callable can only carry IPC handles
            flags = control.gather_obj("fabric/caps", (ok_v, ok_m))
            ok_v = all(f[0] for f in flags)
            ok_m = ok_v and all(f[1] for f in flags)
        if (vmm is True or want == "vmm") and not ok_v:
            raise RuntimeError("a VMM heap was requested but is not available on every rank")
        return ok_v, ok_m

    def _build_ipc(self, exchange) -> None:
        rank, world = self.rank, self.world
        p = C.c_void_p()
        check(self.lib.psb_heap_alloc(self.heap_bytes, C.byref(p)), "psb_heap_alloc")
        local = p.value
        self.heap_base = [0] * world
        self.heap_base[rank] = local
        if world > 1:
            if exchange is None:
                raise ValueError("a multi-rank fabric needs an exchange function")
            handle = C.create_string_buffer(_lib.IPC_HANDLE_BYTES)
            check(self.lib.psb_ipc_export(local, handle), "psb_ipc_export")
            handles = exchange(bytes(handle.raw))
            for r in range(world):
                if r == rank:
                    continue
                q = C.c_void_p()
                check(self.lib.psb_ipc_open(bytes(handles[r]), C.byref(q)),
                      f"psb_ipc_open(peer {r})")
                self.heap_base[r] = q.value
                self._opened.append(q.value)

    def _build_vmm(self, control, use_mc: bool) -> None:
        import os
        rank, world, lib = self.rank, self.world, self.lib
        gran = C.c_size_t()
        check(lib.psb_vmm_granularity(world, int(use_mc), 0, C.byref(gran)), "psb_vmm_granularity")
        g = gran.value
        self.heap_bytes = (self.heap_bytes + g - 1) // g * g
        ptr, handle, fd = C.c_void_p(), C.c_uint64(), C.c_int(-1)
        check(lib.psb_vmm_alloc(self.heap_bytes, C.byref(ptr), C.byref(handle), C.byref(fd)),
              "psb_vmm_alloc")
        self.kind = "vmm"
        self._vmm.append((ptr.value, handle.value))
        self.heap_base = [0] * world
        self.heap_base[rank] = ptr.value
        mine = [fd.value]
        if world == 1:
            os.close(fd.value)
            return
        mc_handle = C.c_uint64(0)
        if use_mc and rank == 0:
            mfd = C.c_int(-1)
            check(lib.psb_mc_create(self.heap_bytes, world, C.byref(mc_handle), C.byref(mfd)),
                  "psb_mc_create")
            mine.append(mfd.value)
        fds = _exchange_fds(control, rank, world, mine)
        for r in range(world):
            if r == rank:
                continue
            q, h = C.c_void_p(), C.c_uint64()
            check(lib.psb_vmm_import(fds[r][0], self.heap_bytes, C.byref(q), C.byref(h)),
                  f"psb_vmm_import(peer {r})")
            self.heap_base[r] = q.value
            self._vmm.append((q.value, h.value))
        if use_mc:
            if rank != 0:
                check(lib.psb_mc_import(fds[0][1], C.byref(mc_handle)), "psb_mc_import")
            self._mc_handle = mc_handle.value
            check(lib.psb_mc_add_device(self._mc_handle), "psb_mc_add_device")
            control.barrier("fabric/mc_added")  # every device is in the team before any bind
            mc = C.c_void_p()
            check(lib.psb_mc_bind_map(self._mc_handle, handle.value, self.heap_bytes, C.byref(mc)),
                  "psb_mc_bind_map")
            self.mc_base = mc.value
            control.barrier("fabric/mc_bound")
        for lst in fds:
            for f in lst:
                try:
                    os.close(f)
                except OSError:
                    pass

    # ------------------------------------------------------------------ construction
    @staticmethod
    def pad_bytes() -> int:
        return _lib.PAD_WORDS * 4

    @classmethod
    def from_coordinator(cls, coord, device: torch.device, heap_bytes: int) -> "Fabric":
        """One rank per process; the handles of the heap segments travel through the
        coordinator's blob exchange (PutHandle / GetHandles + Barrier, proto/psb_ext.proto) —
        the role the reference gives its coordinator for addresses
        (/root/reference/src/worker.cpp:141-186)."""
        ctl = _Control(lambda key, obj: coord.allgather_object(key, obj),
                       lambda name: coord.barrier(name))
        fab = cls(coord.rank, coord.world, device, heap_bytes, control=ctl)
        coord.barrier("fabric/mapped")  # every peer has mapped every segment before first use
        return fab

    @classmethod
    def from_process_group(cls, device: torch.device, heap_bytes: int, group=None) -> "Fabric":
        """One rank per process; IPC handles travel through ``torch.distributed``."""
        import torch.distributed as dist
        if not dist.is_initialized():
            return cls(0, 1, device, heap_bytes)
        rank, world = dist.get_rank(group), dist.get_world_size(group)

        def gather(key, obj):
            out = [None] * world
            dist.all_gather_object(out, obj, group=group)
            return out

        fab = cls(rank, world, device, heap_bytes,
                  control=_Control(gather, lambda name: dist.barrier(group=group)))
        dist.barrier(group=group)  # every peer has mapped every segment before first use
        return fab

    @classmethod
    def local_group(cls, world: int, device: torch.device, heap_bytes: int) -> List["Fabric"]:
        """``world`` emulated ranks inside this process on one GPU (kernel tests: the launches
        of the ranks are serialised by the caller, never concurrent spinning kernels)."""
        device = torch.device(device)
        lib = _lib.load()
        total = (int(heap_bytes) + cls.pad_bytes() + (2 << 20) - 1) & ~((2 << 20) - 1)
        bases = []
        with torch.cuda.device(device):
            for _ in range(world):
                p = C.c_void_p()
                check(lib.psb_heap_alloc(total, C.byref(p)), "psb_heap_alloc")
                bases.append(p.value)
        fabs = [cls(r, world, device, heap_bytes, _local_bases=bases) for r in range(world)]
        fabs[0]._group_bases = bases  # rank 0 frees them all
        return fabs

    # ------------------------------------------------------------------ allocation
    def alloc(self, nbytes: int, align: int = 256) -> SymBuffer:
        """Sub-allocate ``nbytes`` at the same offset on every rank (``HeapAllocator``)."""
        return SymBuffer(self, self._heap.alloc(nbytes, align), int(nbytes))

    def _free(self, buf: SymBuffer) -> None:
        self._heap.free(buf.offset, buf.nbytes)
        buf.nbytes = 0

    def free_bytes(self) -> int:
        """Largest allocation that is certain to succeed (same value on every rank)."""
        return self._heap.largest_free()

    def reset_allocations(self) -> None:
        """Drop everything but the signal pad (benchmarks that rebuild backends on one heap)."""
        self._heap.reset(self.pad.offset + self.pad.nbytes)

    def all_gather(self, t: torch.Tensor) -> List[torch.Tensor]:
        """Copies of every rank's ``t`` (same shape / dtype everywhere), through the symmetric
        heap: stage locally, device barrier, read the peers' staging ranges over NVLink.
        Control-path helper (checkpoints, tests) — replaces ``torch.distributed.all_gather``."""
        if self.world == 1:
            return [t.clone()]
        flat = t.contiguous().view(-1).view(torch.uint8)
        emulated = not self._owns
        if emulated:
            raise RuntimeError("Fabric.all_gather needs one process per rank")
        with torch.cuda.device(self.device):
            stage = self.alloc(flat.numel(), align=256)
            stage.local.copy_(flat)
            if not emulated:
                self.barrier(2)
            out = [stage.peer(r).clone().view(t.dtype).view(t.shape) for r in range(self.world)]
            if not emulated:
                self.barrier(2)
            torch.cuda.synchronize()
            stage.free()
        return out

    def c_fabric(self) -> PsbFabric:
        return self._cfab

    # ------------------------------------------------------------------ K9 helpers
    def barrier(self, slot: int = 0) -> None:
        """Device-side all-rank barrier on the current stream (no host involvement)."""
        with torch.cuda.device(self.device):
            check(self.lib.psb_device_barrier(C.byref(self._cfab), slot,
                                              self._barrier_seq[slot:].data_ptr(),
                                              torch.cuda.current_stream().cuda_stream),
                  "psb_device_barrier")

    def reset_flags(self) -> None:
        """Zero the ready / version / norm / ack words of the local pad between two barriers.

        Flags are monotonic counters compared against per-engine epoch words that start at 0,
        so a new engine on a used fabric must not see the previous engine's values.  Collective:
        every rank calls it while no push / pull kernel is in flight."""
        words = self.pad.view(torch.int32)
        emulated = not self._owns
        with torch.cuda.device(self.device):
            torch.cuda.synchronize()
            if self.world > 1 and not emulated:
                self.barrier(3)
                torch.cuda.synchronize()
            words[:_lib.PAD_BARRIER_OFF].zero_()
            words[_lib.PAD_CLOCK_OFF:].zero_()
            torch.cuda.synchronize()
            if self.world > 1 and not emulated:
                self.barrier(3)
                torch.cuda.synchronize()

    def close(self) -> None:
        with torch.cuda.device(self.device):
            torch.cuda.synchronize()
            for p in self._opened:
                self.lib.psb_ipc_close(p)
            self._opened = []
            if self.kind == "vmm":
                if self.mc_base:
                    self.lib.psb_mc_unbind(self._mc_handle, self.heap_bytes)
                    self.lib.psb_vmm_release(self.mc_base, self.heap_bytes, self._mc_handle)
                    self.mc_base = 0
                for ptr, handle in reversed(self._vmm):
                    self.lib.psb_vmm_release(ptr, self.heap_bytes, handle)
                self._vmm = []
                self.heap_base[self.rank] = 0
            elif self._owns and self.heap_base[self.rank]:
                self.lib.psb_heap_free(self.heap_base[self.rank])
                self.heap_base[self.rank] = 0
            for p in getattr(self, "_group_bases", []):
                self.lib.psb_heap_free(p)
            self._group_bases = []
