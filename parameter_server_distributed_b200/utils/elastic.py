"""Failure handling of the GPU data plane.

The reference detects a dead worker (heartbeats -> eviction, /root/reference/src/coordinator.cpp:
52-67) and then does nothing with it: nobody is told, the parameter server keeps waiting
(SURVEY D9).  Here the coordinator client's heartbeat thread watches the registry; when a peer
disappears it stores 1 into the *abort word* — a host-mapped word every peer spin-wait of the
kernels polls — so kernels blocked on the dead rank's flags return instead of trapping, and
``Trainer.step`` raises :class:`PeerLost`.  The CUDA context of the survivors stays healthy: they
leave cleanly, and the job restarts from the last periodic checkpoint with the smaller world
(checkpoints hold full tensors, so any world size can load them).
"""
from __future__ import annotations

import ctypes as C

from .. import _lib


class PeerLost(RuntimeError):
    """A rank of the job was evicted by the coordinator; parameters of the current step are
    undefined — reload the last checkpoint with the surviving ranks."""


class AbortWord:
    """One 32-bit word of host-mapped memory, wired into the spin-waits of the current device."""

    def __init__(self):
        self.lib = _lib.load()
        p = C.c_void_p()
        _lib.check(self.lib.psb_abort_word_create(C.byref(p)), "psb_abort_word_create")
        self.host_ptr = p.value
        self._word = C.c_uint32.from_address(self.host_ptr)

    @property
    def is_set(self) -> bool:
        return self._word.value != 0

    def set(self) -> None:
        self._word.value = 1

    def clear(self) -> None:
        self._word.value = 0

    def close(self) -> None:
        if self.host_ptr:
            self.lib.psb_abort_word_destroy(self.host_ptr)
            self.host_ptr = 0
