"""ctypes binding of ``libpsb_kernels.so`` (C ABI: ``include/psb_kernels.h``).

The library is built in-tree by :mod:`._build` and loaded from
``parameter_server_distributed_b200/_native/``.  There is no PyTorch fallback for the device
ops: if the library is missing on a GPU box every op raises.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

MAX_PEERS = 8
MAX_BUCKETS = 512
MAX_GATHER = 96
IPC_HANDLE_BYTES = 64

F32, BF16 = 0, 1
OPT_SGD, OPT_MOMENTUM, OPT_ADAM, OPT_ADAMW = 0, 1, 2, 3

PUSH_BROADCAST = 1
PUSH_NO_WAIT = 2
PUSH_REDUCE_ONLY = 4
PUSH_FROM_REDUCED = 8
PUSH_MASKED = 16
PUSH_SEQLOCK = 32
PUSH_TMA = 64
PUSH_NVLS = 128

PAD_READY_OFF = 0
PAD_VERSION_OFF = MAX_BUCKETS * MAX_PEERS
PAD_BARRIER_OFF = 2 * MAX_BUCKETS * MAX_PEERS
PAD_CLOCK_OFF = PAD_BARRIER_OFF + 64 * MAX_PEERS
PAD_NORM_OFF = PAD_CLOCK_OFF + MAX_PEERS
PAD_ACK_OFF = PAD_NORM_OFF + 4 * MAX_PEERS
PAD_DONE_OFF = PAD_ACK_OFF + MAX_BUCKETS * MAX_PEERS
PAD_WORDS = PAD_DONE_OFF + MAX_PEERS


class PsbOptim(C.Structure):
    _fields_ = [("kind", C.c_int32), ("lr", C.c_float), ("momentum", C.c_float),
                ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float),
                ("weight_decay", C.c_float), ("grad_scale", C.c_float)]


class PsbGatherList(C.Structure):
    _fields_ = [("src", C.c_void_p * MAX_GATHER), ("dst_offset", C.c_uint64 * MAX_GATHER),
                ("numel", C.c_uint64 * MAX_GATHER), ("count", C.c_int32),
                ("src_dtype", C.c_int32), ("dst_dtype", C.c_int32), ("scale", C.c_float)]


class PsbFabric(C.Structure):
    _fields_ = [("rank", C.c_int32), ("world", C.c_int32), ("pad", C.c_void_p * MAX_PEERS)]


class PsbPush(C.Structure):
    _fields_ = [("fabric", PsbFabric), ("grad", C.c_void_p * MAX_PEERS),
                ("param", C.c_void_p * MAX_PEERS), ("master", C.c_void_p), ("m", C.c_void_p),
                ("v", C.c_void_p), ("reduced", C.c_void_p), ("sumsq_acc", C.c_void_p),
                ("slice_elems", C.c_uint64), ("bucket", C.c_int32), ("grad_dtype", C.c_int32),
                ("epoch", C.c_void_p), ("step", C.c_void_p), ("clip_coef", C.c_void_p),
                ("peer_mask", C.c_void_p), ("done_counter", C.c_void_p), ("flags", C.c_int32),
                ("grid", C.c_int32), ("opt", PsbOptim), ("grad_off", C.c_void_p),
                ("block", C.c_int32), ("grad_mc", C.c_void_p), ("param_mc", C.c_void_p)]


class PsbPull(C.Structure):
    _fields_ = [("fabric", PsbFabric), ("param", C.c_void_p * MAX_PEERS),
                ("slice_bytes", C.c_uint64), ("bucket", C.c_int32), ("epoch", C.c_void_p),
                ("grid", C.c_int32), ("skip_wait", C.c_int32), ("seqlock", C.c_int32)]


class PsbAsync(C.Structure):
    _fields_ = [("fabric", PsbFabric), ("num_buckets", C.c_int32), ("ring", C.c_int32),
                ("ring_stride_bytes", C.c_uint64), ("applied", C.c_void_p), ("masks", C.c_void_p),
                ("grad_off", C.c_void_p), ("epoch", C.c_void_p)]


class PsbGemmAG(C.Structure):
    _fields_ = [("fabric", PsbFabric), ("w", C.c_void_p * MAX_PEERS), ("x", C.c_void_p),
                ("bias", C.c_void_p), ("y", C.c_void_p), ("M", C.c_int32), ("N", C.c_int32),
                ("K", C.c_int32), ("out_dtype", C.c_int32), ("bucket", C.c_int32),
                ("epoch", C.c_void_p), ("skip_wait", C.c_int32), ("epilogue", C.c_int32),
                ("aux_out", C.c_void_p), ("bias_dtype", C.c_int32)]


EPI_NONE, EPI_GELU, EPI_DGELU = 0, 1, 2


class PsbGemm(C.Structure):
    _fields_ = [("a", C.c_void_p), ("b", C.c_void_p), ("y", C.c_void_p), ("bias", C.c_void_p),
                ("aux_in", C.c_void_p), ("aux_out", C.c_void_p), ("M", C.c_int32),
                ("N", C.c_int32), ("K", C.c_int32), ("trans_a", C.c_int32),
                ("trans_b", C.c_int32), ("out_dtype", C.c_int32), ("epilogue", C.c_int32),
                ("bias_dtype", C.c_int32), ("lda", C.c_int32), ("ldb", C.c_int32)]


class PsbSparse(C.Structure):
    _fields_ = [("fabric", PsbFabric), ("table", C.c_void_p * MAX_PEERS),
                ("stage_idx", C.c_void_p * MAX_PEERS), ("stage_grad", C.c_void_p * MAX_PEERS),
                ("stage_cnt", C.c_void_p * MAX_PEERS), ("vocab", C.c_int64),
                ("rows_per_rank", C.c_int64), ("dim", C.c_int32), ("cap", C.c_int32),
                ("grad_dtype", C.c_int32), ("reserved", C.c_int32), ("master", C.c_void_p),
                ("m", C.c_void_p), ("v", C.c_void_p), ("slot", C.c_void_p),
                ("touched", C.c_void_p), ("n_touched", C.c_void_p), ("acc", C.c_void_p),
                ("step", C.c_void_p), ("opt", PsbOptim)]


class PsbError(RuntimeError):
    pass


_LIB = None
_LIB_PATH = Path(__file__).resolve().parent / "_native" / "libpsb_kernels.so"


def lib_path() -> Path:
    return _LIB_PATH


def _declare(lib) -> None:
    vp, sz, i32, f32, u32 = C.c_void_p, C.c_size_t, C.c_int, C.c_float, C.c_uint32
    sig = {
        "psb_last_error": (C.c_char_p, []),
        "psb_version": (C.c_int, []),
        "psb_launch_count": (C.c_uint64, []),
        "psb_fill_f32": (i32, [vp, sz, f32, vp]),
        "psb_fill_bf16": (i32, [vp, sz, f32, vp]),
        "psb_cast_f32_to_bf16": (i32, [vp, vp, sz, f32, vp]),
        "psb_cast_bf16_to_f32": (i32, [vp, vp, sz, f32, vp]),
        "psb_gather_to_bucket": (i32, [C.POINTER(PsbGatherList), vp, vp]),
        "psb_optim_step": (i32, [C.POINTER(PsbOptim), vp, vp, vp, vp, i32, vp, sz, vp, vp, vp]),
        "psb_counter_add": (i32, [vp, u32, vp]),
        "psb_sumsq_scratch_floats": (sz, []),
        "psb_sumsq": (i32, [vp, i32, sz, vp, vp, vp]),
        "psb_clip_coef": (i32, [vp, i32, f32, f32, vp, vp, vp]),
        "psb_device_count": (i32, [C.POINTER(C.c_int)]),
        "psb_set_device": (i32, [i32]),
        "psb_heap_alloc": (i32, [sz, C.POINTER(vp)]),
        "psb_heap_free": (i32, [vp]),
        "psb_ipc_export": (i32, [vp, C.c_char_p]),
        "psb_ipc_open": (i32, [C.c_char_p, C.POINTER(vp)]),
        "psb_ipc_close": (i32, [vp]),
        "psb_can_access_peer": (i32, [i32, i32, C.POINTER(C.c_int)]),
        "psb_enable_peer_access": (i32, [i32]),
        "psb_memcpy_d2h": (i32, [vp, vp, sz, vp]),
        "psb_memcpy_h2d": (i32, [vp, vp, sz, vp]),
        "psb_memset": (i32, [vp, i32, sz, vp]),
        "psb_stream_sync": (i32, [vp]),
        "psb_device_sync": (i32, []),
        "psb_pad_bytes": (sz, []),
        "psb_push_fused": (i32, [C.POINTER(PsbPush), vp]),
        "psb_push_ready": (i32, [C.POINTER(PsbPush), vp]),
        "psb_pull_allgather": (i32, [C.POINTER(PsbPull), vp]),
        "psb_wait_versions": (i32, [C.POINTER(PsbFabric), vp, i32, i32, vp]),
        "psb_device_barrier": (i32, [C.POINTER(PsbFabric), i32, vp, vp]),
        "psb_norm_allreduce": (i32, [C.POINTER(PsbFabric), vp, vp, f32, f32, vp, vp, vp]),
        "psb_async_signal": (i32, [C.POINTER(PsbAsync), u32, vp]),
        "psb_async_plan": (i32, [C.POINTER(PsbAsync), vp]),
        "psb_async_ack": (i32, [C.POINTER(PsbAsync), vp]),
        "psb_async_gate": (i32, [C.POINTER(PsbAsync), u32, i32, vp]),
        "psb_set_spin_timeout_ms": (i32, [C.c_uint64]),
        "psb_pad_post": (i32, [C.POINTER(PsbFabric), u32, u32, vp]),
        "psb_bcast_slice_nvls": (i32, [C.POINTER(PsbFabric), vp, vp, C.c_uint64, i32, vp, vp, i32, vp]),
        "psb_vmm_support": (i32, [C.POINTER(C.c_int), C.POINTER(C.c_int)]),
        "psb_vmm_granularity": (i32, [i32, i32, i32, C.POINTER(sz)]),
        "psb_vmm_alloc": (i32, [sz, C.POINTER(vp), C.POINTER(C.c_uint64), C.POINTER(C.c_int)]),
        "psb_vmm_import": (i32, [i32, sz, C.POINTER(vp), C.POINTER(C.c_uint64)]),
        "psb_vmm_release": (i32, [vp, sz, C.c_uint64]),
        "psb_mc_create": (i32, [sz, i32, C.POINTER(C.c_uint64), C.POINTER(C.c_int)]),
        "psb_mc_import": (i32, [i32, C.POINTER(C.c_uint64)]),
        "psb_mc_add_device": (i32, [C.c_uint64]),
        "psb_mc_bind_map": (i32, [C.c_uint64, C.c_uint64, sz, C.POINTER(vp)]),
        "psb_mc_unbind": (i32, [C.c_uint64, sz]),
        "psb_abort_word_create": (i32, [C.POINTER(vp)]),
        "psb_abort_word_destroy": (i32, [vp]),
        "psb_sumsq_noalloc": (i32, [vp, i32, sz, vp, vp]),
        "psb_gemm_select": (i32, [i32]),
        "psb_gemm_bf16_tn": (i32, [vp, vp, vp, vp, i32, i32, i32, i32, vp]),
        "psb_gemm_allgather_bf16": (i32, [C.POINTER(PsbGemmAG), vp]),
        "psb_gemm_bf16": (i32, [C.POINTER(PsbGemm), vp]),
        "psb_colsum_scratch_bytes": (sz, [i32, i32]),
        "psb_colsum_bf16": (i32, [vp, vp, i32, i32, i32, vp, vp]),
        "psb_mse_fwd_bf16": (i32, [vp, vp, sz, vp, vp, vp]),
        "psb_mse_bwd_bf16": (i32, [vp, vp, vp, sz, vp, vp]),
        "psb_sparse_init": (i32, [vp, C.c_int64, vp]),
        "psb_sparse_pull": (i32, [C.POINTER(PsbSparse), vp, C.c_int64, vp, vp]),
        "psb_sparse_stage": (i32, [C.POINTER(PsbSparse), vp, vp, i32, C.c_int64, vp]),
        "psb_sparse_apply": (i32, [C.POINTER(PsbSparse), vp]),
    }
    missing = [name for name in sig if not hasattr(lib, name)]
    if missing:  # a library built from older sources: say so instead of an AttributeError
        raise PsbError(f"{_LIB_PATH} is older than this package (no {', '.join(missing)}); "
                       f"rebuild it with `python __graft_entry__.py`")
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args


def load(build_if_missing: bool = True):
    """Return the loaded library (dlopen once per process)."""
    global _LIB
    if _LIB is not None:
        return _LIB
    if not _LIB_PATH.exists():
        if not build_if_missing or os.environ.get("PSB_NO_BUILD"):
            raise PsbError(f"{_LIB_PATH} is missing; run `python __graft_entry__.py` to build it")
        from . import _build
        _build.build_kernels()
    lib = C.CDLL(str(_LIB_PATH), mode=C.RTLD_GLOBAL)
    _declare(lib)
    _LIB = lib
    return lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().psb_last_error().decode(errors="replace")
        raise PsbError(f"{what or 'libpsb_kernels'} failed (code {rc}): {msg}")


def launch_count() -> int:
    return int(load().psb_launch_count())
