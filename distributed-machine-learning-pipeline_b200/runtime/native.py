"""ctypes binding of ``libdmlp_b200.so`` (the C ABI declared in ``include/dmlp_b200.h``).

The library is built in-tree by ``build_native.py`` and loaded from ``_native/``. On a machine
with a GPU a missing library is a hard error (no silent eager fallback): ``lib()`` raises.
ctypes releases the GIL for the duration of every call, so gRPC handler threads can block in
native waits without stalling the interpreter.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO_PATH = os.path.join(os.path.dirname(_HERE), "_native", "libdmlp_b200.so")

OK, ERR_INVALID, ERR_CUDA, ERR_NOT_FOUND, ERR_TIMEOUT, ERR_ABORTED, ERR_UNSUPPORTED = range(7)
F32, BF16, U8, I32, F16 = range(5)
SUM, PROD, MIN, MAX, AVG = range(5)
MAX_RANKS = 8

DTYPE_SIZE = {F32: 4, BF16: 2, U8: 1, I32: 4, F16: 2}


class NativeError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"dmlp native error {code}: {msg}")
        self.code = code
        self.message = msg


class GemmDesc(C.Structure):
    _fields_ = [
        ("dtype", C.c_int), ("a_major", C.c_int), ("b_major", C.c_int),
        ("M", C.c_int), ("N", C.c_int), ("K", C.c_int),
        ("A", C.c_void_p), ("lda", C.c_int64),
        ("B", C.c_void_p), ("ldb", C.c_int64),
        ("out_f32", C.c_void_p), ("ld_out", C.c_int64),
        ("out_bf16", C.c_void_p), ("ld_out_bf16", C.c_int64),
        ("bias", C.c_void_p), ("relu", C.c_int),
        ("mask", C.c_void_p), ("mask_dtype", C.c_int), ("ld_mask", C.c_int64),
        ("colsum", C.c_void_p), ("accumulate", C.c_int), ("block_n", C.c_int),
        ("ce_labels", C.c_void_p), ("ce_classes", C.c_int), ("ce_scale", C.c_float),
        ("ce_logits", C.c_void_p), ("ce_loss_sum", C.c_void_p), ("ce_correct", C.c_void_p),
    ]


class DeviceInfo(C.Structure):
    _fields_ = [
        ("name", C.c_char * 128), ("uuid", C.c_char * 48),
        ("sm_count", C.c_int), ("cc_major", C.c_int), ("cc_minor", C.c_int), ("clock_khz", C.c_int),
        ("multicast_supported", C.c_int), ("smem_per_block_optin", C.c_int),
        ("total_mem", C.c_size_t), ("l2_bytes", C.c_size_t),
    ]


ALGO_AUTO, ALGO_ONESHOT, ALGO_TWOSHOT, ALGO_RING, ALGO_NVLS = range(5)
ALGO_BY_NAME = {"auto": ALGO_AUTO, "oneshot": ALGO_ONESHOT, "twoshot": ALGO_TWOSHOT, "ring": ALGO_RING,
                "nvls": ALGO_NVLS}
ALGO_NAMES = {v: k for k, v in ALGO_BY_NAME.items()}


class AllReduceDesc(C.Structure):
    _fields_ = [
        ("algo", C.c_int), ("dtype", C.c_int), ("op", C.c_int), ("fuse_sgd", C.c_int),
        ("skip_end_barrier", C.c_int), ("scale", C.c_float), ("count", C.c_size_t),
        ("inp", C.c_void_p * MAX_RANKS), ("out", C.c_void_p * MAX_RANKS),
        ("out_bf16", C.c_void_p * MAX_RANKS), ("lr", C.c_void_p * MAX_RANKS),
        ("streams", C.c_void_p * MAX_RANKS), ("prof", C.c_void_p),
    ]


class MlpFusedDesc(C.Structure):
    """dmlp_mlp_fused_desc_t (include/dmlp_b200.h): the cluster-fused step kernel's plan descriptor."""
    _fields_ = [
        ("num_layers", C.c_int), ("dims", C.c_int * 4), ("batch", C.c_int),
        ("x", C.c_void_p), ("labels", C.c_void_p), ("params", C.c_void_p), ("grads", C.c_void_p),
        ("w_off", C.c_size_t * 3), ("b_off", C.c_size_t * 3),
        ("metrics", C.c_void_p), ("debug", C.c_void_p), ("prof", C.c_void_p),
        ("sgd_params", C.c_void_p), ("lr", C.c_void_p), ("x_slots", C.c_int),
        ("xchg_world", C.c_int), ("xchg_rank", C.c_int),
        ("xchg_recv", C.c_void_p * MAX_RANKS), ("xchg_pad", C.c_void_p * MAX_RANKS),
        ("xchg_status", C.c_void_p), ("xchg_timeout_ns", C.c_ulonglong), ("param_count", C.c_size_t),
        ("xchg_reducer_ctas", C.c_int), ("xchg_grads", (C.c_void_p * MAX_RANKS) * 2),
        ("gx_enable", C.c_int),
    ]


_lib = None
_lock = threading.Lock()


def so_path() -> str:
    return _SO_PATH


def available() -> bool:
    return os.path.exists(_SO_PATH)


def _declare(lib) -> None:
    vp, i32, i64, sz, f32 = C.c_void_p, C.c_int, C.c_int64, C.c_size_t, C.c_float
    lib.dmlp_last_error.restype = C.c_char_p
    lib.dmlp_build_info.restype = C.c_char_p
    sigs = {
        "dmlp_gemm": [C.POINTER(GemmDesc), vp],
        "dmlp_softmax_ce": [vp, i64, vp, i32, i32, f32, vp, vp, vp, vp, vp, vp],
        "dmlp_sgd": [vp, vp, sz, vp, f32, vp, vp],
        "dmlp_fill_u32": [vp, C.c_uint32, sz, vp],
        "dmlp_cast_f32_to_bf16": [vp, vp, sz, vp],
        "dmlp_device_count": [C.POINTER(i32)],
        "dmlp_device_info": [i32, C.POINTER(DeviceInfo)],
        "dmlp_peer_matrix": [i32, C.POINTER(i32), C.POINTER(i32), C.POINTER(i32), C.POINTER(i32)],
        "dmlp_enable_peer_access": [i32, C.POINTER(i32)],
        "dmlp_malloc": [i32, sz, C.POINTER(vp)],
        "dmlp_free": [i32, vp],
        "dmlp_host_alloc": [sz, C.POINTER(vp)],
        "dmlp_host_free": [vp],
        "dmlp_memcpy_h2d": [i32, vp, vp, sz, vp],
        "dmlp_memcpy_d2h": [i32, vp, vp, sz, vp],
        "dmlp_memcpy_d2d": [i32, vp, vp, sz, vp],
        "dmlp_memset": [i32, vp, i32, sz, vp],
        "dmlp_stream_create": [i32, C.POINTER(vp)],
        "dmlp_stream_destroy": [i32, vp],
        "dmlp_stream_sync": [i32, vp],
        "dmlp_stream_query": [i32, vp, C.POINTER(i32)],
        "dmlp_stream_wait_event": [i32, vp, vp],
        "dmlp_device_sync": [i32],
        "dmlp_event_create": [i32, i32, C.POINTER(vp)],
        "dmlp_event_destroy": [vp],
        "dmlp_event_record": [i32, vp, vp],
        "dmlp_event_query": [vp, C.POINTER(i32)],
        "dmlp_event_sync": [vp],
        "dmlp_event_elapsed_ms": [vp, vp, C.POINTER(f32)],
        "dmlp_ipc_get_handle": [i32, vp, C.c_char_p],
        "dmlp_ipc_open_handle": [i32, C.c_char_p, C.POINTER(vp)],
        "dmlp_ipc_close_handle": [i32, vp],
        "dmlp_comm_create": [i32, i32, C.POINTER(i32), C.POINTER(i32), sz, sz, C.POINTER(vp)],
        "dmlp_comm_destroy": [vp],
        "dmlp_comm_export": [vp, i32, C.c_char_p],
        "dmlp_comm_import": [vp, i32, C.c_char_p],
        "dmlp_comm_arena_ptr": [vp, i32, i32, C.POINTER(vp), C.POINTER(sz)],
        "dmlp_comm_set_timeout_ms": [vp, C.c_uint64],
        "dmlp_comm_abort": [vp],
        "dmlp_comm_status": [vp, C.POINTER(C.c_uint32)],
        "dmlp_comm_inject_dead_ranks": [vp, C.c_uint32],
        "dmlp_allreduce": [vp, i32, C.POINTER(AllReduceDesc), vp],
        "dmlp_allreduce_pick_algo": [i32, sz],
        "dmlp_allreduce_pick_algo_ex": [i32, sz, i32],
        "dmlp_allreduce_set_table": [C.c_char_p],
        "dmlp_p2p_copy": [i32, vp, vp, sz, vp],
        "dmlp_reduce_n": [i32, vp, C.POINTER(vp), i32, sz, i32, i32, vp],
    }
    for name, args in sigs.items():
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = C.c_int
    # Optional symbols (declared by the modules that use them) are looked up lazily.


def lib():
    """Returns the loaded library, loading it on first use."""
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not os.path.exists(_SO_PATH):
                    raise NativeError(ERR_UNSUPPORTED,
                                      f"{_SO_PATH} is missing; run `python build_native.py` (needs nvcc)")
                l = C.CDLL(_SO_PATH, mode=C.RTLD_GLOBAL)
                _declare(l)
                _lib = l
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        msg = lib().dmlp_last_error().decode("utf-8", "replace")
        raise NativeError(rc, msg)


def build_info() -> str:
    return lib().dmlp_build_info().decode()
