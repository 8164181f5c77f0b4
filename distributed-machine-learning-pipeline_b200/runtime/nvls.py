"""NVLS (NVSwitch multicast) arena: memory every rank binds to one multicast object, so kernels can
use ``multimem.ld_reduce`` / ``multimem.st`` (in-switch reduction, hardware broadcast).

``NvlsArena.local(devices, nbytes)``               one process owning all GPUs
``NvlsArena.from_process_group(device, nbytes)``   one process per GPU: rank 0 creates the multicast
        object and hands its POSIX file descriptor to the peers over a Unix-domain socket
        (``SCM_RIGHTS``); torch.distributed carries only the socket path and the barriers.

Raises ``NvlsUnavailable`` when the device/driver/fabric cannot do multicast — callers then stay on the
unicast algorithms (one-shot / two-shot / ring).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, Optional, Sequence

from . import native
from .comm import ArenaAllocator
from .fdpass import broadcast_fd


class NvlsUnavailable(RuntimeError):
    pass


_declared = False


def _lib():
    global _declared
    lib = native.lib()
    if not _declared:
        vp, i32, sz = C.c_void_p, C.c_int, C.c_size_t
        sigs = {
            "dmlp_nvls_supported": [i32, C.POINTER(i32)],
            "dmlp_nvls_create": [i32, sz, C.POINTER(i32), C.POINTER(vp)],
            "dmlp_nvls_import": [i32, sz, i32, C.POINTER(vp)],
            "dmlp_nvls_size": [vp, C.POINTER(sz)],
            "dmlp_nvls_add_device": [vp, i32],
            "dmlp_nvls_bind": [vp, i32, C.POINTER(vp), C.POINTER(vp)],
            "dmlp_nvls_destroy": [vp],
            "dmlp_allreduce_nvls": [vp, i32, C.POINTER(NvlsAllReduceDesc), vp],
        }
        for name, args in sigs.items():
            fn = getattr(lib, name)
            fn.argtypes, fn.restype = args, C.c_int
        _declared = True
    return lib


class NvlsAllReduceDesc(C.Structure):
    _fields_ = [("count", C.c_size_t), ("mc_in", C.c_void_p), ("mc_out", C.c_void_p), ("mc_out_bf16", C.c_void_p),
                ("uc_out", C.c_void_p), ("lr", C.c_void_p), ("scale", C.c_float), ("fuse_sgd", C.c_int), ("oneshot", C.c_int)]


def supported(dev: int) -> bool:
    v = C.c_int(0)
    _lib().dmlp_nvls_supported(dev, C.byref(v))
    return bool(v.value)


def _check(rc: int) -> None:
    if rc == native.ERR_UNSUPPORTED:
        raise NvlsUnavailable(native.lib().dmlp_last_error().decode())
    native.check(rc)


class NvlsArena:
    def __init__(self, handle, nbytes: int, uc: Dict[int, int], mc: int):
        self._h, self.nbytes, self.uc, self.mc = handle, nbytes, uc, mc
        self.alloc = ArenaAllocator(nbytes)

    # ------------------------------------------------------------------ constructors
    @classmethod
    def local(cls, devices: Sequence[int], nbytes: int) -> "NvlsArena":
        lib = _lib()
        if len(set(devices)) != len(devices):
            raise NvlsUnavailable("multicast needs distinct devices (no emulated ranks)")
        if not all(supported(d) for d in devices):
            raise NvlsUnavailable("device does not report CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED")
        h = C.c_void_p()
        _check(lib.dmlp_nvls_create(len(devices), nbytes, None, C.byref(h)))
        try:
            for d in devices:
                _check(lib.dmlp_nvls_add_device(h, d))
            uc, mc = {}, None
            for d in devices:
                u, m = C.c_void_p(), C.c_void_p()
                _check(lib.dmlp_nvls_bind(h, d, C.byref(u), C.byref(m)))
                uc[d], mc = u.value, m.value
        except Exception:
            lib.dmlp_nvls_destroy(h)
            raise
        size = C.c_size_t()
        lib.dmlp_nvls_size(h, C.byref(size))
        return cls(h, size.value, uc, mc)

    @classmethod
    def from_process_group(cls, device: int, nbytes: int, group=None) -> "NvlsArena":
        import torch.distributed as dist
        lib = _lib()
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        flags = [None] * world
        dist.all_gather_object(flags, supported(device), group=group)
        if not all(flags):
            raise NvlsUnavailable("a rank's device does not support multicast")
        h = C.c_void_p()
        fd, err = C.c_int(-1), None
        if rank == 0:
            if lib.dmlp_nvls_create(world, nbytes, C.byref(fd), C.byref(h)) != 0:
                err = native.lib().dmlp_last_error().decode()
        try:
            got = broadcast_fd(fd.value if rank == 0 else None, group=group, tag="dmlp_nvls", error=err)
        except RuntimeError as e:
            raise NvlsUnavailable(str(e)) from None
        if rank != 0:
            try:
                _check(lib.dmlp_nvls_import(world, nbytes, got, C.byref(h)))
            finally:
                os.close(got)
        _check(lib.dmlp_nvls_add_device(h, device))
        dist.barrier(group=group)  # every device is part of the object before anyone binds
        u, m = C.c_void_p(), C.c_void_p()
        _check(lib.dmlp_nvls_bind(h, device, C.byref(u), C.byref(m)))
        dist.barrier(group=group)
        size = C.c_size_t()
        lib.dmlp_nvls_size(h, C.byref(size))
        return cls(h, size.value, {device: u.value}, m.value)

    # ------------------------------------------------------------------ use
    def allreduce(self, comm, local_idx: int, off_in: int, off_out: int, count: int, device: int, stream: int = 0,
                  fuse_sgd: bool = False, lr: Optional[int] = None, scale: float = 1.0,
                  off_out_bf16: Optional[int] = None, oneshot: bool = False) -> None:
        d = NvlsAllReduceDesc()
        d.count = count
        d.mc_in, d.mc_out = self.mc + off_in, self.mc + off_out
        d.mc_out_bf16 = (self.mc + off_out_bf16) if off_out_bf16 is not None else None
        d.uc_out = self.uc[device] + off_out
        d.lr, d.scale, d.fuse_sgd = lr, scale, int(fuse_sgd)
        d.oneshot = int(oneshot)
        native.check(_lib().dmlp_allreduce_nvls(comm._h, local_idx, C.byref(d), stream or None))

    def destroy(self) -> None:
        if getattr(self, "_h", None):
            _lib().dmlp_nvls_destroy(self._h)
            self._h = None
