"""CUDA device backend: the reference's per-device ``map[uint64][]byte``
(``DSML/gpu_device_service/gpu_device_server.go:26-49``) kept in real GPU memory.

Each address owns one device allocation made by the native runtime. In a single process every
allocation is addressable from every peer GPU once peer access is enabled (UVA), so collectives
and the P2P copy kernel take these pointers as they are. Host<->device copies are
``cudaMemcpy`` through the runtime (C3/K10 in SURVEY.md); device<->device transfers use the
hand-written ``p2p_copy`` kernel on the device's copy stream and report completion via an event.
"""
from __future__ import annotations

import ctypes as C
import struct
import threading
from typing import Optional, Callable, Dict, Tuple

from ..runtime import native
from .backend import DeviceBackend, MissingBufferError


class CudaBackend(DeviceBackend):
    kind = "cuda"

    def __init__(self, device_id: int, mem_size: int, ordinal: int):
        super().__init__(device_id, mem_size)
        self.ordinal = int(ordinal)
        self._lib = native.lib()
        self._lock = threading.RLock()
        self._bufs: Dict[int, Tuple[int, int, int]] = {}  # addr -> (ptr, size, capacity)
        self._scratch: Dict[str, Tuple[int, int]] = {}
        self._scalars: Dict[str, int] = {}
        s = C.c_void_p()
        native.check(self._lib.dmlp_stream_create(self.ordinal, C.byref(s)))
        self.copy_stream = s.value
        self._versions: Dict[int, int] = {}  # bumped whenever a buffer's contents change (Memcpy, collectives)
        self._external: Dict[int, int] = {}  # address -> capacity of buffers that live in somebody else's allocation
        info = native.DeviceInfo()
        native.check(self._lib.dmlp_device_info(self.ordinal, C.byref(info)))
        self._info = {"cuda-ordinal": str(self.ordinal), "uuid": info.uuid.decode(), "name": info.name.decode(),
                      "sm-count": str(info.sm_count), "multicast": str(int(info.multicast_supported)),
                      "interconnect": "nvlink5-nvswitch"}

    def info(self) -> Dict[str, str]:
        d = super().info()
        d.update(self._info)
        return d

    # ------------------------------------------------------------------ allocation table
    def _alloc(self, nbytes: int) -> int:
        p = C.c_void_p()
        native.check(self._lib.dmlp_malloc(self.ordinal, max(nbytes, 16), C.byref(p)))
        return p.value

    def _reserve(self, addr: int, nbytes: int, keep: bool) -> int:
        """Returns a pointer with capacity >= nbytes for ``addr`` (growing keeps old contents if asked)."""
        ent = self._bufs.get(addr)
        if ent is not None and ent[2] >= nbytes:
            self._bufs[addr] = (ent[0], nbytes if not keep else max(ent[1], nbytes), ent[2])
            return ent[0]
        cap = max(256, (nbytes + 255) // 256 * 256)
        ptr = self._alloc(cap)
        if ent is not None:
            if keep and ent[1] > 0:
                native.check(self._lib.dmlp_memcpy_d2d(self.ordinal, ptr, ent[0], ent[1], None))
            if addr in self._external:  # lived in a symmetric heap slot: the heap's owner frees it, the slot is left behind
                self._external.pop(addr)
            else:
                native.check(self._lib.dmlp_free(self.ordinal, ent[0]))
        size = nbytes if not keep or ent is None else max(ent[1], nbytes)
        self._bufs[addr] = (ptr, size, cap)
        return ptr

    # Memcpy H2D / D2H (gpu_device_server.go:195-230): through a pinned staging buffer with cudaMemcpyAsync on the
    # device's copy stream - DMA straight from/to page-locked memory instead of the driver's bounce copy of a pageable
    # buffer. The stream is drained before returning (an RPC reply promises the bytes are there).
    def _staging(self, nbytes: int) -> int:
        if getattr(self, "_pin_cap", 0) < nbytes:
            if getattr(self, "_pin_ptr", None):
                self._lib.dmlp_host_free(self._pin_ptr)
            cap = max(1 << 20, 1 << (nbytes - 1).bit_length())
            p = C.c_void_p()
            native.check(self._lib.dmlp_host_alloc(cap, C.byref(p)))
            self._pin_ptr, self._pin_cap = p.value, cap
        return self._pin_ptr

    def write(self, addr: int, data: bytes) -> None:
        with self._lock:
            n = len(data)
            ptr = self._reserve(addr, n, keep=False)
            self._bufs[addr] = (ptr, n, self._bufs[addr][2])
            self._versions[addr] = self._versions.get(addr, 0) + 1
            if n:
                pin = self._staging(n)
                C.memmove(pin, bytes(data), n)
                native.check(self._lib.dmlp_memcpy_h2d(self.ordinal, ptr, pin, n, self.copy_stream))
                native.check(self._lib.dmlp_stream_sync(self.ordinal, self.copy_stream))

    def read(self, addr: int, nbytes: int = 0) -> bytes:
        with self._lock:
            ent = self._bufs.get(addr)
            if ent is None:
                raise MissingBufferError(f"no data at address 0x{addr:x}")
            n = ent[1] if nbytes <= 0 or nbytes >= ent[1] else nbytes
            if not n:
                return b""
            pin = self._staging(n)
            native.check(self._lib.dmlp_memcpy_d2h(self.ordinal, pin, ent[0], n, self.copy_stream))
            native.check(self._lib.dmlp_stream_sync(self.ordinal, self.copy_stream))
            return C.string_at(pin, n)

    def touched(self, addr: int) -> None:
        """A kernel wrote the buffer at ``addr`` (e.g. the fused all-reduce+SGD updating the weight blob)."""
        with self._lock:
            self._versions[addr] = self._versions.get(addr, 0) + 1

    def has(self, addr: int) -> bool:
        with self._lock:
            return addr in self._bufs

    def size(self, addr: int) -> int:
        with self._lock:
            ent = self._bufs.get(addr)
            if ent is None:
                raise MissingBufferError(f"no data at address 0x{addr:x}")
            return ent[1]

    def ensure(self, addr: int, nbytes: int) -> None:
        with self._lock:
            ent = self._bufs.get(addr)
            if ent is None or ent[1] < nbytes:
                self._reserve(addr, nbytes, keep=True)  # fresh memory is zeroed by the runtime

    # ------------------------------------------------------------------ symmetric heap slots
    def adopt(self, addr: int, ptr: int, cap: int) -> None:
        """Moves the buffer at ``addr`` into caller-owned device memory (a slot of a symmetric, multicast-bound heap that
        has the same offset on every device of a communicator): contents are copied, the private allocation is freed,
        and from now on Memcpy / collectives / RunForward see the slot. The backend never frees ``ptr``."""
        with self._lock:
            ent = self._bufs.get(addr)
            if ent is None:
                raise MissingBufferError(f"no data at address 0x{addr:x}")
            if ent[0] == ptr:
                return
            if ent[1] > cap:
                raise ValueError(f"buffer at 0x{addr:x} ({ent[1]} bytes) does not fit the {cap}-byte slot")
            if ent[1]:
                native.check(self._lib.dmlp_memcpy_d2d(self.ordinal, ptr, ent[0], ent[1], None))
            if addr not in self._external:
                native.check(self._lib.dmlp_free(self.ordinal, ent[0]))
            self._bufs[addr] = (ptr, ent[1], cap)
            self._external[addr] = cap

    def unadopt(self, addr: int) -> None:
        """Back into a private allocation (the heap that held the slot is going away)."""
        with self._lock:
            ent = self._bufs.get(addr)
            if ent is None or addr not in self._external:
                return
            cap = max(256, (ent[1] + 255) // 256 * 256)
            ptr = self._alloc(cap)
            if ent[1]:
                native.check(self._lib.dmlp_memcpy_d2d(self.ordinal, ptr, ent[0], ent[1], None))
            native.check(self._lib.dmlp_device_sync(self.ordinal))
            self._external.pop(addr)
            self._bufs[addr] = (ptr, ent[1], cap)

    def pointer(self, addr: int) -> int:
        with self._lock:
            ent = self._bufs.get(addr)
            if ent is None:
                raise MissingBufferError(f"no data at address 0x{addr:x}")
            return ent[0]

    # ------------------------------------------------------------------ helpers for collectives
    def scratch(self, name: str, nbytes: int) -> int:
        with self._lock:
            ent = self._scratch.get(name)
            if ent is None or ent[1] < nbytes:
                if ent is not None:
                    native.check(self._lib.dmlp_free(self.ordinal, ent[0]))
                ent = (self._alloc(nbytes), nbytes)
                self._scratch[name] = ent
            return ent[0]

    def scalar_f32(self, name: Optional[str], value: float) -> int:
        """Device-resident fp32 scalar (e.g. the learning rate read by the fused SGD kernels). ``name=None`` takes the
        next of 64 rotating slots, so that launches still in flight (a group submitted without waits) keep theirs."""
        with self._lock:
            if name is None:
                self._scalar_seq = (getattr(self, "_scalar_seq", -1) + 1) % 64
                name = f"_rot{self._scalar_seq}"
            ptr = self._scalars.get(name)
            if ptr is None:
                ptr = self._alloc(16)
                self._scalars[name] = ptr
            native.check(self._lib.dmlp_memcpy_h2d(self.ordinal, ptr, struct.pack("<f", value), 4, None))
            return ptr

    def p2p_copy_to(self, dst: "CudaBackend", send_addr: int, recv_addr: int, nbytes: int) -> Callable[[], bool]:
        """Pushes ``nbytes`` from this device's buffer to ``dst``'s buffer with the P2P kernel.
        Returns a poll function (True once the copy's event has completed)."""
        with self._lock:
            ent = self._bufs.get(send_addr)
            if ent is None:
                raise MissingBufferError(f"no data at address 0x{send_addr:x}")
            if ent[1] < nbytes:
                raise MissingBufferError(f"send buffer 0x{send_addr:x} holds {ent[1]} bytes, need {nbytes}")
            src_ptr = ent[0]
        dst.check_addr(recv_addr)
        with dst._lock:
            dptr = dst._reserve(recv_addr, nbytes, keep=False)
            dst._bufs[recv_addr] = (dptr, nbytes, dst._bufs[recv_addr][2])
            dst._versions[recv_addr] = dst._versions.get(recv_addr, 0) + 1
        if dst.ordinal != self.ordinal:
            devs = (C.c_int * 2)(self.ordinal, dst.ordinal)
            native.check(self._lib.dmlp_enable_peer_access(2, devs))
        native.check(self._lib.dmlp_p2p_copy(self.ordinal, dptr, src_ptr, nbytes, self.copy_stream))
        ev = C.c_void_p()
        native.check(self._lib.dmlp_event_create(self.ordinal, 0, C.byref(ev)))
        native.check(self._lib.dmlp_event_record(self.ordinal, ev, self.copy_stream))

        def poll() -> bool:
            busy = C.c_int(1)
            native.check(self._lib.dmlp_event_query(ev, C.byref(busy)))
            if not busy.value:
                self._lib.dmlp_event_destroy(ev)
            return not busy.value

        return poll

    # ------------------------------------------------------------------ RunForward / RunBackward
    # (semantics: services/compute.py). Everything stays on the device: the weight blob, the batch and the result
    # buffers are wrapped as torch tensors over the backend's allocations, the math runs on the tcgen05 layer
    # (models.MLP: fused bias+ReLU GEMMs, softmax-CE in the last GEMM's epilogue, dgrad / wgrad GEMMs).
    def _tensor(self, addr: int, n_floats: int):
        from ..parallel.data_parallel import arena_tensor
        return arena_tensor(self.pointer(addr), n_floats, self.ordinal)

    def _model(self):
        import torch
        from ..models.mlp import MLP
        from .compute import WEIGHT_ADDR, infer_dims
        nbytes = self.size(WEIGHT_ADDR)
        dims = infer_dims(nbytes, self.model_dims)
        mlp = getattr(self, "_mlp", None)
        if mlp is None or mlp.dims != dims:
            mlp = self._mlp = MLP(dims, device=f"cuda:{self.ordinal}")
            self._mlp_version = None
        ver = (self._versions.get(WEIGHT_ADDR, 0), self.pointer(WEIGHT_ADDR))
        if self._mlp_version == ver:
            return mlp  # the resident padded copy still mirrors the blob (RunForward + RunBackward of one step)
        self._mlp_version = ver
        src = self._tensor(WEIGHT_ADDR, nbytes // 4)
        cur = 0
        with torch.no_grad():  # W1|b1|W2|b2|... (unpadded) -> the padded flat layout the kernels use
            for sl, w, b in zip(mlp.layout.layers, mlp.weights, mlp.biases):
                w[:, :sl.fan_out].copy_(src[cur:cur + sl.fan_in * sl.fan_out].view(sl.fan_in, sl.fan_out))
                cur += sl.fan_in * sl.fan_out
                b[:sl.fan_out].copy_(src[cur:cur + sl.fan_out])
                cur += sl.fan_out
        return mlp

    def run_forward(self, input_addr: int, output_addr: int) -> None:
        import torch
        from .compute import batch_of
        with self._lock, torch.cuda.device(self.ordinal):
            mlp = self._model()
            in0, classes = mlp.dims[0], mlp.dims[-1]
            B = batch_of(self.size(input_addr), in0)
            x = self._tensor(input_addr, B * in0).view(B, in0)
            with torch.no_grad():
                logits = mlp(x)[:, :classes]
            self.check_addr(output_addr)
            self._reserve(output_addr, B * classes * 4, keep=False)
            self._bufs[output_addr] = (self._bufs[output_addr][0], B * classes * 4, self._bufs[output_addr][2])
            self._tensor(output_addr, B * classes).view(B, classes).copy_(logits)
            torch.cuda.synchronize()
            self._last_input = input_addr

    def run_backward(self, gradient_addr: int) -> None:
        import torch
        from .compute import LABEL_ADDR, batch_of, labels_from_bytes
        with self._lock, torch.cuda.device(self.ordinal):
            if getattr(self, "_last_input", None) is None:
                raise ValueError("RunBackward needs a preceding RunForward")
            mlp = self._model()
            in0, classes = mlp.dims[0], mlp.dims[-1]
            B = batch_of(self.size(self._last_input), in0)
            x = self._tensor(self._last_input, B * in0).view(B, in0)
            y = torch.from_numpy(labels_from_bytes(self.read(LABEL_ADDR), B, classes)).to(x.device)
            mlp.flat_grad.zero_()
            loss, _ = mlp.loss(x, y)
            loss.backward()
            lay = mlp.layout
            self.check_addr(gradient_addr)
            nbytes = lay.reference_count * 4
            self._reserve(gradient_addr, nbytes, keep=False)
            self._bufs[gradient_addr] = (self._bufs[gradient_addr][0], nbytes, self._bufs[gradient_addr][2])
            dst = self._tensor(gradient_addr, lay.reference_count)
            cur = 0
            for sl in lay.layers:  # padded flat gradients -> the client's blob order
                gw = mlp.flat_grad[sl.w_off:sl.w_off + sl.fan_in * sl.out_pad].view(sl.fan_in, sl.out_pad)[:, :sl.fan_out]
                dst[cur:cur + sl.fan_in * sl.fan_out].view(sl.fan_in, sl.fan_out).copy_(gw)
                cur += sl.fan_in * sl.fan_out
                dst[cur:cur + sl.fan_out].copy_(mlp.flat_grad[sl.b_off:sl.b_off + sl.fan_out])
                cur += sl.fan_out
            torch.cuda.synchronize()

    def close(self) -> None:
        with self._lock:
            for a_, (ptr, _, _) in self._bufs.items():
                if a_ not in self._external:
                    self._lib.dmlp_free(self.ordinal, ptr)
            self._external.clear()
            for ptr, _ in self._scratch.values():
                self._lib.dmlp_free(self.ordinal, ptr)
            for ptr in self._scalars.values():
                self._lib.dmlp_free(self.ordinal, ptr)
            if getattr(self, "_pin_ptr", None):
                self._lib.dmlp_host_free(self._pin_ptr)
                self._pin_ptr, self._pin_cap = None, 0
            self._bufs.clear()
            self._scratch.clear()
            self._scalars.clear()
            if self.copy_stream:
                self._lib.dmlp_stream_destroy(self.ordinal, self.copy_stream)
                self.copy_stream = None
