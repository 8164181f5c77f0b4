"""Device memory backends behind the GPUDevice service.

The reference models device memory as ``map[uint64][]byte``: an address is the *key* of an
independently sized buffer, and ``[minAddr, minAddr+memSize)`` only gates which keys are legal
(``DSML/gpu_device_service/gpu_device_server.go:26-49``). Both backends keep that model:

* ``SimBackend``  — host ``bytearray`` per address. Runs anywhere; it is the CPU oracle and the
  "simulated devices" configuration of the reference.
* ``CudaBackend`` — one CUDA device; every address owns a device allocation made by the native
  runtime (peer-mapped, so collectives and P2P kernels address it from other GPUs directly).
"""
from __future__ import annotations

import threading
from typing import Dict, Optional

MIN_ADDR = 0x1000


class AddressError(ValueError):
    """Address outside [min_addr, max_addr) — surfaces as gRPC InvalidArgument."""


class MissingBufferError(KeyError):
    """No buffer stored at the address — surfaces as gRPC Internal (reference :219)."""


class DeviceBackend:
    kind = "abstract"

    def __init__(self, device_id: int, mem_size: int):
        self.device_id = int(device_id)
        self.mem_size = int(mem_size)
        self.min_addr = MIN_ADDR
        self.max_addr = MIN_ADDR + int(mem_size)
        self.dead = False  # fault injection: a dead device stops answering (services check this)
        self.model_dims = None  # RunForward/RunBackward: model shape; None = inferred from the weight blob's size

    def check_addr(self, addr: int) -> None:
        if addr < self.min_addr or addr >= self.max_addr:
            raise AddressError(f"address 0x{addr:x} out of bounds [0x{self.min_addr:x}, 0x{self.max_addr:x})")

    # -- storage interface -------------------------------------------------------------------
    def write(self, addr: int, data: bytes) -> None:
        raise NotImplementedError

    def read(self, addr: int, nbytes: int = 0) -> bytes:
        raise NotImplementedError

    def has(self, addr: int) -> bool:
        raise NotImplementedError

    def size(self, addr: int) -> int:
        raise NotImplementedError

    def ensure(self, addr: int, nbytes: int) -> None:
        """Make the buffer at ``addr`` at least ``nbytes`` long, zero-extending (the reference's
        padding branch, ``gpu_coordinator_server.go:300-334``)."""
        raise NotImplementedError

    def info(self) -> Dict[str, str]:
        return {"backend": self.kind, "device-id": str(self.device_id)}

    def close(self) -> None:
        pass


class SimBackend(DeviceBackend):
    kind = "sim"

    def __init__(self, device_id: int, mem_size: int):
        super().__init__(device_id, mem_size)
        self._mem: Dict[int, bytearray] = {}
        self._lock = threading.Lock()

    def write(self, addr: int, data: bytes) -> None:
        with self._lock:
            self._mem[addr] = bytearray(data)

    def read(self, addr: int, nbytes: int = 0) -> bytes:
        with self._lock:
            buf = self._mem.get(addr)
            if buf is None:
                raise MissingBufferError(f"no data at address 0x{addr:x}")
            return bytes(buf if nbytes <= 0 or nbytes >= len(buf) else buf[:nbytes])

    def has(self, addr: int) -> bool:
        with self._lock:
            return addr in self._mem

    def size(self, addr: int) -> int:
        with self._lock:
            buf = self._mem.get(addr)
            if buf is None:
                raise MissingBufferError(f"no data at address 0x{addr:x}")
            return len(buf)

    def ensure(self, addr: int, nbytes: int) -> None:
        with self._lock:
            buf = self._mem.setdefault(addr, bytearray())
            if len(buf) < nbytes:
                buf.extend(b"\x00" * (nbytes - len(buf)))

    # -- RunForward / RunBackward (services/compute.py): NumPy oracle ------------------------
    def _compute(self):
        if getattr(self, "_oracle", None) is None:
            from .compute import OracleCompute
            self._oracle = OracleCompute(self)
        return self._oracle

    def run_forward(self, input_addr: int, output_addr: int) -> None:
        self._compute().run_forward(input_addr, output_addr)

    def run_backward(self, gradient_addr: int) -> None:
        self._compute().run_backward(gradient_addr)


def make_backend(kind: str, device_id: int, mem_size: int, cuda_ordinal: Optional[int] = None) -> DeviceBackend:
    if kind == "sim":
        return SimBackend(device_id, mem_size)
    if kind == "cuda":
        from .cuda_backend import CudaBackend
        return CudaBackend(device_id, mem_size, cuda_ordinal if cuda_ordinal is not None else 0)
    raise ValueError(f"unknown backend {kind!r} (expected 'sim' or 'cuda')")
