"""Python face of the native communicator (``csrc/collectives.cu``) and device helpers.

``Communicator`` is the object the reference calls a communicator
(``DSML/gpu_coordinator_service/gpu_coordinator_server.go:25-52``): an ordered set of ranks, each
bound to one GPU. Here every rank owns a peer-visible memory *segment* (signal pad, ring staging,
symmetric arena) and the collectives are kernels that read/write the peers' segments directly.

Three ways to build one:
  ``Communicator.local([0, 1, 2, 3])``      one process drives all ranks (gRPC coordinator, sweeps)
  ``Communicator.local([0, 0, 0])``         several logical ranks on ONE GPU ("emulated": a single
                                            cooperative launch plays all ranks; used on 1-GPU boxes)
  ``Communicator.from_process_group(...)``  one process per GPU (torchrun); segments are exchanged
                                            as CUDA-IPC handles over torch.distributed
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence

from . import native

PHASES = {0: "-", 1: "barrier-in", 2: "barrier-out", 3: "ring-data", 4: "ring-ack"}


class CommFailure(RuntimeError):
    """A collective failed (peer never arrived / communicator aborted). Sticky, like the
    reference's FAILED communicator state (gpu_coordinator_server.go:114-118,281-285)."""

    def __init__(self, status: Dict[str, int]):
        self.status = status
        super().__init__(
            "communicator FAILED: rank {rank} waited for rank {peer} in phase {phase_name} (code {code})".format(**status))


def device_count() -> int:
    n = C.c_int(0)
    rc = native.lib().dmlp_device_count(C.byref(n))
    return n.value if rc == 0 else 0


@dataclass
class DeviceProps:
    ordinal: int
    name: str
    uuid: str
    sm_count: int
    cc: str
    clock_khz: int
    multicast: bool
    total_mem: int
    l2_bytes: int


def device_props(dev: int) -> DeviceProps:
    info = native.DeviceInfo()
    native.check(native.lib().dmlp_device_info(dev, C.byref(info)))
    return DeviceProps(dev, info.name.decode(), info.uuid.decode(), info.sm_count,
                       f"{info.cc_major}.{info.cc_minor}", info.clock_khz, bool(info.multicast_supported),
                       info.total_mem, info.l2_bytes)


def peer_matrix(devs: Sequence[int]) -> Dict[str, List[List[int]]]:
    """P2P topology of the given devices: access / native atomics / performance rank (0 = best,
    i.e. NVLink). On an NVSwitch box every off-diagonal entry is uniform."""
    n = len(devs)
    arr = (C.c_int * n)(*devs)
    acc, atom, perf = (C.c_int * (n * n))(), (C.c_int * (n * n))(), (C.c_int * (n * n))()
    native.check(native.lib().dmlp_peer_matrix(n, arr, acc, atom, perf))
    sq = lambda a: [[a[i * n + j] for j in range(n)] for i in range(n)]
    return {"access": sq(acc), "atomics": sq(atom), "perf_rank": sq(perf)}


def set_allreduce_table(spec: Optional[str]) -> None:
    """Overrides rows of the ``algo=auto`` dispatcher table for this process: ``"world:oneshot_max:nvls_min[,...]"`` with
    K / M / G suffixes and ``never`` (e.g. ``"8:256K:32M,2:1M:never"``); ``None`` / ``""`` restores the built-in table plus
    ``DMLP_ALLREDUCE_TABLE``. The built-in thresholds come from sweeps on one NVSwitch box (profiles/r2_allreduce_sweep_*)."""
    native.check(native.lib().dmlp_allreduce_set_table(spec.encode() if spec else None))


def pick_algo(world: int, nbytes: int, nvls_available: bool = False) -> int:
    """``algo=auto`` for ``world`` ranks and a message of ``nbytes`` (no communicator needed)."""
    return native.lib().dmlp_allreduce_pick_algo_ex(world, nbytes, int(nvls_available))


class ArenaAllocator:
    """Bump allocator over the symmetric arena: every rank performs the same sequence of
    allocations, so an offset names the same logical buffer on every rank."""

    def __init__(self, capacity: int, align: int = 256):
        self.capacity, self.align, self.top = capacity, align, 0

    def alloc(self, nbytes: int) -> int:
        off = (self.top + self.align - 1) // self.align * self.align
        if off + nbytes > self.capacity:
            raise MemoryError(f"symmetric arena exhausted: need {nbytes} at {off}, capacity {self.capacity}")
        self.top = off + nbytes
        return off


class Communicator:
    def __init__(self, world: int, local_ranks: Sequence[int], local_devs: Sequence[int],
                 staging_bytes: int = 8 << 20, arena_bytes: int = 0):
        self.world = world
        self.local_ranks = list(local_ranks)
        self.local_devs = list(local_devs)
        self.staging_bytes, self.arena_bytes = staging_bytes, arena_bytes
        n = len(self.local_ranks)
        h = C.c_void_p()
        native.check(native.lib().dmlp_comm_create(
            world, n, (C.c_int * n)(*self.local_ranks), (C.c_int * n)(*self.local_devs),
            staging_bytes, arena_bytes, C.byref(h)))
        self._h = h
        self.arena = ArenaAllocator(arena_bytes)
        self.emulated = n == world and world > 1 and len(set(self.local_devs)) == 1
        self._arena_base: Dict[tuple, int] = {}

    # ------------------------------------------------------------------ constructors
    @classmethod
    def local(cls, devices: Sequence[int], staging_bytes: int = 8 << 20, arena_bytes: int = 0) -> "Communicator":
        return cls(len(devices), list(range(len(devices))), list(devices), staging_bytes, arena_bytes)

    @classmethod
    def from_process_group(cls, device: int, staging_bytes: int = 8 << 20, arena_bytes: int = 0,
                           group=None) -> "Communicator":
        """One rank per process. torch.distributed is used only to exchange 64-byte IPC handles."""
        import torch.distributed as dist
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        comm = cls(world, [rank], [device], staging_bytes, arena_bytes)
        if world > 1:
            handles: List[Optional[bytes]] = [None] * world
            dist.all_gather_object(handles, comm.export_handle(0), group=group)
            for r, hd in enumerate(handles):
                if r != rank:
                    comm.import_handle(r, hd)
            dist.barrier(group=group)
        return comm

    # ------------------------------------------------------------------ plumbing
    def export_handle(self, local_idx: int) -> bytes:
        buf = C.create_string_buffer(64)
        native.check(native.lib().dmlp_comm_export(self._h, local_idx, buf))
        return buf.raw

    def import_handle(self, rank: int, handle: bytes) -> None:
        native.check(native.lib().dmlp_comm_import(self._h, rank, handle))

    def arena_ptr(self, local_idx: int, rank: int, offset: int = 0) -> int:
        key = (local_idx, rank)
        if key not in self._arena_base:
            p, sz = C.c_void_p(), C.c_size_t()
            native.check(native.lib().dmlp_comm_arena_ptr(self._h, local_idx, rank, C.byref(p), C.byref(sz)))
            self._arena_base[key] = p.value
        return self._arena_base[key] + offset

    def set_timeout_ms(self, ms: int) -> None:
        native.check(native.lib().dmlp_comm_set_timeout_ms(self._h, int(ms)))

    def abort(self) -> None:
        native.check(native.lib().dmlp_comm_abort(self._h))

    def inject_dead_ranks(self, mask: int) -> None:
        """Fault injection: the ranks in ``mask`` stop showing up at collectives."""
        native.check(native.lib().dmlp_comm_inject_dead_ranks(self._h, mask))

    def status(self) -> Dict[str, int]:
        out = (C.c_uint32 * 5)()
        native.check(native.lib().dmlp_comm_status(self._h, out))
        return {"code": out[0], "rank": out[1], "peer": out[2], "phase": out[3],
                "phase_name": PHASES.get(out[3], str(out[3])), "aborted": out[4]}

    def check(self) -> None:
        st = self.status()
        if st["code"] != 0 or st["aborted"]:
            raise CommFailure(st)

    def sync(self) -> None:
        for d in sorted(set(self.local_devs)):
            native.check(native.lib().dmlp_device_sync(d))

    # ------------------------------------------------------------------ collectives
    def allreduce(self, inputs: Sequence[Optional[int]], outputs: Sequence[Optional[int]], count: int,
                  dtype: int = native.F32, op: int = native.SUM, algo: int = native.ALGO_AUTO,
                  local_idx: int = -1, stream: int = 0, streams: Optional[Sequence[int]] = None,
                  fuse_sgd: bool = False, lr: Optional[Sequence[Optional[int]]] = None, scale: float = 1.0,
                  out_bf16: Optional[Sequence[Optional[int]]] = None, skip_end_barrier: bool = False,
                  prof: Optional[int] = None) -> None:
        """Launches an all-reduce over raw device pointers (one entry per rank). Asynchronous."""
        d = native.AllReduceDesc()
        d.algo, d.dtype, d.op = algo, dtype, op
        d.fuse_sgd, d.skip_end_barrier, d.scale, d.count = int(fuse_sgd), int(skip_end_barrier), scale, count
        for r in range(self.world):
            d.inp[r] = inputs[r] if r < len(inputs) else None
            d.out[r] = outputs[r] if r < len(outputs) else None
            if lr is not None:
                d.lr[r] = lr[r]
            if out_bf16 is not None:
                d.out_bf16[r] = out_bf16[r]
        if streams is not None:
            for i, s in enumerate(streams):
                d.streams[i] = s
        d.prof = prof
        native.check(native.lib().dmlp_allreduce(self._h, local_idx, C.byref(d), stream or None))

    def pick_algo(self, nbytes: int, nvls_available: bool = False) -> int:
        """What ``algo=auto`` runs for a message of ``nbytes`` (table in csrc/collectives.cu, from the sweeps under
        profiles/); with ``nvls_available`` (buffers in a multicast-bound arena) the table includes NVLS."""
        return native.lib().dmlp_allreduce_pick_algo_ex(self.world, nbytes, int(nvls_available))

    def destroy(self) -> None:
        if getattr(self, "_h", None):
            native.lib().dmlp_comm_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass
