"""Data-parallel MLP trainer — the public training API.

    from dmlp_b200.parallel.trainer import Trainer, TrainerConfig
    t = Trainer(TrainerConfig(dims=(784, 128, 64, 10), batch=64), device=0)          # one GPU
    t = Trainer(cfg, device=local_rank, process_group=True)                          # under torchrun
    stats = t.fit(x_host, y_host, steps=1000)     # x float32 [N,784] (pinned for async copies), y int32 [N]

Each ``fit`` call runs natively (``csrc/trainer.cu``): per step the batch is copied host→device on a
copy stream, a captured CUDA graph runs forward / softmax-CE / backward / fused all-reduce+SGD, and
{loss, correct} are copied back. Replaces the reference client's loop
(``DSML/client/client.go:516-659``); log lines keep the reference wording.
"""
from __future__ import annotations

import ctypes as C
import logging
import time
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

from ..models.layout import ParamLayout
from ..runtime import native
from ..runtime.comm import Communicator
from ..utils.tracing import traced
from . import schedule as sched

log = logging.getLogger("dmlp.trainer")


def shard_schedule(global_step: int, rank: int, world: int, batch_stride: Optional[int] = None) -> Tuple[int, int]:
    """(first batch, stride) of a rank at ``global_step``: step s of rank r consumes batch ``(global_step + s) * world + r``
    of the shared host dataset, i.e. the ranks split every group of ``world`` consecutive batches (sequential, no
    shuffle, like the reference's single stream of batches, client.go:596-647). ``batch_stride`` overrides the stride
    for rank-private datasets that are walked densely (then every rank starts at ``global_step * stride``)."""
    if batch_stride is None:
        return global_step * world + rank, world
    return global_step * int(batch_stride), int(batch_stride)


class _NativeCfg(C.Structure):
    _fields_ = [("num_layers", C.c_int), ("dims", C.c_int * 9), ("batch", C.c_int), ("dtype", C.c_int),
                ("lr", C.c_float), ("allreduce_algo", C.c_int), ("overlap", C.c_int),
                ("nvls_uc", C.c_void_p), ("nvls_mc", C.c_void_p), ("nvls_bytes", C.c_size_t), ("fused", C.c_int)]


@dataclass
class TrainerConfig:
    dims: Sequence[int] = (784, 128, 64, 10)   # north-star MLP (README shape); reference code: (784,128,10)
    batch: int = 64                            # per GPU (reference batchSize, client.go:23)
    lr: float = 0.01                           # reference learningRate (client.go:27)
    precision: str = "tf32"                    # "tf32" (fp32 storage) | "bf16" (fp32 master + bf16 shadow)
    allreduce: str = "auto"                    # auto | oneshot | twoshot | ring | nvls (falls back to auto if unavailable)
    overlap: bool = False                      # per-layer buckets reduced during backward
    use_graph: bool = True
    fused: int = -1                            # cluster-fused step kernel: -1 auto, 0 off, 1 require, 2 require + separate
                                               # update kernel, 3 require + gradient exchange inside the step kernel (world > 1),
                                               # 4 require + persistent launch: compute cluster + reducer CTAs (world > 1),
                                               # 5 require + global-batch mode: dZ1 / input exchange instead of the dW1
                                               #   all-reduce (world > 1; opt-in, see profiles/r2_dp_modes_scaling.md)
    schedule: str = "constant"
    seed: int = 0
    watchdog_ms: int = 10000
    staging_bytes: int = 8 << 20

    def native(self) -> _NativeCfg:
        c = _NativeCfg()
        c.num_layers = len(self.dims) - 1
        for i, d in enumerate(self.dims):
            c.dims[i] = int(d)
        c.batch, c.lr = int(self.batch), float(self.lr)
        c.dtype = native.BF16 if self.precision == "bf16" else native.F32
        c.allreduce_algo = native.ALGO_BY_NAME[self.allreduce]
        c.overlap = int(self.overlap)
        c.fused = int(self.fused)
        return c


_declared = False


def _declare(lib) -> None:
    global _declared
    if _declared:
        return
    vp, i32, sz = C.c_void_p, C.c_int, C.c_size_t
    cfgp = C.POINTER(_NativeCfg)
    sigs = {
        "dmlp_trainer_param_count": [cfgp, C.POINTER(sz)],
        "dmlp_trainer_arena_bytes": [cfgp, C.POINTER(sz)],
        "dmlp_trainer_layout": [cfgp, i32, C.POINTER(sz), C.POINTER(sz), C.POINTER(i32)],
        "dmlp_trainer_create": [cfgp, i32, vp, i32, i32, i32, sz, C.POINTER(vp)],
        "dmlp_trainer_destroy": [vp],
        "dmlp_trainer_params": [vp, C.POINTER(vp), C.POINTER(sz)],
        "dmlp_trainer_params_updated": [vp],
        "dmlp_trainer_set_lr": [vp, C.c_float],
        "dmlp_trainer_kernels_per_step": [vp, C.POINTER(i32)],
        "dmlp_trainer_launches_last_run": [vp, C.POINTER(C.c_longlong)],
        "dmlp_trainer_run": [vp, vp, vp, sz, sz, sz, i32, vp, vp, i32, vp, vp],
        "dmlp_trainer_load_inputs": [vp, i32, vp, vp],
        "dmlp_trainer_run_resident": [vp, i32, i32, vp, vp],
        "dmlp_trainer_eval": [vp, vp, vp, sz, C.POINTER(C.c_double), C.POINTER(C.c_longlong)],
        "dmlp_trainer_buffer": [vp, i32, i32, C.POINTER(vp)],
        "dmlp_trainer_set_fault": [vp, C.c_longlong],
        "dmlp_trainer_state": [vp, C.POINTER(i32), C.POINTER(i32)],
        "dmlp_trainer_pipelined_layers": [vp, C.POINTER(i32)],
    }
    for name, args in sigs.items():
        fn = getattr(lib, name)
        fn.argtypes, fn.restype = args, C.c_int
    _declared = True


class PinnedArray:
    """NumPy array backed by pinned (page-locked) host memory from the native runtime."""

    def __init__(self, shape, dtype):
        self.dtype = np.dtype(dtype)
        self.shape = tuple(shape)
        nbytes = int(np.prod(self.shape)) * self.dtype.itemsize
        p = C.c_void_p()
        native.check(native.lib().dmlp_host_alloc(max(nbytes, 16), C.byref(p)))
        self.ptr = p.value
        buf = (C.c_char * max(nbytes, 16)).from_address(self.ptr)
        self.array = np.frombuffer(buf, dtype=self.dtype, count=int(np.prod(self.shape))).reshape(self.shape)

    def free(self):
        if self.ptr:
            self.array = None
            native.lib().dmlp_host_free(self.ptr)
            self.ptr = None


class DeviceDataset:
    """A dataset resident in device memory (kernel-only timing: inputs stream from HBM, not PCIe)."""

    def __init__(self, device: int, x: np.ndarray, y: np.ndarray):
        self.device, self.n = device, x.shape[0]
        self.shape_x = x.shape
        lib = native.lib()
        px, py = C.c_void_p(), C.c_void_p()
        native.check(lib.dmlp_malloc(device, x.nbytes, C.byref(px)))
        native.check(lib.dmlp_malloc(device, y.nbytes, C.byref(py)))
        x = np.ascontiguousarray(x, dtype=np.float32)
        y = np.ascontiguousarray(y, dtype=np.int32)
        native.check(lib.dmlp_memcpy_h2d(device, px, x.ctypes.data, x.nbytes, None))
        native.check(lib.dmlp_memcpy_h2d(device, py, y.ctypes.data, y.nbytes, None))
        self.x_ptr, self.y_ptr, self.nbytes = px.value, py.value, x.nbytes + y.nbytes

    def free(self):
        if self.x_ptr:
            native.lib().dmlp_free(self.device, self.x_ptr)
            native.lib().dmlp_free(self.device, self.y_ptr)
            self.x_ptr = self.y_ptr = None


def pin(arr: np.ndarray) -> PinnedArray:
    out = PinnedArray(arr.shape, arr.dtype)
    out.array[...] = arr
    return out


class Trainer:
    def __init__(self, cfg: TrainerConfig, device: int = 0, process_group: bool = False,
                 comm: Optional[Communicator] = None, local_idx: int = 0):
        self.cfg = cfg
        self.device = device
        self.lib = native.lib()
        _declare(self.lib)
        self.layout = ParamLayout(cfg.dims)
        ncfg = cfg.native()
        self.rank, self.world = 0, 1
        self._owns_comm = False
        self.nvls = None
        self.comm = comm
        if process_group:
            import torch.distributed as dist
            self.rank, self.world = dist.get_rank(), dist.get_world_size()
            if self.world > 1 and comm is None:
                need = C.c_size_t()
                native.check(self.lib.dmlp_trainer_arena_bytes(C.byref(ncfg), C.byref(need)))
                self.comm = Communicator.from_process_group(device, staging_bytes=cfg.staging_bytes,
                                                            arena_bytes=need.value)
                self.comm.set_timeout_ms(cfg.watchdog_ms)
                self._owns_comm = True
            if self.world > 1 and cfg.allreduce == "nvls":
                from ..runtime.nvls import NvlsArena, NvlsUnavailable
                need = C.c_size_t()
                native.check(self.lib.dmlp_trainer_arena_bytes(C.byref(ncfg), C.byref(need)))
                try:
                    self.nvls = NvlsArena.from_process_group(device, need.value)
                    ncfg.nvls_uc, ncfg.nvls_mc, ncfg.nvls_bytes = self.nvls.uc[device], self.nvls.mc, self.nvls.nbytes
                except NvlsUnavailable as e:
                    log.warning("NVLS unavailable (%s); using the unicast all-reduce", e)
                    ncfg.allreduce_algo = native.ALGO_AUTO
                    self.cfg.allreduce = "auto"
        elif comm is not None:
            self.rank, self.world = comm.local_ranks[local_idx], comm.world
        h = C.c_void_p()
        native.check(self.lib.dmlp_trainer_create(
            C.byref(ncfg), device, self.comm._h if (self.comm is not None and self.world > 1) else None,
            local_idx, self.rank, self.world, 0, C.byref(h)))
        self._h = h
        self.global_step = 0
        self._pinned: Dict[int, Tuple[PinnedArray, PinnedArray]] = {}
        self.set_params(self.layout.init_reference(cfg.seed))
        if process_group and self.world > 1:
            # every rank's receive areas / signal words are initialised before anybody's first step writes into them
            import torch.distributed as dist
            native.check(self.lib.dmlp_device_sync(self.device))
            dist.barrier()

    # ------------------------------------------------------------------ parameters
    def _param_ptr(self) -> Tuple[int, int]:
        p, n = C.c_void_p(), C.c_size_t()
        native.check(self.lib.dmlp_trainer_params(self._h, C.byref(p), C.byref(n)))
        return p.value, n.value

    def get_params(self) -> np.ndarray:
        """Flat padded fp32 parameter buffer (see ``models.layout.ParamLayout``)."""
        ptr, n = self._param_ptr()
        out = np.empty(n, dtype=np.float32)
        native.check(self.lib.dmlp_device_sync(self.device))
        native.check(self.lib.dmlp_memcpy_d2h(self.device, out.ctypes.data, ptr, n * 4, None))
        return out

    def set_params(self, flat: np.ndarray) -> None:
        ptr, n = self._param_ptr()
        flat = np.ascontiguousarray(flat, dtype=np.float32)
        assert flat.size == n, f"expected {n} floats, got {flat.size}"
        native.check(self.lib.dmlp_memcpy_h2d(self.device, ptr, flat.ctypes.data, n * 4, None))
        native.check(self.lib.dmlp_trainer_params_updated(self._h))

    def weights(self) -> List[Tuple[np.ndarray, np.ndarray]]:
        return [(w.copy(), b.copy()) for w, b in self.layout.views(self.get_params())]

    def save_weights(self, path: str) -> None:
        """Writes the reference client's weight blob (W1‖b1‖W2‖b2…, LE float32; client.go:204-218)."""
        with open(path, "wb") as f:
            f.write(self.layout.to_reference_bytes(self.get_params()))

    def load_weights(self, path: str) -> None:
        with open(path, "rb") as f:
            self.set_params(self.layout.from_reference_bytes(f.read()))

    def save_checkpoint(self, path: str, epoch: int = 0, schedule: str = "constant", **extra) -> None:
        """Weight blob in the reference layout + ``<path>.json`` (step counter, epoch, lr, schedule, shape)."""
        from ..utils import checkpoint as ck
        ck.write_checkpoint(path, self.layout.to_reference_bytes(self.get_params()), dims=self.cfg.dims,
                            batch=self.cfg.batch, global_step=self.global_step, epoch=epoch, lr=self.cfg.lr,
                            schedule=schedule, world=self.world, extra=extra or None)

    def load_checkpoint(self, path: str) -> dict:
        """Restores weights and the step counter (LR schedules are functions of it); returns the sidecar dict."""
        from ..utils import checkpoint as ck
        blob, meta = ck.read_checkpoint(path, expect_dims=self.cfg.dims)
        self.set_params(self.layout.from_reference_bytes(blob))
        self.global_step = int(meta.get("global_step", self.global_step))
        return meta

    # ------------------------------------------------------------------ failure handling
    def inject_fault(self, step: int) -> None:
        """Fault injection: this rank stops taking part in the gradient exchange from global step ``step`` on (counted
        from this trainer's creation; -1 switches it off), as if its GPU had died. The peers' in-kernel watchdogs fire
        after ``watchdog_ms``; their ``fit`` raises ``native.DmlpError`` (TIMEOUT) and the trainer turns FAILED — the
        training-path analogue of the reference's device-failure test (gpu_coordinator_server_test.go:370-429)."""
        native.check(self.lib.dmlp_trainer_set_fault(self._h, int(step)))
        self._fault_step = int(step)  # (per-layer / two-kernel paths: applied by fit() through the communicator's dead mask)

    @property
    def failed(self) -> bool:
        f, m = C.c_int(), C.c_int()
        native.check(self.lib.dmlp_trainer_state(self._h, C.byref(f), C.byref(m)))
        return bool(f.value)

    @property
    def dp_mode(self) -> int:
        """Data-parallel variant in use: 0 none/per-layer, 1 two-kernel chain, 3 in-cluster exchange, 4 persistent
        compute + reducer launch, 5 global-batch mode."""
        f, m = C.c_int(), C.c_int()
        native.check(self.lib.dmlp_trainer_state(self._h, C.byref(f), C.byref(m)))
        return m.value

    def set_lr(self, lr: float) -> None:
        native.check(self.lib.dmlp_trainer_set_lr(self._h, float(lr)))

    @property
    def pipelined_layers(self) -> int:
        """Layers whose weight gradient goes through the fused wgrad GEMM -> reduce-scatter -> SGD -> all-gather
        pipeline (``csrc/wgrad_rs.cuh``) instead of a bucket all-reduce."""
        n = C.c_int32()
        native.check(self.lib.dmlp_trainer_pipelined_layers(self._h, C.byref(n)))
        return n.value

    @property
    def kernels_per_step(self) -> int:
        n = C.c_int()
        native.check(self.lib.dmlp_trainer_kernels_per_step(self._h, C.byref(n)))
        return n.value

    @property
    def launches_last_run(self) -> int:
        """Kernel launches of the most recent ``fit`` on the cluster-fused path (one launch may run several
        steps); 0 on the per-layer path, where it is ``kernels_per_step * steps``."""
        n = C.c_longlong()
        native.check(self.lib.dmlp_trainer_launches_last_run(self._h, C.byref(n)))
        return n.value

    # ------------------------------------------------------------------ training
    def _host_ptrs(self, x, y) -> Tuple[int, int, int]:
        """Accepts NumPy arrays, PinnedArray or a DeviceDataset (as x; y ignored); returns (x_ptr, y_ptr, n)."""
        if isinstance(x, DeviceDataset):
            assert x.shape_x[1] == self.cfg.dims[0]
            return x.x_ptr, x.y_ptr, x.n
        xa = x.array if isinstance(x, PinnedArray) else x
        ya = y.array if isinstance(y, PinnedArray) else y
        assert xa.dtype == np.float32 and ya.dtype == np.int32 and xa.flags["C_CONTIGUOUS"] and ya.flags["C_CONTIGUOUS"]
        assert xa.shape[1] == self.cfg.dims[0] and xa.shape[0] == ya.shape[0]
        xp = x.ptr if isinstance(x, PinnedArray) else xa.ctypes.data
        yp = y.ptr if isinstance(y, PinnedArray) else ya.ctypes.data
        return xp, yp, xa.shape[0]

    @traced("dmlp.Trainer.fit")
    def fit(self, x, y, steps: int, first_batch: Optional[int] = None, lr_schedule: Optional[sched.Schedule] = None,
            return_metrics: bool = True, device_timing: bool = False,
            batch_stride: Optional[int] = None) -> Dict[str, float]:
        """Runs ``steps`` optimisation steps. Rank r consumes batches r, r+world, r+2·world, … of the
        host dataset (sequential, no shuffle, partial last batch dropped — like the reference).
        ``batch_stride=1`` walks a rank-private dataset densely instead."""
        fs = getattr(self, "_fault_step", -1)
        if (fs >= 0 and self.comm is not None and self.world > 1 and self.dp_mode in (0, 1)
                and self.global_step <= fs < self.global_step + steps and first_batch is None):
            # Fault injection on the paths whose collectives are separate kernels (per-layer buckets, two-kernel chain):
            # run up to the fault step, then mark this rank dead in its communicator - its all-reduce kernels return
            # at once from then on and the peers' watchdogs fire. (The fused data-parallel kernels take the step natively.)
            k = fs - self.global_step
            self._fault_step = -1
            if k > 0:
                self.fit(x, y, k, lr_schedule=lr_schedule, return_metrics=return_metrics, batch_stride=batch_stride)
            self.comm.inject_dead_ranks(1 << self.rank)
            self.cfg.use_graph = False  # captured graphs froze the collectives' arguments: the dead rank runs eagerly
            return self.fit(x, y, steps - k, lr_schedule=lr_schedule, return_metrics=return_metrics,
                            device_timing=device_timing, batch_stride=batch_stride)
        xp, yp, n = self._host_ptrs(x, y)
        nb = n // self.cfg.batch
        if nb == 0:
            raise ValueError("dataset smaller than one batch")
        first, stride = shard_schedule(self.global_step, self.rank, self.world, batch_stride)
        if first_batch is not None:
            first = first_batch
        lrs = None
        if lr_schedule is not None:
            lrs = np.array([lr_schedule(self.global_step + s) for s in range(steps)], dtype=np.float32)
        metrics = np.zeros(2 * steps, dtype=np.float32) if return_metrics else None
        e0 = e1 = None
        if device_timing:
            e0, e1 = C.c_void_p(), C.c_void_p()
            native.check(self.lib.dmlp_event_create(self.device, 1, C.byref(e0)))
            native.check(self.lib.dmlp_event_create(self.device, 1, C.byref(e1)))
        t0 = time.perf_counter()
        native.check(self.lib.dmlp_trainer_run(
            self._h, xp, yp, nb, first % nb, stride, steps, lrs.ctypes.data if lrs is not None else None,
            metrics.ctypes.data if metrics is not None else None, int(self.cfg.use_graph), e0, e1))
        dt = time.perf_counter() - t0
        self.global_step += steps
        out = {"steps": steps, "seconds": dt, "samples_per_s": steps * self.cfg.batch * self.world / dt if dt > 0 else 0.0}
        if device_timing:
            ms = C.c_float()
            native.check(self.lib.dmlp_event_elapsed_ms(e0, e1, C.byref(ms)))
            out["device_ms"] = ms.value
            self.lib.dmlp_event_destroy(e0)
            self.lib.dmlp_event_destroy(e1)
        if metrics is not None and steps:
            m = metrics.reshape(steps, 2)
            out["loss"] = float(m[:, 0].mean() / self.cfg.batch)
            out["accuracy"] = float(m[:, 1].sum() / (steps * self.cfg.batch))
            out["last_loss"] = float(m[-1, 0] / self.cfg.batch)
            out["per_step_loss"] = (m[:, 0] / self.cfg.batch)
        return out

    def train_epochs(self, x, y, epochs: int, steps_per_epoch: Optional[int] = None,
                     lr_schedule: Optional[sched.Schedule] = None) -> List[Dict[str, float]]:
        """Epoch loop with the reference's per-epoch log line (client.go:650-652)."""
        _, _, n = self._host_ptrs(x, y)
        spe = steps_per_epoch or (n // (self.cfg.batch * self.world))
        hist = []
        if self.rank == 0:
            log.info("Starting MLP training...")
        for ep in range(1, epochs + 1):
            st = self.fit(x, y, spe, lr_schedule=lr_schedule)
            hist.append(st)
            if self.rank == 0:
                log.info("Epoch %d complete: Avg Loss: %.4f, Accuracy: %.2f%%", ep, st["loss"], 100 * st["accuracy"])
        if self.rank == 0:
            log.info("Training complete.")
        return hist

    @traced("dmlp.Trainer.evaluate")
    def evaluate(self, x, y, label: str = "Final Test Accuracy") -> Dict[str, float]:
        """Forward-only pass over all full batches (reference ``testModel``, client.go:462-502)."""
        xp, yp, n = self._host_ptrs(x, y)
        nb = n // self.cfg.batch
        loss, correct = C.c_double(), C.c_longlong()
        native.check(self.lib.dmlp_trainer_eval(self._h, xp, yp, nb, C.byref(loss), C.byref(correct)))
        tot = nb * self.cfg.batch
        acc = correct.value / max(tot, 1)
        if self.rank == 0:
            log.info("%s: %.2f%%", label, 100 * acc)  # client.go:501
        return {"loss": loss.value / max(tot, 1), "accuracy": acc, "samples": tot}

    # ------------------------------------------------------------------ kernel-only timing helpers
    def load_resident_inputs(self, x: np.ndarray, y: np.ndarray) -> None:
        B = self.cfg.batch
        for par in range(2):
            xs = np.ascontiguousarray(x[par * B:(par + 1) * B], dtype=np.float32)
            ys = np.ascontiguousarray(y[par * B:(par + 1) * B], dtype=np.int32)
            native.check(self.lib.dmlp_trainer_load_inputs(self._h, par, xs.ctypes.data, ys.ctypes.data))

    def run_resident(self, steps: int) -> float:
        """Runs ``steps`` steps on the device-resident inputs; returns device milliseconds (CUDA events)."""
        e0, e1 = C.c_void_p(), C.c_void_p()
        native.check(self.lib.dmlp_event_create(self.device, 1, C.byref(e0)))
        native.check(self.lib.dmlp_event_create(self.device, 1, C.byref(e1)))
        native.check(self.lib.dmlp_trainer_run_resident(self._h, steps, int(self.cfg.use_graph), e0, e1))
        ms = C.c_float()
        native.check(self.lib.dmlp_event_elapsed_ms(e0, e1, C.byref(ms)))
        self.lib.dmlp_event_destroy(e0)
        self.lib.dmlp_event_destroy(e1)
        self.global_step += steps
        return ms.value

    def debug_buffer(self, kind: int, index: int, count: int) -> np.ndarray:
        p = C.c_void_p()
        native.check(self.lib.dmlp_trainer_buffer(self._h, kind, index, C.byref(p)))
        out = np.empty(count, dtype=np.float32)
        native.check(self.lib.dmlp_device_sync(self.device))
        native.check(self.lib.dmlp_memcpy_d2h(self.device, out.ctypes.data, p.value, count * 4, None))
        return out

    def close(self) -> None:
        if getattr(self, "_h", None):
            self.lib.dmlp_trainer_destroy(self._h)
            self._h = None
        if self.nvls is not None:
            self.nvls.destroy()
            self.nvls = None
        if self._owns_comm and self.comm is not None:
            self.comm.destroy()
            self.comm = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
