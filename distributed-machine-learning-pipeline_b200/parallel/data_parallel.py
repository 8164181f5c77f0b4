"""``DataParallel`` — DP wrapper for ``models.MLP`` in the PyTorch layer.

The reference's "data parallelism" copies one gradient to every device, ring-reduces it and applies
SGD on the client (``DSML/client/client.go:596-647``). Here every rank computes on its own shard;
parameters and gradients live in the communicator's symmetric arena, and ``step()`` runs ONE kernel
per bucket that all-reduces the gradients over peer memory, averages them and applies SGD — no NCCL,
no separate optimizer kernel. With ``overlap=True`` each layer's bucket is launched on a side stream
as soon as that layer's backward has produced its gradients (layer-by-layer overlap).
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch

from ..models.layout import ParamLayout
from ..models.mlp import MLP
from ..runtime import native
from ..runtime.comm import Communicator


class _CudaView:
    """Exposes raw device memory to torch through ``__cuda_array_interface__``."""

    def __init__(self, ptr: int, n_floats: int):
        self.__cuda_array_interface__ = {"shape": (n_floats,), "typestr": "<f4", "data": (ptr, False), "version": 2}


def arena_tensor(ptr: int, n_floats: int, device: int) -> torch.Tensor:
    return torch.as_tensor(_CudaView(ptr, n_floats), device=torch.device("cuda", device))


class DataParallel:
    def __init__(self, dims: Sequence[int], device: int, lr: float = 0.01, overlap: bool = True,
                 allreduce: str = "auto", seed: int = 0, comm: Optional[Communicator] = None):
        import torch.distributed as dist
        self.device, self.lr, self.overlap = device, lr, overlap
        self.algo = native.ALGO_BY_NAME[allreduce]
        lay = ParamLayout(dims)
        nbytes = (lay.total * 4 + 255) // 256 * 256
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.comm = comm
        if self.world > 1 and comm is None:
            self.comm = Communicator.from_process_group(device, arena_bytes=2 * nbytes)
        if self.comm is not None and self.world > 1:
            self.off_w, self.off_g = self.comm.arena.alloc(nbytes), self.comm.arena.alloc(nbytes)
            flat = arena_tensor(self.comm.arena_ptr(0, self.rank, self.off_w), lay.total, device)
            grad = arena_tensor(self.comm.arena_ptr(0, self.rank, self.off_g), lay.total, device)
        else:
            flat = torch.zeros(lay.total, device=f"cuda:{device}")
            grad = torch.zeros_like(flat)
        self.module = MLP(dims, device=f"cuda:{device}", seed=seed, flat_storage=flat, grad_storage=grad)
        self.lr_dev = torch.tensor([lr], device=f"cuda:{device}")
        self.comm_stream = torch.cuda.Stream(device=device)
        self._launched: List[int] = []
        if overlap:
            self.module.layer_ready_hook = self._layer_ready

    # bucket l = [W_l | b_l] (contiguous in the flat layout)
    def _bucket(self, l: int):
        layers = self.module.layout.layers
        lo = layers[l].w_off
        hi = layers[l + 1].w_off if l + 1 < len(layers) else self.module.layout.total
        return lo, hi - lo

    def _reduce(self, lo: int, n: int, stream: torch.cuda.Stream) -> None:
        if self.world == 1:
            from ..ops import kernels
            with torch.cuda.stream(stream):
                kernels.sgd(self.module.flat[lo:lo + n], self.module.flat_grad[lo:lo + n], self.lr_dev)
            return
        ins = [self.comm.arena_ptr(0, r, self.off_g) + 4 * lo for r in range(self.world)]
        outs = [self.comm.arena_ptr(0, r, self.off_w) + 4 * lo for r in range(self.world)]
        lrs = [self.lr_dev.data_ptr() if r == self.rank else None for r in range(self.world)]
        self.comm.allreduce(ins, outs, n, algo=self.algo, local_idx=0, stream=stream.cuda_stream, fuse_sgd=True,
                            lr=lrs, scale=1.0 / self.world)

    def _layer_ready(self, l: int) -> None:
        # Backward of layer l has produced dW_l, db_l and (dgrad) no longer needs W_l.
        self.comm_stream.wait_stream(torch.cuda.current_stream(self.device))
        lo, n = self._bucket(l)
        self._reduce(lo, n, self.comm_stream)
        self._launched.append(l)

    def loss(self, x, y):
        return self.module.loss(x, y)

    def step(self) -> None:
        """Completes the update after ``loss.backward()``: launches whatever buckets are still pending
        (all of them without overlap) and makes the compute stream wait for the collectives."""
        cur = torch.cuda.current_stream(self.device)
        if not self.overlap:
            self.comm_stream.wait_stream(cur)
            self._reduce(0, self.module.layout.total, self.comm_stream)
        cur.wait_stream(self.comm_stream)
        self._launched.clear()
        if self.comm is not None and self.world > 1:
            # the next backward overwrites the gradient buffer peers may still be reading (one-shot):
            # the fused kernels end with a rendezvous unless double-buffered, which this path is not.
            pass

    def set_lr(self, lr: float) -> None:
        self.lr = lr
        self.lr_dev.fill_(lr)
