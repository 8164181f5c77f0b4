"""One launch (after a warm-up launch) of every kernel family of the library on ONE GPU: the ncu target for the
per-kernel evidence under profiles/ncu/ (multi-rank collectives run as emulated ranks: one cooperative launch over
two logical ranks on this GPU; the multi-GPU-only kernels - NVLS, the persistent data-parallel launch, the input
exchange - cannot run under ncu, which is single-GPU, single-process here).

    python scripts/ncu_kernel_zoo.py            # plain run (must exit 0 before the same command runs under ncu)
"""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dmlp_b200.ops import kernels as k  # noqa: E402
from dmlp_b200.runtime import native  # noqa: E402
from dmlp_b200.runtime.comm import Communicator  # noqa: E402

torch.manual_seed(0)
dev = "cuda"
B = 64
x = torch.rand(B, 784, device=dev)
w1 = torch.randn(784, 128, device=dev) * 0.05
b1 = torch.zeros(128, device=dev)
h1 = torch.empty(B, 128, device=dev)
w3 = torch.randn(64, 16, device=dev) * 0.05
b3 = torch.zeros(16, device=dev)
h2 = torch.rand(B, 64, device=dev)
y = torch.randint(0, 10, (B,), device=dev, dtype=torch.int32)
dl = torch.empty(B, 16, device=dev)
loss = torch.zeros(1, device=dev)
hit = torch.zeros(1, device=dev, dtype=torch.int32)
db = torch.zeros(16, device=dev)
dz1 = torch.randn(B, 128, device=dev)
w2 = torch.randn(128, 64, device=dev) * 0.05
dz2 = torch.randn(B, 64, device=dev)
dh1 = torch.empty(B, 128, device=dev)
gb1 = torch.zeros(128, device=dev)
gw1 = torch.empty(784, 128, device=dev)
logits = torch.randn(B, 16, device=dev)
lr = torch.tensor([0.01], device=dev)
wflat, gflat = torch.randn(109440, device=dev), torch.randn(109440, device=dev)
xl = torch.randn(8192, 784, device=dev).to(torch.bfloat16)
wl = (torch.randn(784, 8192, device=dev) * 0.05).to(torch.bfloat16)
bl = torch.zeros(8192, device=dev)
ol = torch.empty(8192, 8192, device=dev, dtype=torch.bfloat16)
src, dst = torch.randn(1 << 20, device=dev), torch.empty(1 << 20, device=dev)
comm = Communicator.local([0, 0], staging_bytes=4 << 20)
ar_in = [torch.randn(109440, device=dev) for _ in range(2)]
ar_out = [torch.empty(109440, device=dev) for _ in range(2)]
lib = native.lib()


def once():
    k.linear_forward(x, w1, b1, True, out=h1)                                   # tf32 forward + bias + ReLU (64 x 784 x 128)
    k.linear_softmax_ce(h2, w3, b3, y, 10, 1.0 / B, dl, loss, hit, dbias=db)    # last layer with the softmax-CE epilogue
    k.linear_dgrad(dz2, w2, h1, out=dh1, colsum=gb1)                            # dgrad + ReLU mask + column sums
    k.linear_wgrad(x, dz1, gw1)                                                 # wgrad (A and B MN-major)
    k.softmax_ce(logits, y, 10, 1.0 / B, dl, loss, hit, colsum=db)              # stand-alone softmax-CE
    k.sgd(wflat, gflat, lr)                                                     # stand-alone SGD
    k.linear_forward(xl, wl, bl, True, out_bf16=ol)                             # bf16 CTA-pair GEMM, TMA-store epilogue (K = 784)
    native.check(lib.dmlp_p2p_copy(0, dst.data_ptr(), src.data_ptr(), src.numel() * 4, None))
    srcs = (C.c_void_p * 2)(ar_in[0].data_ptr(), ar_in[1].data_ptr())
    native.check(lib.dmlp_reduce_n(0, ar_out[0].data_ptr(), srcs, 2, 109440, native.F32, native.SUM, None))
    for algo in ("oneshot", "twoshot", "ring"):                                 # emulated ranks: one cooperative launch
        comm.allreduce([t.data_ptr() for t in ar_in], [t.data_ptr() for t in ar_out], 109440,
                       algo=native.ALGO_BY_NAME[algo])
    torch.cuda.synchronize()


once()
once()
comm.check()
# the cluster-fused step (one launch = 8 steps) through the trainer
from dmlp_b200.parallel.trainer import Trainer, TrainerConfig  # noqa: E402
from dmlp_b200.utils.data import synthetic_mnist  # noqa: E402
t = Trainer(TrainerConfig(dims=(784, 128, 64, 10), batch=64, lr=0.05), device=0)
xs, ys = synthetic_mnist(64 * 16, seed=0)
st = t.fit(xs, ys, 16)
assert np.isfinite(st["loss"])
t.close()
comm.destroy()
print("ok")
