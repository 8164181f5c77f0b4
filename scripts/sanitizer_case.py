"""Small program for compute-sanitizer: elementwise kernels + the three peer-memory all-reduce kernels with two
emulated ranks on one GPU (one cooperative launch each). No tcgen05 kernels: the sanitizer's support for TMEM / TMA is
not something this repo depends on."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dmlp_b200.ops import kernels as k  # noqa: E402
from dmlp_b200.runtime import native  # noqa: E402
from dmlp_b200.runtime.comm import Communicator  # noqa: E402

torch.manual_seed(0)
B, N = 64, 16
logits = torch.randn(B, N, device="cuda")
labels = torch.randint(0, 10, (B,), device="cuda", dtype=torch.int32)
dl = torch.empty_like(logits)
loss = torch.zeros(1, device="cuda")
hit = torch.zeros(1, device="cuda", dtype=torch.int32)
k.softmax_ce(logits, labels, 10, 1.0 / B, dl, loss, hit)
w = torch.ones(5000, device="cuda")
g = torch.ones(5000, device="cuda")
lr = torch.tensor([0.1], device="cuda")
k.sgd(w, g, lr)
h = torch.empty(5000, device="cuda", dtype=torch.bfloat16)
k.cast_bf16(w, h)
torch.cuda.synchronize()
assert abs(w[0].item() - 0.9) < 1e-6

comm = Communicator.local([0, 0])
n = 16384 + 3
for algo in ("oneshot", "twoshot", "ring"):
    ins = [torch.full((n,), float(r + 1), device="cuda") for r in range(2)]
    outs = [torch.zeros(n, device="cuda") for _ in range(2)]
    comm.allreduce([t.data_ptr() for t in ins], [t.data_ptr() for t in outs], n, algo=native.ALGO_BY_NAME[algo])
    torch.cuda.synchronize()
    comm.check()
    assert torch.equal(outs[0], torch.full_like(outs[0], 3.0)) and torch.equal(outs[0], outs[1]), algo
comm.destroy()
print("sanitizer case ok")
