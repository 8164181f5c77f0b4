"""ncu target for the collectives: the one-shot all-reduce + average + SGD kernel on the small model's gradient
(437 544 bytes), two EMULATED ranks on one GPU (a single cooperative launch plays both ranks, so one process on one GPU
can be profiled; real multi-GPU timings are in profiles/r1_allreduce_sweep_*.jsonl)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dmlp_b200.runtime import native  # noqa: E402
from dmlp_b200.runtime.comm import Communicator  # noqa: E402

n = 109386
comm = Communicator.local([0, 0])
g = [torch.randn(n, device="cuda") for _ in range(2)]
w = [torch.ones(n, device="cuda") for _ in range(2)]
lr = [torch.tensor([0.01], device="cuda") for _ in range(2)]
for algo in ("oneshot", "twoshot", "ring"):
    for _ in range(3):
        comm.allreduce([t.data_ptr() for t in g], [t.data_ptr() for t in w], n, algo=native.ALGO_BY_NAME[algo],
                       fuse_sgd=True, lr=[t.data_ptr() for t in lr], scale=0.5)
    torch.cuda.synchronize()
    comm.check()
assert torch.equal(w[0], w[1])
comm.destroy()
print("ok")
