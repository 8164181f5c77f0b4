"""Smallest program that runs the CTA-pair GEMM (8192^3 bf16, + bias + ReLU -> bf16) a few times: the ncu target."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dmlp_b200.ops import kernels as k  # noqa: E402

M = K = N = 8192
x = torch.randn(M, K, device="cuda").to(torch.bfloat16)
w = (torch.randn(K, N, device="cuda") * 0.05).to(torch.bfloat16)
b = torch.zeros(N, device="cuda")
o = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
for _ in range(6):
    k.linear_forward(x, w, b, True, out_bf16=o)
torch.cuda.synchronize()
print("ok")
