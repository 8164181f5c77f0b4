"""Warm, in-graph timings of the GEMM kernel for the training-step shapes (1 GPU).

Each case captures `reps` back-to-back launches of the same GEMM into a CUDA graph and times graph
replays with CUDA events: the number is the steady-state cost of one dependent kernel node
(kernel time + inter-node gap), which is what the training step's critical path is made of.
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dmlp_b200.ops import kernels as k  # noqa: E402


def timed_graph(fn, reps=50, iters=20):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        fn()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for _ in range(reps):
                fn()
        for _ in range(3):
            g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(iters):
            g.replay()
        e1.record(s)
        torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (reps * iters)  # us per launch


def main():
    dt = torch.float32
    out = []
    B = 64

    def fwd(K, N, bn=0):
        x = torch.randn(B, K, device="cuda", dtype=dt)
        w = torch.randn(K, N, device="cuda", dtype=dt)
        b = torch.randn(N, device="cuda")
        o = torch.empty(B, N, device="cuda")
        return lambda: k.linear_forward(x, w, b, True, out=o, block_n=bn)

    def dgrad(K, N, bn=0):  # dz [B,K] @ w[N,K]^T -> [B,N]
        dz = torch.randn(B, K, device="cuda", dtype=dt)
        w = torch.randn(N, K, device="cuda", dtype=dt)
        h = torch.rand(B, N, device="cuda")
        o = torch.empty(B, N, device="cuda")
        cs = torch.zeros(N, device="cuda")
        return lambda: k.linear_dgrad(dz, w, h, out=o, colsum=cs, block_n=bn)

    def wgrad(M, N, bn=0):
        a = torch.randn(B, M, device="cuda", dtype=dt)
        dz = torch.randn(B, N, device="cuda", dtype=dt)
        o = torch.empty(M, N, device="cuda")
        return lambda: k.linear_wgrad(a, dz, o, block_n=bn)

    def empty_kernel():
        t = torch.zeros(4, device="cuda", dtype=torch.int32)
        return lambda: k.fill_u32(t, 0)

    cases = [("empty fill kernel", empty_kernel())]
    for bn in (0, 16, 32, 64, 128):
        cases.append((f"fwd1 64x784x128 bn={bn}", fwd(784, 128, bn)))
    for K in (32, 64, 128, 256, 512, 1024, 2048):
        cases.append((f"fwd 64x{K}x128 bn=32", fwd(K, 128, 32)))
    cases += [("fwd2 64x128x64", fwd(128, 64)), ("fwd3 64x64x16", fwd(64, 16)),
              ("dgrad3 64x16x64", dgrad(16, 64)), ("dgrad2 64x64x128", dgrad(64, 128)),
              ("wgrad3 64->64x16", wgrad(64, 16)), ("wgrad2 128x64", wgrad(128, 64)), ("wgrad1 784x128", wgrad(784, 128)),
              ("wgrad1 784x128 bn=128", wgrad(784, 128, 128)), ("wgrad1 784x128 bn=64", wgrad(784, 128, 64))]
    for name, fn in cases:
        us = timed_graph(fn)
        out.append({"case": name, "us_per_launch": round(us, 3)})
        print(json.dumps(out[-1]), flush=True)


if __name__ == "__main__":
    main()
