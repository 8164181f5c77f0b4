"""Per-tile time of the persistent GEMM kernel as a function of K (mainloop length) and epilogue flavour.

t_tile = t_kernel / ceil(tiles / 148). With a double-buffered accumulator the expectation is
t_tile ~ max(mainloop(K), epilogue); the intercept at small K is the epilogue, the slope the cost per k-block.
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dmlp_b200.ops import kernels as k  # noqa: E402


def best(fn, n=7):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    b = 1e9
    for _ in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        b = min(b, e0.elapsed_time(e1))
    return b * 1e3


def main():
    dt = torch.bfloat16
    M = N = 8192
    waves = (M // 128) * (N // 256) / 148
    out_h = torch.empty(M, N, device="cuda", dtype=dt)
    out_f = torch.empty(M, N, device="cuda")
    bias = torch.zeros(N, device="cuda")
    out_pad = torch.empty(M, N + 64, device="cuda", dtype=dt)[:, :N]
    out_fpad = torch.empty(M, N + 32, device="cuda")[:, :N]
    mask = torch.randn(M, N, device="cuda").to(dt)
    cs = torch.zeros(N, device="cuda")
    for K in (64, 784):
        x = torch.randn(M, K, device="cuda").to(dt)
        w = (torch.randn(K, N, device="cuda") * 0.05).to(dt)
        wt = (torch.randn(N, K, device="cuda") * 0.05).to(dt)
        cases = {
            "bf16 out": lambda: k.linear_forward(x, w, None, False, out_bf16=out_h),
            "bf16 out + bias + relu": lambda: k.linear_forward(x, w, bias, True, out_bf16=out_h),
            "f32 out": lambda: k.linear_forward(x, w, None, False, out=out_f),
            "colsum only": lambda: k.gemm(x, w, a_major=0, b_major=1, M=M, N=N, K=K, colsum=cs),
            "dgrad: bf16 out + mask": lambda: k.linear_dgrad(x, wt, mask, out_bf16=out_h),
            "bf16 out, ld padded by 64": lambda: k.linear_forward(x, w, None, False, out_bf16=out_pad),
            "f32 out, ld padded by 32": lambda: k.linear_forward(x, w, None, False, out=out_fpad),
        }
        for name, fn in cases.items():
            if name.startswith("dgrad"):
                # dz [M, K] . wt[N, K]^T -> [M, N]
                pass
            us = best(fn)
            print(json.dumps({"K": K, "k_blocks": (K + 63) // 64, "case": name, "us": round(us, 1),
                              "us_per_tile": round(us / waves, 2)}), flush=True)


if __name__ == "__main__":
    main()
