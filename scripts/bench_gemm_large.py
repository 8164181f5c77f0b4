"""Large-GEMM throughput of the tcgen05 kernel against the library (torch.matmul -> cuBLAS), 1 GPU.

Same recipe as MEASURED_PEAKS.json: bf16, best of 10 single launches timed with CUDA events
("burst") and the mean of back-to-back launches over ~1.5 s ("sustained", power-limited).
Shapes are the wide model's (784-8192-8192-10, batch 8192): forward / dgrad / wgrad.
"""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dmlp_b200.ops import kernels as k  # noqa: E402


def burst(fn, n=10):
    best = 1e9
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    for _ in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


def sustained(fn, seconds=1.5):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.time()
    n = 0
    e0.record()
    while time.time() - t0 < seconds:
        for _ in range(20):
            fn()
        n += 20
        torch.cuda.synchronize()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def main():
    dt = torch.bfloat16
    sizes = [(8192, 8192, 8192), (4096, 4096, 4096), (8192, 784, 8192)]
    if len(sys.argv) > 1:
        sizes = [tuple(int(v) for v in a.split("x")) for a in sys.argv[1:]]
    for (M, K, N) in sizes:
        x = torch.randn(M, K, device="cuda").to(dt)          # activations [batch, in]
        w = (torch.randn(K, N, device="cuda") * 0.05).to(dt)  # weights [in, out]
        dz = torch.randn(M, N, device="cuda").to(dt)         # dZ [batch, out]
        bias = torch.zeros(N, device="cuda")
        out_f = torch.empty(M, N, device="cuda")
        out_h = torch.empty(M, N, device="cuda", dtype=dt)
        dx_h = torch.empty(M, K, device="cuda", dtype=dt)
        dw_f = torch.empty(K, N, device="cuda")
        wt = w.t().contiguous()
        cases = {
            "fwd  X.W  -> bf16 (+bias, relu)": (lambda: k.linear_forward(x, w, bias, True, out_bf16=out_h),
                                               lambda: torch.matmul(x, w, out=out_h), 2 * M * N * K),
            "fwd  X.W  -> f32": (lambda: k.linear_forward(x, w, None, False, out=out_f), None, 2 * M * N * K),
            "dgrad dZ.Wt -> bf16 (mask)": (lambda: k.linear_dgrad(dz, w, x, out_bf16=dx_h),
                                           lambda: torch.matmul(dz, wt, out=dx_h), 2 * M * N * K),
            "wgrad Xt.dZ -> f32": (lambda: k.linear_wgrad(x, dz, dw_f), lambda: torch.matmul(x.t(), dz), 2 * M * N * K),
        }
        for name, (ours, lib, flops) in cases.items():
            rec = {"shape_MKN": [M, K, N], "case": name}
            b = burst(ours)
            s = sustained(ours)
            rec["ours_burst_tflops"] = round(flops / b / 1e9, 1)
            rec["ours_sustained_tflops"] = round(flops / s / 1e9, 1)
            if lib is not None:
                b = burst(lib)
                s = sustained(lib)
                rec["cublas_burst_tflops"] = round(flops / b / 1e9, 1)
                rec["cublas_sustained_tflops"] = round(flops / s / 1e9, 1)
            print(json.dumps(rec), flush=True)


if __name__ == "__main__":
    main()
