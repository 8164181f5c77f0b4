"""Per-tile timeline of CTA 0 of the persistent GEMM kernel (DMLP_GEMM_PROF=1)."""
import ctypes as C
import os
import sys

os.environ["DMLP_GEMM_PROF"] = "1"
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dmlp_b200.ops import kernels as k  # noqa: E402
from dmlp_b200.runtime import native  # noqa: E402

NAMES = ["prod:slot", "prod:last_tma", "mma:acc_free", "mma:first_full", "mma:last_commit", "epi:acc_ready",
         "epi:drained", "epi:done"]


def main():
    dt = torch.bfloat16
    M = N = 8192
    Ks = [int(a) for a in sys.argv[1:]] or [64, 784, 8192]
    out_h = torch.empty(M, N, device="cuda", dtype=dt)
    for K in Ks:
        x = torch.randn(M, K, device="cuda").to(dt)
        w = (torch.randn(K, N, device="cuda") * 0.05).to(dt)
        for _ in range(3):
            k.linear_forward(x, w, None, False, out_bf16=out_h)
        torch.cuda.synchronize()
        buf = (C.c_ulonglong * (14 * 16))()
        lib = native.lib()
        lib.dmlp_gemm_prof_read.argtypes = [C.POINTER(C.c_ulonglong), C.c_int]
        native.check(lib.dmlp_gemm_prof_read(buf, 14 * 16))
        t0 = buf[0]
        print(f"K={K}: CTA 0 timeline, us since its first slot acquisition")
        print("tile " + " ".join(f"{n:>10s}" for n in NAMES))
        for t in range(14):
            print(f"{t:4d} " + " ".join(f"{(buf[t * 16 + i] - t0) / 1e3:10.2f}" for i in range(8)))


if __name__ == "__main__":
    main()
