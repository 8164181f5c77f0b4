"""GPU diagnostic: runs every GEMM variant over a shape list and prints errors without asserting.

    python scripts/gpu_diag_gemm.py            # spawns one subprocess per (kind, dtype) group
    python scripts/gpu_diag_gemm.py fwd f32    # one group in-process

A failing/trapping kernel poisons only its own group's process.
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

SHAPES = [(64, 784, 128), (64, 128, 64), (64, 64, 16), (128, 32, 16), (128, 64, 32), (128, 256, 256),
          (200, 104, 72), (512, 1024, 384), (4096, 4096, 4096)]


def trunc(t):
    import torch
    if t.dtype != torch.float32:
        return t.double()
    return (t.contiguous().view(torch.int32) & ~0x1FFF).view(torch.float32).double()


def run_group(kind, dt):
    import torch
    from dmlp_b200.ops import kernels as k
    dtype = torch.float32 if dt == "f32" else torch.bfloat16
    print(f"== group {kind} {dt} on {torch.cuda.get_device_name(0)}", flush=True)
    for (M, K, N) in SHAPES:
        torch.manual_seed(0)
        if kind == "fwd":
            a = torch.randn(M, K, device="cuda").to(dtype)
            b = (torch.randn(K, N, device="cuda") * 0.1).to(dtype)
            out = torch.full((M, N), float("nan"), device="cuda")
            k.linear_forward(a, b, None, False, out=out)
            ref = trunc(a) @ trunc(b)
        elif kind == "dgrad":
            a = torch.randn(M, N, device="cuda").to(dtype)
            b = (torch.randn(K, N, device="cuda") * 0.1).to(dtype)
            out = torch.full((M, K), float("nan"), device="cuda")
            k.linear_dgrad(a, b, None, out=out)
            ref = trunc(a) @ trunc(b).T
        else:
            a = torch.randn(M, K, device="cuda").to(dtype)  # act [B=M, in=K]
            b = torch.randn(M, N, device="cuda").to(dtype)  # dz  [B=M, out=N]
            out = torch.full((K, N), float("nan"), device="cuda")
            k.linear_wgrad(a, b, out)
            ref = trunc(a).T @ trunc(b)
        torch.cuda.synchronize()
        err = (out.double() - ref).abs().max().item()
        nan = torch.isnan(out).sum().item()
        print(f"{kind} {dt} M={M} K={K} N={N}: max_abs_err={err:.3e} ref_max={ref.abs().max().item():.3e} nan={nan}", flush=True)
        if kind == "fwd" and M * N * K >= 4096 ** 3:
            # quick timing vs torch.matmul (library) for orientation
            for name, fn in (("dmlp", lambda: k.linear_forward(a, b, None, False, out=out)),
                             ("torch", lambda: torch.matmul(a, b))):
                for _ in range(3):
                    fn()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(10):
                    fn()
                e1.record()
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / 10
                print(f"   {name}: {ms:.3f} ms  {2 * M * N * K / ms / 1e9:.1f} TFLOP/s", flush=True)


if __name__ == "__main__":
    if len(sys.argv) == 3:
        run_group(sys.argv[1], sys.argv[2])
    else:
        for kind in ("fwd", "dgrad", "wgrad"):
            for dt in ("f32", "bf16"):
                r = subprocess.run([sys.executable, __file__, kind, dt], capture_output=True, text=True, timeout=300)
                print(r.stdout)
                if r.returncode != 0:
                    print(f"!! group {kind} {dt} exited {r.returncode}\n{r.stderr[-2000:]}")
