"""Probe: cost per K-block of the GEMM mainloop vs pipeline depth / tile width / dtype (1 GPU)."""
import json, os, sys, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

def run():
    import torch
    from dmlp_b200.ops import kernels as k
    from gpu_gemm_microbench import timed_graph
    for dt in (torch.float32, torch.bfloat16):
        for (M, N, bn) in ((64, 128, 32), (128, 128, 128), (128, 256, 256)):
            res = {}
            for K in (512, 4096):
                x = torch.randn(M, K, device="cuda").to(dt)
                w = torch.randn(K, N, device="cuda").to(dt)
                o = torch.empty(M, N, device="cuda")
                res[K] = timed_graph(lambda: k.linear_forward(x, w, None, False, out=o, block_n=bn))
            ks = 32 if dt == torch.float32 else 64
            per = (res[4096] - res[512]) / ((4096 - 512) / ks)
            print(json.dumps({"stages": os.environ.get("DMLP_GEMM_STAGES", "auto"), "dtype": str(dt)[6:], "M": M, "N": N, "bn": bn,
                              "us_K512": round(res[512], 2), "us_K4096": round(res[4096], 2), "us_per_kblock": round(per, 4)}), flush=True)

if __name__ == "__main__":
    if len(sys.argv) > 1:
        sys.path.insert(0, os.path.join(ROOT, "scripts"))
        run()
    else:
        for st in ("1", "2", "4", "8"):
            env = dict(os.environ, DMLP_GEMM_STAGES=st)
            subprocess.run([sys.executable, __file__, "child"], env=env)
