"""Ad-hoc edge check of the CTA-pair GEMM path: M such that the second CTA of the last pair is entirely out of range,
N not a multiple of 256, K not a multiple of the stage width; forward (+bias, ReLU), dgrad (+mask, colsum), wgrad."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from dmlp_b200.ops import kernels as k  # noqa: E402
from test_gpu_gemm import _operand_ref, _rand  # noqa: E402

worst = 0.0
for dtype in (torch.float32, torch.bfloat16):
    for (M, K, N) in ((2100, 200, 2320), (2305, 72, 2304), (4352, 136, 1104)):
        torch.manual_seed(0)
        x = _rand((M, K), dtype, (K + 7) // 8 * 8)
        w = _rand((K, N), dtype, (N + 7) // 8 * 8) * 0.1
        b = torch.randn(N, device="cuda")
        out = torch.full((M, N), float("nan"), device="cuda")
        k.linear_forward(x, w, b, True, out=out)
        ref = torch.relu(_operand_ref(x) @ _operand_ref(w) + b.double())
        e1 = (out.double() - ref).abs().max().item() / max(1.0, ref.abs().max().item())
        # dgrad: dz [M, N2] . w2[K2, N2]^T -> [M, K2] with K2 large (pair path over the GEMM's N = K2)
        dz = _rand((M, 200), dtype, 200)
        w2 = _rand((N, 200), dtype, 200) * 0.1
        h = torch.relu(torch.randn(M, N, device="cuda"))
        o2 = torch.full((M, N), float("nan"), device="cuda")
        cs = torch.zeros(N, device="cuda")
        k.linear_dgrad(dz, w2, h, out=o2, colsum=cs)
        r2 = (_operand_ref(dz) @ _operand_ref(w2).T) * (h > 0)
        e2 = (o2.double() - r2).abs().max().item() / max(1.0, r2.abs().max().item())
        e2c = (cs.double() - r2.sum(0)).abs().max().item() / (max(1.0, r2.abs().max().item()) * M ** 0.5)
        # wgrad: act [B, M] ^T . dz [B, N] -> [M, N]
        act = _rand((200, M), dtype, (M + 7) // 8 * 8)
        dz3 = _rand((200, N), dtype, (N + 7) // 8 * 8)
        o3 = torch.full((M, N), float("nan"), device="cuda")
        k.linear_wgrad(act, dz3, o3)
        r3 = _operand_ref(act).T @ _operand_ref(dz3)
        e3 = (o3.double() - r3).abs().max().item() / max(1.0, r3.abs().max().item())
        torch.cuda.synchronize()
        print(dtype, (M, K, N), f"fwd {e1:.2e} dgrad {e2:.2e} colsum {e2c:.2e} wgrad {e3:.2e}", flush=True)
        worst = max(worst, e1, e2, e2c, e3)
assert worst < 2e-3, worst
print("PAIR_EDGES_OK")
