"""Numerics of the tcgen05 GEMM kernels against plain PyTorch references (GPU only).

Two references per case: (a) the same product computed in fp64 on inputs truncated to the
precision the tensor core consumes (tf32: 10-bit mantissa, truncated) — this pins layout /
descriptor correctness tightly; (b) the plain fp32 product, with the tolerance tf32/bf16 allows.
"""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _trunc_tf32(t: torch.Tensor) -> torch.Tensor:
    return (t.contiguous().view(torch.int32) & ~0x1FFF).view(torch.float32)


def _operand_ref(t: torch.Tensor) -> torch.Tensor:
    return _trunc_tf32(t).double() if t.dtype == torch.float32 else t.double()


def _rand(shape, dtype, ld=None):
    rows, cols = shape
    ld = ld or cols
    buf = torch.randn(rows, ld, device="cuda", dtype=torch.float32).to(dtype)
    return buf[:, :cols]


CASES = [
    # (M(batch), K(in), N(out)) — reference shapes, north-star small, ragged, multi-tile
    (64, 784, 128), (64, 128, 64), (64, 64, 16), (64, 128, 16),
    (128, 256, 256), (200, 104, 72), (512, 1024, 384), (1024, 512, 1024),
    # >= 74 tiles of 256 x 256: the CTA-pair (tcgen05.mma.cta_group::2) path, ragged in M, N and K
    (2304, 200, 2320), (2200, 328, 2560), (2304, 2320, 200), (200, 2304, 2320),
    (2100, 200, 2320),  # the second CTA of the last pair lies entirely past M (scripts/check_gemm_pair_edges.py has more)
]


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("M,K,N", CASES)
def test_forward_bias_relu(dtype, M, K, N):
    from dmlp_b200.ops import kernels as k
    torch.manual_seed(0)
    ldk = (K + 7) // 8 * 8
    ldn = (N + 7) // 8 * 8
    x = _rand((M, K), dtype, ldk)
    w = _rand((K, N), dtype, ldn) * 0.1
    b = torch.randn(N, device="cuda")
    out = torch.full((M, N), float("nan"), device="cuda")
    out16 = torch.zeros((M, N), device="cuda", dtype=torch.bfloat16)
    k.linear_forward(x, w, b, True, out=out, out_bf16=out16)
    torch.cuda.synchronize()
    ref = torch.relu(_operand_ref(x) @ _operand_ref(w) + b.double())
    err = (out.double() - ref).abs().max().item()
    assert err < 2e-3 * max(1.0, ref.abs().max().item()), f"max err {err}"
    assert (out16.double() - ref).abs().max().item() < 2e-2 * max(1.0, ref.abs().max().item())


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("M,K,N", CASES)
def test_dgrad_mask_colsum(dtype, M, K, N):
    """dX[M, K_in] = dZ[M, N_out] · W[K_in, N_out]ᵀ, masked; colsum = column sums of the result."""
    from dmlp_b200.ops import kernels as k
    torch.manual_seed(1)
    ldn = (N + 7) // 8 * 8
    dz = _rand((M, N), dtype, ldn)
    w = _rand((K, N), dtype, ldn) * 0.1
    h = torch.relu(torch.randn(M, K, device="cuda"))
    out = torch.full((M, K), float("nan"), device="cuda")
    cs = torch.zeros(K, device="cuda")
    k.linear_dgrad(dz, w, h, out=out, colsum=cs)
    torch.cuda.synchronize()
    ref = (_operand_ref(dz) @ _operand_ref(w).T) * (h > 0)
    scale = max(1.0, ref.abs().max().item())
    assert (out.double() - ref).abs().max().item() < 2e-3 * scale
    assert (cs.double() - ref.sum(0)).abs().max().item() < 2e-3 * scale * (M ** 0.5)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("B,M,N", CASES)
def test_wgrad(dtype, B, M, N):
    """dW[M_in, N_out] = act[B, M_in]ᵀ · dZ[B, N_out] (both operands MN-major)."""
    from dmlp_b200.ops import kernels as k
    torch.manual_seed(2)
    ldm = (M + 7) // 8 * 8
    ldn = (N + 7) // 8 * 8
    act = _rand((B, M), dtype, ldm)
    dz = _rand((B, N), dtype, ldn)
    out = torch.full((M, N), float("nan"), device="cuda")
    k.linear_wgrad(act, dz, out)
    torch.cuda.synchronize()
    ref = _operand_ref(act).T @ _operand_ref(dz)
    scale = max(1.0, ref.abs().max().item())
    assert (out.double() - ref).abs().max().item() < 2e-3 * scale
    # accumulate mode adds on top
    k.linear_wgrad(act, dz, out, accumulate=True)
    torch.cuda.synchronize()
    assert (out.double() - 2 * ref).abs().max().item() < 4e-3 * scale


@pytest.mark.parametrize("bn", [16, 32, 64, 128, 256])
def test_forward_explicit_block_n(bn):
    from dmlp_b200.ops import kernels as k
    torch.manual_seed(3)
    x = _rand((300, 160), torch.float32)
    w = _rand((160, 512), torch.float32)
    out = torch.empty(300, 512, device="cuda")
    k.linear_forward(x, w, None, False, out=out, block_n=bn)
    torch.cuda.synchronize()
    ref = _operand_ref(x) @ _operand_ref(w)
    assert (out.double() - ref).abs().max().item() < 1e-3 * ref.abs().max().item()


@pytest.mark.parametrize("dtype,out_kind", [(torch.float32, "f32"), (torch.bfloat16, "f32"), (torch.bfloat16, "bf16")])
@pytest.mark.parametrize("M,K,N,bn", [(640, 192, 768, 256), (300, 160, 512, 128), (384, 96, 640, 64),
                                      (2304, 200, 2320, 0), (2100, 264, 2560, 0)])
@pytest.mark.parametrize("bias_relu", [False, True])
def test_tma_store_epilogue(dtype, out_kind, M, K, N, bn, bias_relu):
    """One output tensor and nothing but bias / ReLU after the accumulator: the eight-warp TMA-store epilogue
    (interior tiles) next to the bounds-checked path (edge rows / columns of the same launch), single CTAs and CTA pairs.
    References: (a) fp64 product of the operands as the tensor core sees them, (b) the plain fp32 PyTorch product."""
    if out_kind == "f32" and bn == 64 and dtype == torch.bfloat16:
        pytest.skip("covered by the fp32-operand case")
    from dmlp_b200.ops import kernels as k
    torch.manual_seed(5)
    x = _rand((M, K), dtype, (K + 7) // 8 * 8)
    w = _rand((K, N), dtype, (N + 7) // 8 * 8) * 0.1
    b = torch.randn(N, device="cuda") if bias_relu else None
    out = (torch.full((M, N), float("nan"), device="cuda") if out_kind == "f32"
           else torch.zeros((M, N), device="cuda", dtype=torch.bfloat16))
    if out_kind == "f32":
        k.linear_forward(x, w, b, bias_relu, out=out, block_n=bn)
    else:
        k.linear_forward(x, w, b, bias_relu, out_bf16=out, block_n=bn)
    torch.cuda.synchronize()
    ref = _operand_ref(x) @ _operand_ref(w)
    plain = x.float() @ w.float()
    if bias_relu:
        ref, plain = torch.relu(ref + b.double()), torch.relu(plain + b)
    scale = max(1.0, ref.abs().max().item())
    tol = 2e-3 if out_kind == "f32" else 2e-2
    assert (out.double() - ref).abs().max().item() < tol * scale
    assert (out.double() - plain.double()).abs().max().item() < (2e-2 if dtype == torch.float32 or out_kind == "bf16" else 5e-3) * scale


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("B,K,C,N", [(64, 64, 10, 16), (100, 128, 10, 16), (64, 96, 27, 32), (300, 64, 10, 16)])
def test_fused_linear_softmax_ce(dtype, B, K, C, N):
    """Last-layer GEMM with the softmax-CE epilogue: dlogits, db, loss and hits vs PyTorch."""
    from dmlp_b200.ops import kernels as k
    torch.manual_seed(4)
    x = _rand((B, K), dtype)
    w = torch.zeros(K, N, device="cuda", dtype=dtype)
    w[:, :C] = (torch.randn(K, C, device="cuda") * 0.3).to(dtype)
    bias = torch.zeros(N, device="cuda")
    bias[:C] = torch.randn(C, device="cuda")
    labels = torch.randint(0, C, (B,), device="cuda", dtype=torch.int32)
    dl = torch.full((B, N), float("nan"), device="cuda")
    logits = torch.zeros(B, N, device="cuda")
    loss = torch.zeros(1, device="cuda")
    correct = torch.zeros(1, device="cuda", dtype=torch.int32)
    db = torch.zeros(N, device="cuda")
    k.linear_softmax_ce(x, w, bias, labels, C, 1.0 / B, dl, loss, correct, dbias=db, logits=logits)
    torch.cuda.synchronize()
    z = (_operand_ref(x) @ _operand_ref(w) + bias.double())[:, :C]
    p = torch.softmax(z, 1)
    oh = torch.nn.functional.one_hot(labels.long(), C).double()
    ref_d = (p - oh) / B
    assert (logits[:, :C].double() - z).abs().max().item() < 2e-3
    assert (dl[:, :C].double() - ref_d).abs().max().item() < 1e-4
    assert (dl[:, C:] == 0).all()
    ref_loss = -(torch.log((p * oh).sum(1) + 1e-10)).sum().item()
    assert abs(loss.item() - ref_loss) < 2e-3 * max(1.0, abs(ref_loss))
    hits = (z.argmax(1) == labels.long()).sum().item()
    assert abs(correct.item() - hits) <= 1  # tf32/bf16 rounding may flip a near-tie
    assert (db[:C].double() - ref_d.sum(0)).abs().max().item() < 1e-4
