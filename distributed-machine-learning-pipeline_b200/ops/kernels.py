"""Tensor-level wrappers over the native sm_100a kernels.

Every function takes CUDA ``torch.Tensor`` arguments, launches on the current torch stream and
returns without synchronising. There is no PyTorch fallback: if the native library is missing on a
GPU machine these raise (``runtime.native.NativeError``).

Layout convention (reference ``DSML/client/client.go:36-58,118,134``): weights are ``[in][out]``
row-major. With activations ``[batch][features]`` row-major this makes
  forward  X·W      : A K-major,  B MN-major
  dgrad    dZ·Wᵀ    : A K-major,  B K-major
  wgrad    Aᵀ·dZ    : A MN-major, B MN-major
and the tcgen05 shared-memory descriptors express all three without materialising a transpose.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from ..runtime import native

_DT = {torch.float32: native.F32, torch.bfloat16: native.BF16}


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _check_2d(t: torch.Tensor, name: str) -> None:
    if t.dim() != 2 or t.stride(1) != 1:
        raise ValueError(f"{name} must be 2-D with unit inner stride, got shape {tuple(t.shape)} strides {t.stride()}")
    if not t.is_cuda:
        raise ValueError(f"{name} must be a CUDA tensor")


def gemm(a: torch.Tensor, b: torch.Tensor, *, a_major: int, b_major: int, M: int, N: int, K: int,
         out: Optional[torch.Tensor] = None, out_bf16: Optional[torch.Tensor] = None,
         bias: Optional[torch.Tensor] = None, relu: bool = False,
         mask: Optional[torch.Tensor] = None, colsum: Optional[torch.Tensor] = None,
         accumulate: bool = False, block_n: int = 0, ce_labels: Optional[torch.Tensor] = None,
         ce_classes: int = 0, ce_scale: float = 1.0, ce_logits: Optional[torch.Tensor] = None,
         ce_loss_sum: Optional[torch.Tensor] = None, ce_correct: Optional[torch.Tensor] = None) -> None:
    """D[M,N] = epilogue(A·B) on tcgen05. See ``include/dmlp_b200.h`` (dmlp_gemm) for semantics.
    ``ce_labels`` switches the epilogue to fused softmax cross-entropy (out/out_bf16 receive dlogits)."""
    _check_2d(a, "a")
    _check_2d(b, "b")
    if a.dtype != b.dtype or a.dtype not in _DT:
        raise ValueError(f"operands must both be float32 (tf32 math) or bfloat16, got {a.dtype}/{b.dtype}")
    d = native.GemmDesc()
    d.dtype = _DT[a.dtype]
    d.a_major, d.b_major = a_major, b_major
    d.M, d.N, d.K = M, N, K
    d.A, d.lda = a.data_ptr(), a.stride(0)
    d.B, d.ldb = b.data_ptr(), b.stride(0)
    if out is not None:
        _check_2d(out, "out")
        assert out.dtype == torch.float32
        d.out_f32, d.ld_out = out.data_ptr(), out.stride(0)
    if out_bf16 is not None:
        _check_2d(out_bf16, "out_bf16")
        assert out_bf16.dtype == torch.bfloat16
        d.out_bf16, d.ld_out_bf16 = out_bf16.data_ptr(), out_bf16.stride(0)
    if bias is not None:
        assert bias.dtype == torch.float32 and bias.numel() >= N
        d.bias = bias.data_ptr()
    d.relu = int(relu)
    if mask is not None:
        _check_2d(mask, "mask")
        d.mask, d.mask_dtype, d.ld_mask = mask.data_ptr(), _DT[mask.dtype], mask.stride(0)
    if colsum is not None:
        assert colsum.dtype == torch.float32 and colsum.numel() >= N
        d.colsum = colsum.data_ptr()
    d.accumulate = int(accumulate)
    d.block_n = block_n
    if ce_labels is not None:
        assert ce_labels.dtype == torch.int32 and ce_labels.numel() >= M
        d.ce_labels, d.ce_classes, d.ce_scale = ce_labels.data_ptr(), ce_classes, float(ce_scale)
        d.ce_logits, d.ce_loss_sum, d.ce_correct = _ptr(ce_logits), _ptr(ce_loss_sum), _ptr(ce_correct)
    native.check(native.lib().dmlp_gemm(C.byref(d), _stream()))


def linear_softmax_ce(x: torch.Tensor, w: torch.Tensor, bias: torch.Tensor, labels: torch.Tensor,
                      num_classes: int, scale: float, dlogits: Optional[torch.Tensor],
                      loss_sum: Optional[torch.Tensor], correct: Optional[torch.Tensor],
                      dbias: Optional[torch.Tensor] = None, logits: Optional[torch.Tensor] = None,
                      dlogits_bf16: Optional[torch.Tensor] = None) -> None:
    """Last layer + loss in one kernel: logits = x @ w + bias stay in TMEM/registers; the epilogue
    emits dlogits=(softmax-onehot)*scale, the bias gradient, the loss sum and the hit count.
    Reference client.go:130-165,352-382."""
    B, K = x.shape
    N = w.shape[1]
    gemm(x, w, a_major=0, b_major=1, M=B, N=N, K=K, out=dlogits, out_bf16=dlogits_bf16, bias=bias,
         colsum=dbias, ce_labels=labels, ce_classes=num_classes, ce_scale=scale, ce_logits=logits,
         ce_loss_sum=loss_sum, ce_correct=correct)


def linear_forward(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor], relu: bool,
                   out: Optional[torch.Tensor] = None, out_bf16: Optional[torch.Tensor] = None,
                   block_n: int = 0) -> None:
    """out = act(x @ w + bias); x [B, in], w [in, out]. Reference client.go:112-141."""
    B, K = x.shape
    N = w.shape[1]
    assert w.shape[0] == K
    gemm(x, w, a_major=0, b_major=1, M=B, N=N, K=K, out=out, out_bf16=out_bf16, bias=bias,
         relu=relu, block_n=block_n)


def linear_dgrad(dz: torch.Tensor, w: torch.Tensor, mask: Optional[torch.Tensor],
                 out: Optional[torch.Tensor] = None, out_bf16: Optional[torch.Tensor] = None,
                 colsum: Optional[torch.Tensor] = None, block_n: int = 0) -> None:
    """out = (dz @ w.T) * (mask > 0); dz [B, out], w [in, out] → [B, in]. Reference client.go:178-189.

    ``colsum`` (+= column sums of the masked result) is the bias gradient of the *previous* layer.
    """
    B, K = dz.shape
    N = w.shape[0]
    assert w.shape[1] == K
    gemm(dz, w, a_major=0, b_major=0, M=B, N=N, K=K, out=out, out_bf16=out_bf16, mask=mask,
         colsum=colsum, block_n=block_n)


def linear_wgrad(act: torch.Tensor, dz: torch.Tensor, out: torch.Tensor, accumulate: bool = False,
                 out_bf16: Optional[torch.Tensor] = None, block_n: int = 0) -> None:
    """out = act.T @ dz; act [B, in], dz [B, out] → [in, out]. Reference client.go:167-176,191-199."""
    B, M = act.shape
    N = dz.shape[1]
    assert dz.shape[0] == B
    gemm(act, dz, a_major=1, b_major=1, M=M, N=N, K=B, out=out, out_bf16=out_bf16,
         accumulate=accumulate, block_n=block_n)


def softmax_ce(logits: torch.Tensor, labels: torch.Tensor, num_classes: int, scale: float,
               dlogits: Optional[torch.Tensor], loss_sum: Optional[torch.Tensor],
               correct: Optional[torch.Tensor], colsum: Optional[torch.Tensor] = None,
               dlogits_bf16: Optional[torch.Tensor] = None) -> None:
    """Fused softmax + cross-entropy forward/backward. Reference client.go:76-93,144-165,352-382."""
    _check_2d(logits, "logits")
    assert logits.dtype == torch.float32 and labels.dtype == torch.int32
    B, ld = logits.shape[0], logits.stride(0)
    for t in (dlogits, dlogits_bf16):
        if t is not None:
            assert t.stride(0) == ld, "dlogits must share the logits leading dimension"
    native.check(native.lib().dmlp_softmax_ce(
        logits.data_ptr(), ld, labels.data_ptr(), B, num_classes, float(scale), _ptr(dlogits),
        _ptr(dlogits_bf16), _ptr(loss_sum), _ptr(correct), _ptr(colsum), _stream()))


def sgd(w: torch.Tensor, g: torch.Tensor, lr: torch.Tensor, scale: float = 1.0,
        w_bf16: Optional[torch.Tensor] = None) -> None:
    """w -= lr[0] * scale * g (lr is a device scalar). Reference client.go:254-267."""
    assert w.is_contiguous() and g.is_contiguous() and w.numel() == g.numel()
    native.check(native.lib().dmlp_sgd(w.data_ptr(), g.data_ptr(), w.numel(), lr.data_ptr(),
                                       float(scale), _ptr(w_bf16), _stream()))


def fill_u32(t: torch.Tensor, value: int) -> None:
    assert t.is_contiguous() and (t.numel() * t.element_size()) % 4 == 0
    native.check(native.lib().dmlp_fill_u32(t.data_ptr(), value, t.numel() * t.element_size() // 4, _stream()))


def cast_bf16(src: torch.Tensor, dst: torch.Tensor) -> None:
    assert src.is_contiguous() and dst.is_contiguous() and src.numel() == dst.numel()
    native.check(native.lib().dmlp_cast_f32_to_bf16(src.data_ptr(), dst.data_ptr(), src.numel(), _stream()))
