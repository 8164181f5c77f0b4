"""Autograd functions over the native kernels — the PyTorch framework layer.

``FusedLinear``    y = act(x @ W + b) with the tcgen05 GEMM (+bias+ReLU epilogue); backward runs the
                   dgrad / wgrad GEMMs. Gradients of W and b are written straight into caller-owned
                   flat gradient buffers (no autograd accumulation pass) and a callback announces
                   "layer l is ready" so a data-parallel wrapper can start that layer's all-reduce
                   while backward continues.
``LinearSoftmaxCE`` last layer + softmax cross-entropy in one kernel (logits never leave the SM).

Reference math: ``DSML/client/client.go:112-202``.
"""
from __future__ import annotations

from typing import Callable, Optional

import torch

from . import kernels as K


def _bf16(t: torch.Tensor) -> bool:
    return t.dtype == torch.bfloat16


class FusedLinear(torch.autograd.Function):
    """x [B, in] · W [in, out_pad] (+ b) → [B, out_pad]; optional ReLU. fp32 (tf32 math) or bf16 operands.

    Chained layers keep the ReLU backward and the bias gradients inside the GEMM epilogues:
    ``mask_input``  this layer's input IS the ReLU output of the layer before it, so its dgrad kernel applies that
                    layer's ReLU mask (``x > 0``) and sums the masked result's columns into ``prev_grad_b`` - the
                    previous layer's bias gradient (reference client.go:178-189);
    ``pre_masked``  the consumer did exactly that for THIS layer: the incoming gradient is already dZ and the bias
                    gradient is already in ``grad_b``. Without these flags (stand-alone use) both are eager torch ops."""

    @staticmethod
    def forward(ctx, x, w, b, relu: bool, grad_w: Optional[torch.Tensor], grad_b: Optional[torch.Tensor],
                on_ready: Optional[Callable[[], None]], mask_input: bool = False,
                prev_grad_b: Optional[torch.Tensor] = None, pre_masked: bool = False):
        B, N = x.shape[0], w.shape[1]
        out = torch.empty(B, N, device=x.device, dtype=x.dtype)
        if _bf16(x):
            K.linear_forward(x, w, b, relu, out_bf16=out)
        else:
            K.linear_forward(x, w, b, relu, out=out)
        ctx.save_for_backward(x, w, out if (relu and not pre_masked) else None)
        ctx.relu, ctx.grad_w, ctx.grad_b, ctx.on_ready = relu, grad_w, grad_b, on_ready
        ctx.mask_input, ctx.prev_grad_b, ctx.pre_masked = mask_input, prev_grad_b, pre_masked
        ctx.has_bias = b is not None
        return out

    @staticmethod
    def backward(ctx, dy):
        x, w, out = ctx.saved_tensors
        dy = dy.contiguous()
        if ctx.relu and not ctx.pre_masked:  # ReLU backward on this layer's own output (client.go:104-110)
            dy = dy * (out > 0)
        B, N = dy.shape
        gw = ctx.grad_w if ctx.grad_w is not None else torch.empty(w.shape, device=w.device, dtype=torch.float32)
        K.linear_wgrad(x, dy, gw)
        gb = None
        if ctx.has_bias and not ctx.pre_masked:
            gb = dy.float().sum(0)
            if ctx.grad_b is not None:
                ctx.grad_b.copy_(gb)
        dx = None
        if ctx.needs_input_grad[0]:
            dx = _dgrad(dy, w, x, ctx.mask_input, ctx.prev_grad_b)
        if ctx.on_ready is not None:
            ctx.on_ready()
        ret_w = None if ctx.grad_w is not None else gw.to(w.dtype)
        ret_b = None if (ctx.grad_b is not None or not ctx.has_bias or ctx.pre_masked) else gb
        return dx, ret_w, ret_b, None, None, None, None, None, None, None


def _dgrad(dz: torch.Tensor, w: torch.Tensor, x: torch.Tensor, mask_input: bool,
           prev_grad_b: Optional[torch.Tensor]) -> torch.Tensor:
    """dX = dZ · Wᵀ; with ``mask_input`` the epilogue multiplies by (x > 0) and writes the column sums of the result
    (the previous layer's bias gradient) - no separate elementwise or reduction kernel."""
    B = dz.shape[0]
    dx = torch.empty(B, w.shape[0], device=x.device, dtype=x.dtype)
    mask = x if mask_input else None
    colsum = prev_grad_b if mask_input else None
    if colsum is not None and B > 128:
        colsum.zero_()  # several M tiles accumulate their partial sums
    if _bf16(x):
        K.linear_dgrad(dz, w, mask, out_bf16=dx, colsum=colsum)
    else:
        K.linear_dgrad(dz, w, mask, out=dx, colsum=colsum)
    return dx


class LinearSoftmaxCE(torch.autograd.Function):
    """loss = mean CE(softmax(x @ W + b), labels) with everything after the GEMM fused in its epilogue.
    Returns (loss, correct_count). ``mask_input`` / ``prev_grad_b``: see ``FusedLinear``."""

    @staticmethod
    def forward(ctx, x, w, b, labels, num_classes: int, grad_w, grad_b, on_ready, mask_input: bool = False,
                prev_grad_b: Optional[torch.Tensor] = None):
        B, N = x.shape[0], w.shape[1]
        dl = torch.empty(B, N, device=x.device, dtype=x.dtype)
        loss = torch.zeros(1, device=x.device)
        correct = torch.zeros(1, device=x.device, dtype=torch.int32)
        db = grad_b if grad_b is not None else torch.zeros(N, device=x.device)
        if B > 128:
            db.zero_()
        if _bf16(x):
            K.linear_softmax_ce(x, w, b, labels, num_classes, 1.0 / B, None, loss, correct, dbias=db, dlogits_bf16=dl)
        else:
            K.linear_softmax_ce(x, w, b, labels, num_classes, 1.0 / B, dl, loss, correct, dbias=db)
        ctx.save_for_backward(x, w, dl, db)
        ctx.grad_w, ctx.grad_b, ctx.on_ready = grad_w, grad_b, on_ready
        ctx.mask_input, ctx.prev_grad_b = mask_input, prev_grad_b
        ctx.mark_non_differentiable(correct)
        return loss / B, correct

    @staticmethod
    def backward(ctx, dloss, _dcorrect):
        x, w, dl, db = ctx.saved_tensors
        # dl already holds (p - y)/B and db its column sums; the incoming gradient (ones for loss.backward()) scales
        # both ON THE DEVICE: reading it on the host would stall every backward pass on a synchronisation
        dl.mul_(dloss.to(dl.dtype))
        db.mul_(dloss)
        gw = ctx.grad_w if ctx.grad_w is not None else torch.empty(w.shape, device=w.device, dtype=torch.float32)
        K.linear_wgrad(x, dl, gw)
        dx = None
        if ctx.needs_input_grad[0]:
            dx = _dgrad(dl, w, x, ctx.mask_input, ctx.prev_grad_b)
        if ctx.on_ready is not None:
            ctx.on_ready()
        ret_w = None if ctx.grad_w is not None else gw.to(w.dtype)
        ret_b = None if ctx.grad_b is not None else db
        return dx, ret_w, ret_b, None, None, None, None, None, None, None


def fused_linear(x, w, b=None, relu=False, grad_w=None, grad_b=None, on_ready=None, mask_input=False, prev_grad_b=None,
                 pre_masked=False):
    return FusedLinear.apply(x, w, b, relu, grad_w, grad_b, on_ready, mask_input, prev_grad_b, pre_masked)


def linear_softmax_ce(x, w, b, labels, num_classes, grad_w=None, grad_b=None, on_ready=None, mask_input=False,
                      prev_grad_b=None):
    return LinearSoftmaxCE.apply(x, w, b, labels, num_classes, grad_w, grad_b, on_ready, mask_input, prev_grad_b)
