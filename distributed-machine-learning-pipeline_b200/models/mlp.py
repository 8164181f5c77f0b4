"""``MLP`` — the reference client's model as a PyTorch module over the native kernels.

Reference: ``DSML/client/client.go:36-58`` (two linear layers 784-128-10, U(-0.05,0.05) init) and the
README's 784-128-64-10 shape (``README.md:134-146``). Weights are ``[in][out]`` row-major like the
reference; all parameters are views into ONE flat fp32 buffer in ``ParamLayout`` order, so the
data-parallel wrapper can all-reduce / update them as a single vector and ``save_weights`` /
``load_weights`` produce the reference client's blob.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence

import torch
from torch import nn

from ..ops import functional as F
from .layout import ParamLayout


class MLP(nn.Module):
    def __init__(self, dims: Sequence[int] = (784, 128, 64, 10), device: str = "cuda", seed: int = 0,
                 flat_storage: Optional[torch.Tensor] = None, grad_storage: Optional[torch.Tensor] = None):
        super().__init__()
        self.dims = tuple(dims)
        self.layout = ParamLayout(dims)
        self.num_classes = self.dims[-1]
        flat = flat_storage if flat_storage is not None else torch.zeros(self.layout.total, device=device)
        assert flat.numel() == self.layout.total and flat.dtype == torch.float32
        self.flat = flat
        self.flat_grad = grad_storage if grad_storage is not None else torch.zeros_like(flat)
        self.weights, self.biases = nn.ParameterList(), nn.ParameterList()
        self._gw: List[torch.Tensor] = []
        self._gb: List[torch.Tensor] = []
        for s in self.layout.layers:
            w = flat[s.w_off:s.w_off + s.fan_in * s.out_pad].view(s.fan_in, s.out_pad)
            b = flat[s.b_off:s.b_off + s.out_pad]
            self.weights.append(nn.Parameter(w))
            self.biases.append(nn.Parameter(b))
            self._gw.append(self.flat_grad[s.w_off:s.w_off + s.fan_in * s.out_pad].view(s.fan_in, s.out_pad))
            self._gb.append(self.flat_grad[s.b_off:s.b_off + s.out_pad])
        self.layer_ready_hook: Optional[Callable[[int], None]] = None
        with torch.no_grad():
            self.flat.copy_(torch.from_numpy(self.layout.init_reference(seed)))

    def _ready(self, l: int):
        hook = self.layer_ready_hook
        return (lambda: hook(l)) if hook is not None else None

    def _chain(self, l: int) -> dict:
        """Layer l > 0 consumes the ReLU output of layer l-1: its dgrad epilogue applies that ReLU's mask and writes
        layer l-1's bias gradient (no eager mask / column-sum kernels in backward)."""
        return dict(mask_input=l > 0, prev_grad_b=self._gb[l - 1] if l > 0 else None)

    def features(self, x: torch.Tensor, head_follows: bool = False) -> torch.Tensor:
        """All layers but the last (ReLU after each). A layer that is consumed by the next layer of this module gets
        its gradient masked and its bias gradient produced there (``pre_masked``); the last of them only when the
        module's own head follows (``forward`` / ``loss``) - a caller that puts its own head on ``features(x)`` gets an
        ordinary autograd tensor."""
        h = x
        n = len(self.weights)
        for l in range(n - 1):
            h = F.fused_linear(h, self.weights[l], self.biases[l], True, self._gw[l], self._gb[l], self._ready(l),
                               pre_masked=(l < n - 2) or head_follows, **self._chain(l))
        return h

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        """Logits ``[B, out_pad]`` (columns >= num_classes are padding and stay zero)."""
        l = len(self.weights) - 1
        return F.fused_linear(self.features(x, head_follows=True), self.weights[l], self.biases[l], False, self._gw[l], self._gb[l],
                              self._ready(l), **self._chain(l))

    def loss(self, x: torch.Tensor, labels: torch.Tensor):
        """Mean cross-entropy and hit count; the last layer and the loss run as one kernel.
        ``loss.backward()`` leaves all gradients in ``flat_grad`` (same layout as ``flat``)."""
        l = len(self.weights) - 1
        return F.linear_softmax_ce(self.features(x, head_follows=True), self.weights[l], self.biases[l], labels.to(torch.int32),
                                   self.num_classes, self._gw[l], self._gb[l], self._ready(l), **self._chain(l))

    # ------------------------------------------------------------------ checkpoints (reference layout)
    def reference_blob(self) -> bytes:
        return self.layout.to_reference_bytes(self.flat.detach().cpu().numpy())

    def load_reference_blob(self, blob: bytes) -> None:
        with torch.no_grad():
            self.flat.copy_(torch.from_numpy(self.layout.from_reference_bytes(blob)))

    def save_weights(self, path: str) -> None:
        with open(path, "wb") as f:
            f.write(self.reference_blob())

    def load_weights(self, path: str) -> None:
        with open(path, "rb") as f:
            self.load_reference_blob(f.read())
