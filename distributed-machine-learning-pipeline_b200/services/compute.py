"""``RunForward`` / ``RunBackward``: MLP compute on the device that holds the data.

The reference's generated stubs still carry these two GPUDevice RPCs from an older proto
(``DSML/proto/gpu_sim.pb.go:1225-1425``, ``gpu_sim_grpc.pb.go:28-29,122-140``) but no server implements them
(``gpu_sim_grpc.pb.go:187-192``). They are the natural hook for "compute where the buffers live", so both
backends provide them here with semantics built from the memory addresses the reference client defines
(``DSML/client/client.go:29-32``: gradients 0x1000, weights 0x2000, data 0x3000, labels 0x4000):

  RunForward(inputAddr, outputAddr)   X = float32 [B][in] at ``inputAddr``, weights = the client's blob
                                      (W1|b1|W2|b2|..., ``client.go:204-218``) at 0x2000
                                      -> logits float32 [B][classes] written at ``outputAddr``
  RunBackward(gradientAddr)           labels at 0x4000 (one-hot float32 [B][classes] like the reference loader,
                                      ``client.go:315-350``, or int32 [B]) for the batch of the last RunForward
                                      -> mean softmax-CE gradients (``client.go:143-202``) as a blob in the weight
                                      layout at ``gradientAddr`` — ready for AllReduceRing

The model shape is inferred from the size of the weight blob (the reference's 784-128-10, the README's
784-128-64-10) or set explicitly with ``DeviceBackend.model_dims``. ``sim`` computes with the NumPy oracle,
``cuda`` with the tcgen05 layer of ``models.MLP`` (same kernels as training).
"""
from __future__ import annotations

from typing import Optional, Sequence, Tuple

import numpy as np

from ..models.layout import ParamLayout

WEIGHT_ADDR = 0x2000
LABEL_ADDR = 0x4000
KNOWN_DIMS = ((784, 128, 10), (784, 128, 64, 10))


def infer_dims(blob_bytes: int, hint: Optional[Sequence[int]] = None) -> Tuple[int, ...]:
    cands = [tuple(hint)] if hint else list(KNOWN_DIMS)
    for dims in cands:
        if ParamLayout(dims).reference_count * 4 == blob_bytes:
            return tuple(dims)
    raise ValueError(f"weight blob of {blob_bytes} bytes matches none of the model shapes {cands}; "
                     f"set model_dims on the backend")


def batch_of(x_bytes: int, in_features: int) -> int:
    if x_bytes == 0 or x_bytes % (4 * in_features):
        raise ValueError(f"input buffer of {x_bytes} bytes is not a whole number of float32[{in_features}] rows")
    return x_bytes // (4 * in_features)


def labels_from_bytes(raw: bytes, batch: int, classes: int) -> np.ndarray:
    """int32 [B] class indices from either the reference's one-hot float32 [B][classes] or int32 [B]."""
    if len(raw) == batch * classes * 4:
        onehot = np.frombuffer(raw, dtype="<f4").reshape(batch, classes)
        return onehot.argmax(1).astype(np.int32)
    if len(raw) == batch * 4:
        y = np.frombuffer(raw, dtype="<i4").astype(np.int32)
        if y.min(initial=0) < 0 or y.max(initial=0) >= classes:
            raise ValueError("label index out of range")
        return y
    raise ValueError(f"label buffer of {len(raw)} bytes is neither one-hot float32[{batch}][{classes}] nor int32[{batch}]")


class OracleCompute:
    """NumPy implementation (fp64 math, fp32 results): the ``sim`` backend and the reference for the GPU test."""

    def __init__(self, backend):
        self.backend = backend
        self.last_input: Optional[int] = None

    def _params(self):
        from ..models import oracle  # noqa: F401  (import check)
        blob = self.backend.read(WEIGHT_ADDR)
        dims = infer_dims(len(blob), getattr(self.backend, "model_dims", None))
        lay = ParamLayout(dims)
        params = [(w.astype(np.float64), b.astype(np.float64)) for w, b in lay.views(lay.from_reference_bytes(blob))]
        return dims, lay, params

    def run_forward(self, input_addr: int, output_addr: int) -> None:
        from ..models import oracle
        dims, _, params = self._params()
        raw = self.backend.read(input_addr)
        x = np.frombuffer(raw, dtype="<f4").reshape(batch_of(len(raw), dims[0]), dims[0]).astype(np.float64)
        logits, _ = oracle.forward(params, x)
        self.backend.check_addr(output_addr)
        self.backend.write(output_addr, np.ascontiguousarray(logits, dtype="<f4").tobytes())
        self.last_input = input_addr

    def run_backward(self, gradient_addr: int) -> None:
        from ..models import oracle
        if self.last_input is None:
            raise ValueError("RunBackward needs a preceding RunForward")
        dims, lay, params = self._params()
        raw = self.backend.read(self.last_input)
        B = batch_of(len(raw), dims[0])
        x = np.frombuffer(raw, dtype="<f4").reshape(B, dims[0]).astype(np.float64)
        y = labels_from_bytes(self.backend.read(LABEL_ADDR), B, dims[-1])
        _, _, grads = oracle.loss_and_grads(params, x, y)
        flat = np.zeros(lay.total, dtype=np.float32)
        for (w, b), (gw, gb) in zip(lay.views(flat), grads):
            w[...] = gw
            b[...] = gb
        self.backend.check_addr(gradient_addr)
        self.backend.write(gradient_addr, lay.to_reference_bytes(flat))
