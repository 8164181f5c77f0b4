"""Parameter layout shared by the native trainer, the PyTorch modules and the checkpoints.

In memory (device): one flat fp32 buffer; per layer ``W[in][out_pad]`` then ``b[out_pad]`` with
``out_pad = ceil16(out)`` (padding stays zero) and every tensor starting on a 64-float (256 B)
boundary — what TMA (16-byte row pitch), the tcgen05 N granularity and the 16-byte all-reduce
packs need. This mirrors ``make_layout`` in ``csrc/trainer.cu``.

On the wire / on disk: the reference client's serialisation (``DSML/client/client.go:60-66,
204-218``): ``W1‖b1‖W2‖b2(‖W3‖b3…)``, each matrix ``[in][out]`` row-major *without* padding,
little-endian float32 — the blob the reference uploads to ``MemAddr 0x2000``.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Sequence, Tuple

import numpy as np


def _up(v: int, a: int) -> int:
    return (v + a - 1) // a * a


@dataclass(frozen=True)
class LayerSlot:
    fan_in: int
    fan_out: int
    out_pad: int
    w_off: int  # floats
    b_off: int


class ParamLayout:
    def __init__(self, dims: Sequence[int]):
        self.dims = tuple(int(d) for d in dims)
        cur = 0
        slots: List[LayerSlot] = []
        for fi, fo in zip(self.dims[:-1], self.dims[1:]):
            pad = _up(fo, 16)
            w_off = cur
            cur = _up(cur + fi * pad, 64)
            b_off = cur
            cur = _up(cur + pad, 64)
            slots.append(LayerSlot(fi, fo, pad, w_off, b_off))
        self.layers = tuple(slots)
        self.total = cur

    @property
    def reference_count(self) -> int:
        """Number of floats in the reference serialisation (101 770 for 784-128-10)."""
        return sum(s.fan_in * s.fan_out + s.fan_out for s in self.layers)

    def views(self, flat: np.ndarray) -> List[Tuple[np.ndarray, np.ndarray]]:
        """(W[in, out], b[out]) views (unpadded columns) into a flat padded buffer."""
        out = []
        for s in self.layers:
            w = flat[s.w_off:s.w_off + s.fan_in * s.out_pad].reshape(s.fan_in, s.out_pad)[:, :s.fan_out]
            b = flat[s.b_off:s.b_off + s.fan_out]
            out.append((w, b))
        return out

    def init_reference(self, seed: int = 0) -> np.ndarray:
        """U(-0.05, 0.05) for every weight and bias like ``NewMLP`` (client.go:43-58); padding zero."""
        rng = np.random.default_rng(seed)
        flat = np.zeros(self.total, dtype=np.float32)
        for w, b in self.views(flat):
            w[...] = (rng.random(w.shape, dtype=np.float32) - 0.5) * 0.1
            b[...] = (rng.random(b.shape, dtype=np.float32) - 0.5) * 0.1
        return flat

    def to_reference_bytes(self, flat: np.ndarray) -> bytes:
        parts = []
        for w, b in self.views(np.asarray(flat, dtype=np.float32)):
            parts.append(np.ascontiguousarray(w, dtype="<f4").tobytes())
            parts.append(np.ascontiguousarray(b, dtype="<f4").tobytes())
        return b"".join(parts)

    def from_reference_bytes(self, blob: bytes) -> np.ndarray:
        vals = np.frombuffer(blob, dtype="<f4")
        if vals.size != self.reference_count:
            raise ValueError(f"weight blob has {vals.size} floats, model {self.dims} needs {self.reference_count}")
        flat = np.zeros(self.total, dtype=np.float32)
        cur = 0
        for w, b in self.views(flat):
            w[...] = vals[cur:cur + w.size].reshape(w.shape)
            cur += w.size
            b[...] = vals[cur:cur + b.size]
            cur += b.size
        return flat
