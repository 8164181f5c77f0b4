"""Checkpoint = the reference client's weight blob + a JSON sidecar with what is needed to resume.

The reference has no disk checkpoint; the nearest thing is the ``W1‖b1‖W2‖b2`` little-endian float32 blob
it pushes to every device after each step (``DSML/client/client.go:204-218,60-66``). A checkpoint here is that
very blob (so any tool that understands the reference layout can read it) and ``<path>.json`` next to it:
step counter (the LR schedule is a pure function of it), epoch, learning-rate schedule, model shape.
"""
from __future__ import annotations

import json
import os
import tempfile
from typing import Any, Dict, Optional, Sequence

FORMAT = "dmlp_b200.checkpoint.v1"


def sidecar_path(path: str) -> str:
    return path + ".json"


def write_checkpoint(path: str, blob: bytes, *, dims: Sequence[int], batch: int, global_step: int, epoch: int,
                     lr: float, schedule: str = "constant", world: int = 1, extra: Optional[Dict[str, Any]] = None) -> None:
    """Atomic: both files are written to temporaries and renamed, the sidecar last (a sidecar always has its blob)."""
    meta = {"format": FORMAT, "dims": list(dims), "batch": int(batch), "global_step": int(global_step),
            "epoch": int(epoch), "lr": float(lr), "schedule": schedule, "world": int(world),
            "blob_bytes": len(blob), "layout": "W1|b1|W2|b2|... little-endian float32, W[in][out] row-major"}
    if extra:
        meta["extra"] = extra
    d = os.path.dirname(os.path.abspath(path)) or "."
    for target, payload, mode in ((path, blob, "wb"), (sidecar_path(path), json.dumps(meta, indent=1).encode(), "wb")):
        fd, tmp = tempfile.mkstemp(dir=d, prefix=".ckpt-")
        with os.fdopen(fd, mode) as f:
            f.write(payload)
        os.replace(tmp, target)


def read_checkpoint(path: str, expect_dims: Optional[Sequence[int]] = None):
    """Returns (blob, meta). A blob without a sidecar is accepted as a plain weight file (meta = {})."""
    with open(path, "rb") as f:
        blob = f.read()
    meta: Dict[str, Any] = {}
    sp = sidecar_path(path)
    if os.path.exists(sp):
        with open(sp) as f:
            meta = json.load(f)
        if meta.get("format") != FORMAT:
            raise ValueError(f"{sp}: unknown checkpoint format {meta.get('format')!r}")
        if meta.get("blob_bytes") not in (None, len(blob)):
            raise ValueError(f"{path}: {len(blob)} bytes, sidecar says {meta['blob_bytes']}")
        if expect_dims is not None and list(expect_dims) != list(meta.get("dims", expect_dims)):
            raise ValueError(f"{path}: checkpoint is for dims {meta['dims']}, model has {list(expect_dims)}")
    return blob, meta
