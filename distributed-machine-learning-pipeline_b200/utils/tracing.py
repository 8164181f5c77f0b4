"""NVTX ranges for the Python side of the framework (``DMLP_NVTX=1``).

The reference has no tracing (``log.Printf`` only). Here the native runtime wraps trainer runs, graph
capture and collective launches in NVTX ranges (``csrc/common.h::NvtxRange``); this module does the same for the RPC
handlers and the Python entry points, so a timeline taken with Nsight Systems or ``ncu --nvtx`` shows which RPC or
``fit`` call a kernel belongs to. Without ``DMLP_NVTX=1`` (or without a CUDA build of torch) the ranges cost one
attribute lookup.
"""
from __future__ import annotations

import contextlib
import functools
import os
from typing import Callable, Iterator

_PUSH = _POP = None


def enabled() -> bool:
    return os.environ.get("DMLP_NVTX", "0") not in ("", "0")


def _bind() -> bool:
    global _PUSH, _POP
    if _PUSH is None:
        try:
            import torch
            if not torch.cuda.is_available():
                raise RuntimeError("no CUDA device")
            _PUSH, _POP = torch.cuda.nvtx.range_push, torch.cuda.nvtx.range_pop
        except Exception:  # no torch / no CUDA: ranges are no-ops
            _PUSH = _POP = False
    return bool(_PUSH)


@contextlib.contextmanager
def nvtx_range(name: str) -> Iterator[None]:
    if not (enabled() and _bind()):
        yield
        return
    _PUSH(name)
    try:
        yield
    finally:
        _POP()


def traced(name: str) -> Callable:
    """Decorator form: ``@traced("dmlp.rpc.AllReduceRing")``."""
    def deco(fn):
        @functools.wraps(fn)
        def wrapper(*a, **kw):
            if not enabled():
                return fn(*a, **kw)
            with nvtx_range(name):
                return fn(*a, **kw)
        return wrapper
    return deco
