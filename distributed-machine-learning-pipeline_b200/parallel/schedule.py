"""Learning-rate schedules. The reference trains with a constant 0.01 (``DSML/client/client.go:27``)
although its README advertises an "adaptive learning rate scheduler" (``README.md:144``); both are
available here. The value is written to a device scalar that the fused all-reduce+SGD kernel reads,
so changing it never re-captures the CUDA graph."""
from __future__ import annotations

import math
from typing import Callable

Schedule = Callable[[int], float]


def constant(lr: float) -> Schedule:
    return lambda step: lr


def step_decay(lr: float, every: int, gamma: float = 0.5) -> Schedule:
    return lambda step: lr * (gamma ** (step // max(every, 1)))


def cosine(lr: float, total_steps: int, min_lr: float = 0.0) -> Schedule:
    return lambda step: min_lr + 0.5 * (lr - min_lr) * (1 + math.cos(math.pi * min(step, total_steps) / max(total_steps, 1)))


def warmup(inner: Schedule, warmup_steps: int) -> Schedule:
    return lambda step: inner(step) * min(1.0, (step + 1) / max(warmup_steps, 1))


def make(name: str, lr: float, total_steps: int = 0, **kw) -> Schedule:
    if name == "constant":
        return constant(lr)
    if name == "step":
        return step_decay(lr, kw.get("every", max(total_steps // 3, 1)), kw.get("gamma", 0.5))
    if name == "cosine":
        return cosine(lr, total_steps, kw.get("min_lr", 0.0))
    if name == "warmup-cosine":
        return warmup(cosine(lr, total_steps, kw.get("min_lr", 0.0)), kw.get("warmup_steps", max(total_steps // 20, 1)))
    raise ValueError(f"unknown schedule {name!r}")


def parse_fault(spec: str):
    """``"dead-rank:<r>@step:<s>"`` -> (r, s); empty / None -> (-1, -1). The fault-injection syntax of ``Config.fault``
    (utils/config.py), ``train.py`` and ``bench.py --fault``."""
    if not spec:
        return -1, -1
    try:
        left, right = spec.split("@")
        kind, r = left.split(":")
        what, s = right.split(":")
        if kind != "dead-rank" or what != "step":
            raise ValueError(spec)
        return int(r), int(s)
    except Exception as e:  # noqa: BLE001
        raise ValueError(f"fault spec {spec!r} is not of the form dead-rank:<r>@step:<s>") from e
