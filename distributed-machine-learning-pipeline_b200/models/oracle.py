"""CPU restatement of the reference's model math, used as the numerical oracle and as the timed CPU
baseline. Same formulas as ``DSML/client/client.go``: forward (:112-141), softmax + loss with the
1e-10 epsilon and dLogits=(p-y)/B (:76-93,144-165), backward (:167-200), ReLU mask on the
activation (:104-110,189), SGD (:254-267), accuracy (:352-382) — vectorised in NumPy float64/32,
and generalised from two layers to any depth.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import numpy as np


def forward(params: Sequence[Tuple[np.ndarray, np.ndarray]], x: np.ndarray):
    """Returns (logits, activations) where activations[l] is the input of layer l."""
    acts = [x]
    h = x
    for i, (w, b) in enumerate(params):
        z = h @ w + b
        h = np.maximum(z, 0) if i < len(params) - 1 else z
        acts.append(h)
    return acts[-1], acts[:-1]


def softmax(z: np.ndarray) -> np.ndarray:
    e = np.exp(z - z.max(axis=1, keepdims=True))
    return e / e.sum(axis=1, keepdims=True)


def loss_and_grads(params, x: np.ndarray, labels: np.ndarray):
    """→ (mean loss, #correct, [(dW, db)] per layer)."""
    B = x.shape[0]
    logits, acts = forward(params, x)
    p = softmax(logits)
    onehot = np.zeros_like(p)
    onehot[np.arange(B), labels] = 1
    loss = float(np.mean(-np.log((p * onehot).sum(1) + 1e-10)))
    correct = int((logits.argmax(1) == labels).sum())
    dz = (p - onehot) / B
    grads: List[Tuple[np.ndarray, np.ndarray]] = [None] * len(params)
    for l in range(len(params) - 1, -1, -1):
        grads[l] = (acts[l].T @ dz, dz.sum(0))
        if l > 0:
            dz = (dz @ params[l][0].T) * (acts[l] > 0)
    return loss, correct, grads


def sgd_step(params, grads, lr: float):
    return [(w - lr * dw, b - lr * db) for (w, b), (dw, db) in zip(params, grads)]


def train_step(params, x, labels, lr: float):
    loss, correct, grads = loss_and_grads(params, x, labels)
    return sgd_step(params, grads, lr), loss, correct


def dp_train_step(params, shards: Sequence[Tuple[np.ndarray, np.ndarray]], lr: float):
    """Data-parallel oracle: every rank computes grads on its shard; the update uses their mean
    (what the fused all-reduce+SGD kernel must produce on every replica)."""
    all_grads, losses, corrects = [], [], 0
    for x, y in shards:
        l, c, g = loss_and_grads(params, x, y)
        all_grads.append(g)
        losses.append(l)
        corrects += c
    n = len(shards)
    mean = [(sum(g[l][0] for g in all_grads) / n, sum(g[l][1] for g in all_grads) / n) for l in range(len(params))]
    return sgd_step(params, mean, lr), float(np.mean(losses)), corrects


def ring_allreduce(buffers: Sequence[np.ndarray], op=np.add) -> List[np.ndarray]:
    """Correct ring all-reduce schedule on host arrays (scatter-reduce then all-gather), including
    sizes not divisible by the rank count — the intended semantics of the reference's
    ``AllReduceRing`` (gpu_coordinator_server.go:272-376; see SURVEY §3.2 for its defects)."""
    n = len(buffers)
    bufs = [np.array(b, copy=True) for b in buffers]
    if n < 2:
        return bufs
    size = bufs[0].shape[0]
    seg = -(-size // n)
    bnd = [(min(s * seg, size), min((s + 1) * seg, size)) for s in range(n)]
    for step in range(n - 1):
        sends = [bufs[r][slice(*bnd[(r - step) % n])].copy() for r in range(n)]
        for r in range(n):
            lo, hi = bnd[(r - 1 - step) % n]
            bufs[r][lo:hi] = op(bufs[r][lo:hi], sends[(r - 1) % n])
    for step in range(n - 1):
        sends = [bufs[r][slice(*bnd[(r + 1 - step) % n])].copy() for r in range(n)]
        for r in range(n):
            lo, hi = bnd[(r - step) % n]
            bufs[r][lo:hi] = sends[(r - 1) % n]
    return bufs


def naive_allreduce(buffers: Sequence[np.ndarray], op=np.add) -> List[np.ndarray]:
    """Gather → reduce → broadcast (gpu_coordinator_server.go:611-717)."""
    acc = np.array(buffers[0], copy=True)
    for b in buffers[1:]:
        acc = op(acc, b)
    return [acc.copy() for _ in buffers]
