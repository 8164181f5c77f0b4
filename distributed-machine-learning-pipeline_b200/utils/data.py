"""Datasets: gzip IDX reader (MNIST files) and a synthetic MNIST-shaped generator.

Reference loader: ``DSML/client/client.go:270-313`` (images: magic 2051, big-endian header, pixels
/255 → float32) and ``:315-350`` (labels: magic 2049 → one-hot float32[10]). The reference's
training images are missing from its repo (``.MISSING_LARGE_BLOBS``) and there is no network, so
training runs use ``synthetic_mnist``; the t10k split that does exist loads through ``load_idx_*``.

Labels are class indices (int32); ``one_hot`` gives the reference's float one-hot form.
"""
from __future__ import annotations

import gzip
import os
import struct
from typing import Optional, Tuple

import numpy as np

IMAGE_MAGIC, LABEL_MAGIC = 2051, 2049


def _open(path: str):
    return gzip.open(path, "rb") if path.endswith(".gz") else open(path, "rb")


def load_idx_images(path: str) -> np.ndarray:
    """→ float32 [N, rows*cols] in [0, 1]."""
    with _open(path) as f:
        head = f.read(16)
        magic = struct.unpack(">I", head[:4])[0] if len(head) >= 4 else -1
        if magic != IMAGE_MAGIC or len(head) < 16:
            raise ValueError(f"{path}: invalid magic number {magic} for an image file (expected {IMAGE_MAGIC})")
        _, n, rows, cols = struct.unpack(">IIII", head)
        raw = f.read(n * rows * cols)
    if len(raw) != n * rows * cols:
        raise ValueError(f"{path}: truncated image payload")
    return (np.frombuffer(raw, dtype=np.uint8).astype(np.float32) / 255.0).reshape(n, rows * cols)


def load_idx_labels(path: str) -> np.ndarray:
    """→ int32 [N] class indices."""
    with _open(path) as f:
        head = f.read(8)
        magic, n = struct.unpack(">II", head) if len(head) == 8 else (-1, 0)
        if magic != LABEL_MAGIC:
            raise ValueError(f"{path}: invalid magic number {magic} for a label file (expected {LABEL_MAGIC})")
        raw = f.read(n)
    if len(raw) != n:
        raise ValueError(f"{path}: truncated label payload")
    return np.frombuffer(raw, dtype=np.uint8).astype(np.int32)


def one_hot(labels: np.ndarray, num_classes: int = 10) -> np.ndarray:
    out = np.zeros((labels.shape[0], num_classes), dtype=np.float32)
    out[np.arange(labels.shape[0]), labels] = 1.0
    return out


def write_idx_images(path: str, images_u8: np.ndarray) -> None:
    n, rows, cols = images_u8.shape
    with gzip.open(path, "wb") if path.endswith(".gz") else open(path, "wb") as f:
        f.write(struct.pack(">IIII", IMAGE_MAGIC, n, rows, cols))
        f.write(images_u8.astype(np.uint8).tobytes())


def write_idx_labels(path: str, labels: np.ndarray) -> None:
    with gzip.open(path, "wb") if path.endswith(".gz") else open(path, "wb") as f:
        f.write(struct.pack(">II", LABEL_MAGIC, labels.shape[0]))
        f.write(labels.astype(np.uint8).tobytes())


def synthetic_mnist(n: int, seed: int = 0, num_classes: int = 10, features: int = 784,
                    noise: float = 0.35) -> Tuple[np.ndarray, np.ndarray]:
    """Learnable MNIST-shaped data: each class has a fixed random "stroke" template in [0,1];
    samples are the template under pixel noise, clipped to [0,1]. float32 [n, features], int32 [n]."""
    rng = np.random.default_rng(seed)
    side = int(round(features ** 0.5))
    templates = np.zeros((num_classes, features), dtype=np.float32)
    trng = np.random.default_rng(12345)  # templates do not depend on the sample seed
    for c in range(num_classes):
        t = np.zeros((side, side), dtype=np.float32) if side * side == features else None
        if t is not None:
            for _ in range(4):  # a few random bars per class
                r0, c0 = trng.integers(2, side - 8, 2)
                if trng.random() < 0.5:
                    t[r0:r0 + 2, c0:c0 + int(trng.integers(5, 12))] = 1.0
                else:
                    t[r0:r0 + int(trng.integers(5, 12)), c0:c0 + 2] = 1.0
            templates[c] = t.reshape(-1)
        else:
            templates[c] = (trng.random(features) < 0.15).astype(np.float32)
    labels = rng.integers(0, num_classes, n).astype(np.int32)
    x = templates[labels] + noise * rng.standard_normal((n, features)).astype(np.float32)
    np.clip(x, 0.0, 1.0, out=x)
    return x.astype(np.float32), labels


def find_reference_data(root: str = "/root/reference/DSML/data") -> Optional[dict]:
    """Paths of the MNIST files shipped with the reference, if mounted (t10k only is complete)."""
    p = {k: os.path.join(root, v) for k, v in {
        "test_images": "t10k-images-idx3-ubyte.gz", "test_labels": "t10k-labels-idx1-ubyte.gz",
        "train_images": "train-images-idx3-ubyte.gz", "train_labels": "train-labels-idx1-ubyte.gz"}.items()}
    return {k: v for k, v in p.items() if os.path.exists(v)} or None
