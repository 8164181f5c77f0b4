"""dmlp_b200 — a B200-native (sm_100a) collective runtime and data-parallel MLP trainer.

Capabilities and gRPC contract follow Helenbzbz/Distributed-Machine-Learning-Pipeline (a Go
simulator); the implementation is new: hand-written tcgen05/TMEM/TMA GEMMs,
peer-memory all-reduce kernels fused with the SGD update, a C++/CUDA runtime behind a C ABI
(``include/dmlp_b200.h``), and a PyTorch framework layer on top.

Sub-packages
    proto      the ``gpu_sim`` wire contract, built at runtime (no protoc needed)
    services   GPUDevice / GPUCoordinator gRPC servers (backends: ``sim`` on CPU, ``cuda`` on GPUs)
    runtime    ctypes binding of the native library (devices, communicators, collectives)
    ops        tensor-level ops over the native kernels (+ autograd functions)
    models     MLP (reference layout) and oracle
    parallel   data-parallel wrapper, fused all-reduce+SGD, LR schedules
    utils      data (IDX reader, synthetic MNIST), checkpoints, config, logging, timing
"""
__version__ = "0.1.0"

from . import utils  # noqa: F401  (pure Python, always importable)
