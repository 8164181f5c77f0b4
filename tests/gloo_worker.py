"""CPU multi-process worker (gloo): host-side plumbing that needs more than one process but no GPU.

  1. runtime.fdpass.broadcast_fd: rank 0's file descriptor reaches every rank as SCM_RIGHTS (the NVLS multicast
     object travels this way); an error on rank 0 raises on every rank.
  2. data-parallel semantics: every rank follows parallel.trainer.shard_schedule through a shared dataset, computes
     oracle gradients of its shard, averages them with an all-reduce and applies SGD -> identical replicas that match
     models.oracle.dp_train_step (the reference of the GPU trainer's multi-GPU tests).
Launched by tests/test_cpu_multiprocess.py through torch.distributed.run.
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    from dmlp_b200.models import oracle
    from dmlp_b200.models.layout import ParamLayout
    from dmlp_b200.parallel.trainer import shard_schedule
    from dmlp_b200.runtime.fdpass import broadcast_fd
    from dmlp_b200.utils.data import synthetic_mnist

    # ---- 1. descriptor passing
    path = sys.argv[1]
    fd = os.open(path, os.O_RDONLY) if rank == 0 else None
    got = broadcast_fd(fd, tag="dmlp_test")
    assert os.pread(got, 64, 0) == b"multicast handle stand-in", os.pread(got, 64, 0)
    os.close(got)
    try:
        broadcast_fd(None, tag="dmlp_test", error="rank 0 could not create the object" if rank == 0 else None)
        raise AssertionError("expected RuntimeError on every rank")
    except RuntimeError as e:
        assert "could not create" in str(e)

    # ---- 2. data-parallel schedule + averaged SGD == oracle
    dims, batch, lr, steps = (784, 128, 64, 10), 16, 0.1, 3
    lay = ParamLayout(dims)
    params = [(w.astype(np.float64), b.astype(np.float64)) for w, b in lay.views(lay.init_reference(7))]
    ref = [(w.copy(), b.copy()) for w, b in params]
    x, y = synthetic_mnist(batch * world * steps, seed=3)
    gstep = 0
    for s in range(steps):
        first, stride = shard_schedule(gstep, rank, world)
        assert (first, stride) == (gstep * world + rank, world)
        xs, ys = x[first * batch:(first + 1) * batch].astype(np.float64), y[first * batch:(first + 1) * batch]
        _, _, grads = oracle.loss_and_grads(params, xs, ys)
        new = []
        for (w, b), (gw, gb) in zip(params, grads):
            tw, tb = torch.from_numpy(np.ascontiguousarray(gw)), torch.from_numpy(np.ascontiguousarray(gb))
            dist.all_reduce(tw)
            dist.all_reduce(tb)
            new.append((w - lr * tw.numpy() / world, b - lr * tb.numpy() / world))
        params = new
        shards = [(x[(gstep * world + r) * batch:(gstep * world + r + 1) * batch].astype(np.float64),
                   y[(gstep * world + r) * batch:(gstep * world + r + 1) * batch]) for r in range(world)]
        ref, _, _ = oracle.dp_train_step(ref, shards, lr)
        gstep += 1
    for (w, b), (rw, rb) in zip(params, ref):
        assert np.abs(w - rw).max() < 1e-12 and np.abs(b - rb).max() < 1e-12
    flat = torch.from_numpy(np.concatenate([np.concatenate([w.ravel(), b.ravel()]) for w, b in params]))
    gathered = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(gathered, flat)
    assert all(torch.equal(g, gathered[0]) for g in gathered), "replicas diverged"
    # a rank-private dataset is walked densely
    assert shard_schedule(5, rank, world, batch_stride=1) == (5, 1)
    dist.barrier()
    if rank == 0:
        print("GLOO_WORKER_OK")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
