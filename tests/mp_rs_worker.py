"""Worker for the wgrad GEMM -> reduce-scatter -> SGD -> all-gather pipeline (csrc/wgrad_rs.cuh; torch.distributed.run, one
rank per GPU).

A model with one layer big enough for the pipeline and one that stays on the bucket all-reduce is trained twice from the
same seed: with the pipeline and with ``DMLP_WGRAD_RS=0`` (GEMM, then the bucket all-reduce - the unfused form of the same
math). Parameters must agree to rounding, replicas must be bit-identical, the loss
must fall; both precisions, graph and eager launches. Then a rank dies mid-run: the survivors' consumer kernels time
out waiting for its tiles and name it. Prints 'MP_RS_OK' on rank 0.
"""
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from dmlp_b200.parallel.trainer import Trainer, TrainerConfig
    from dmlp_b200.runtime import native
    from dmlp_b200.utils.data import synthetic_mnist

    dims, batch, steps = (512, 2048, 4096, 10), 512, 12
    x, y = synthetic_mnist(batch * world * steps, seed=3, features=dims[0])

    def replicas_equal(t):
        flat = torch.from_numpy(t.get_params()).cuda()
        gathered = [torch.empty_like(flat) for _ in range(world)]
        dist.all_gather(gathered, flat)
        return all(torch.equal(g, gathered[0]) for g in gathered)

    def run(precision, use_graph, pipelined):
        os.environ["DMLP_WGRAD_RS"] = "1" if pipelined else "0"
        cfg = TrainerConfig(dims=dims, batch=batch, lr=0.05, seed=5, precision=precision, overlap=True, fused=0,
                            use_graph=use_graph)
        t = Trainer(cfg, device=local, process_group=True)
        assert t.pipelined_layers == (1 if pipelined and 128 % world == 0 else 0), t.pipelined_layers
        st = t.fit(x, y, 5)
        st2 = t.fit(x, y, steps - 5)  # a second call: counters / epochs carry over
        losses = np.concatenate([st["per_step_loss"], st2["per_step_loss"]])
        assert np.isfinite(losses).all(), losses
        assert replicas_equal(t), "replicas diverged"
        flat = t.get_params()
        t.close()
        dist.barrier()
        return flat, losses

    for precision in ("bf16", "tf32"):
        for use_graph in (True, False):
            a, la = run(precision, use_graph, True)
            b, lb = run(precision, use_graph, False)
            diff = float(np.abs(a - b).max())
            moved = float(np.abs(a).max())
            assert diff <= 1e-5 * max(1.0, moved), (precision, use_graph, diff)  # (sum order / fma contraction)
            assert np.allclose(la, lb, rtol=1e-4, atol=1e-5), (la, lb)
            assert la[-3:].mean() < la[:3].mean(), la
            if rank == 0:
                print(f"pipeline ok: {precision} graph={use_graph} max |dW| vs bucket all-reduce {diff:.1e}, loss "
                      f"{la[0]:.3f} -> {la[-1]:.3f}", flush=True)

    # a rank dies mid-run
    os.environ["DMLP_WGRAD_RS"] = "1"
    cfg = TrainerConfig(dims=dims, batch=batch, lr=0.05, seed=5, precision="bf16", overlap=True, fused=0, watchdog_ms=400)
    t = Trainer(cfg, device=local, process_group=True)
    assert t.pipelined_layers == (1 if 128 % world == 0 else 0)
    t.fit(x, y, 4)
    dead = world - 1
    if rank == dead:
        t.inject_fault(4 + 3)
    t0 = time.time()
    err = None
    try:
        t.fit(x, y, 8)
    except native.NativeError as e:
        err = e
    dt = time.time() - t0
    if rank != dead:
        assert err is not None and err.code == native.ERR_TIMEOUT, err
        assert f"waited for rank {dead}" in str(err), str(err)
        assert dt < 0.4 + 1.5, f"survivor returned after {dt:.2f} s"
        assert t.failed
        print(f"failure ok on rank {rank}: {err} after {dt * 1e3:.0f} ms", flush=True)
    t.close()
    dist.barrier()
    os.environ.pop("DMLP_WGRAD_RS", None)
    if rank == 0:
        print("MP_RS_OK", flush=True)
    dist.destroy_process_group()
    os._exit(0)


if __name__ == "__main__":
    main()
