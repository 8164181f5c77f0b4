"""Worker for multi-process GPU tests (launched by torch.distributed.run, one rank per GPU).

Checks, with rank-private data and the NumPy oracle:
  1. peer-memory all-reduce over CUDA-IPC mapped arenas (one-shot / two-shot / ring), SUM and fused SGD
  2. the data-parallel trainer: N ranks x batch 64 must follow the oracle's mean-gradient SGD exactly
     and keep replicas bit-identical
Prints 'MP_WORKER_OK' on rank 0 when everything passed.
"""
import os
import sys

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
    from dmlp_b200.models import oracle
    from dmlp_b200.parallel.trainer import Trainer, TrainerConfig
    from dmlp_b200.runtime import native
    from dmlp_b200.runtime.comm import Communicator
    from dmlp_b200.utils.data import synthetic_mnist

    # ---- 1. collectives over IPC-mapped arenas
    count = 109386
    nbytes = (count * 4 + 255) // 256 * 256
    comm = Communicator.from_process_group(local, staging_bytes=1 << 20, arena_bytes=4 * nbytes)
    off_in, off_out, off_w = comm.arena.alloc(nbytes), comm.arena.alloc(nbytes), comm.arena.alloc(nbytes)
    rng = np.random.default_rng(5)
    all_in = [rng.standard_normal(count).astype(np.float32) for _ in range(world)]
    ref = np.sum(np.stack(all_in).astype(np.float64), axis=0)
    lib = native.lib()
    mine = comm.arena_ptr(0, rank)
    lr = torch.tensor([0.1], device="cuda")
    for algo in ("oneshot", "twoshot", "ring"):
        native.check(lib.dmlp_memcpy_h2d(local, mine + off_in, all_in[rank].ctypes.data, count * 4, None))
        ins = [comm.arena_ptr(0, r, off_in) for r in range(world)]
        outs = [comm.arena_ptr(0, r, off_out) for r in range(world)]
        comm.allreduce(ins, outs, count, algo=native.ALGO_BY_NAME[algo], local_idx=0)
        torch.cuda.synchronize()
        comm.check()
        got = np.empty(count, np.float32)
        native.check(lib.dmlp_memcpy_d2h(local, got.ctypes.data, mine + off_out, count * 4, None))
        assert np.abs(got - ref).max() < 1e-4, (algo, np.abs(got - ref).max())
        # fused average + SGD on a weight replica
        w0 = np.ones(count, np.float32)
        native.check(lib.dmlp_memcpy_h2d(local, mine + off_w, w0.ctypes.data, count * 4, None))
        dist.barrier()
        ws = [comm.arena_ptr(0, r, off_w) for r in range(world)]
        lrs = [lr.data_ptr() if r == rank else None for r in range(world)]
        comm.allreduce(ins, ws, count, algo=native.ALGO_BY_NAME[algo], local_idx=0, fuse_sgd=True, lr=lrs,
                       scale=1.0 / world)
        torch.cuda.synchronize()
        comm.check()
        native.check(lib.dmlp_memcpy_d2h(local, got.ctypes.data, mine + off_w, count * 4, None))
        assert np.abs(got - (1.0 - 0.1 * ref / world)).max() < 1e-4, algo
        dist.barrier()
    comm.destroy()

    # ---- 1b. NVLS (multicast) all-reduce, when the fabric supports it
    from dmlp_b200.runtime.nvls import NvlsArena, NvlsUnavailable
    nvls_ok = True
    try:
        arena = NvlsArena.from_process_group(local, 6 * nbytes)
    except NvlsUnavailable as e:
        nvls_ok = False
        if rank == 0:
            print("NVLS unavailable:", e)
    if nvls_ok:
        comm2 = Communicator.from_process_group(local, staging_bytes=1 << 20)
        count4 = count // 4 * 4
        o_in, o_out, o_w = arena.alloc.alloc(nbytes), arena.alloc.alloc(nbytes), arena.alloc.alloc(nbytes)
        uc = arena.uc[local]
        native.check(lib.dmlp_memcpy_h2d(local, uc + o_in, all_in[rank].ctypes.data, count * 4, None))
        w0 = np.ones(count, np.float32)
        native.check(lib.dmlp_memcpy_h2d(local, uc + o_w, w0.ctypes.data, count * 4, None))
        dist.barrier()
        arena.allreduce(comm2, 0, o_in, o_out, count, local)
        arena.allreduce(comm2, 0, o_in, o_w, count, local, fuse_sgd=True, lr=lr.data_ptr(), scale=1.0 / world)
        torch.cuda.synchronize()
        comm2.check()
        got = np.empty(count, np.float32)
        native.check(lib.dmlp_memcpy_d2h(local, got.ctypes.data, uc + o_out, count * 4, None))
        assert np.abs(got - ref).max() < 1e-4, ("nvls", np.abs(got - ref).max())
        native.check(lib.dmlp_memcpy_d2h(local, got.ctypes.data, uc + o_w, count * 4, None))
        assert np.abs(got - (1.0 - 0.1 * ref / world)).max() < 1e-4, "nvls sgd"
        # one-shot flavour: every rank pulls the switch-reduced sum of everything into its own replica
        o_in2, o_w2 = arena.alloc.alloc(nbytes), arena.alloc.alloc(nbytes)
        native.check(lib.dmlp_memcpy_h2d(local, uc + o_in2, all_in[rank].ctypes.data, count4 * 4, None))
        native.check(lib.dmlp_memcpy_h2d(local, uc + o_w2, w0.ctypes.data, count4 * 4, None))
        torch.cuda.synchronize()
        dist.barrier()
        arena.allreduce(comm2, 0, o_in2, o_w2, count4, local, fuse_sgd=True, lr=lr.data_ptr(), scale=1.0 / world, oneshot=True)
        torch.cuda.synchronize()
        comm2.check()
        got4 = np.empty(count4, np.float32)
        native.check(lib.dmlp_memcpy_d2h(local, got4.ctypes.data, uc + o_w2, count4 * 4, None))
        assert np.abs(got4 - (1.0 - 0.1 * ref[:count4] / world)).max() < 1e-4, "nvls one-shot sgd"
        rep = torch.from_numpy(got4).cuda()
        reps = [torch.empty_like(rep) for _ in range(world)]
        dist.all_gather(reps, rep)
        assert all(torch.equal(r_, reps[0]) for r_ in reps), "nvls one-shot: replicas differ"
        dist.barrier()
        arena.destroy()
        comm2.destroy()
        if rank == 0:
            print("NVLS_OK")

    # ---- 2. data-parallel trainer vs oracle
    # (fused=3: the step kernel exchanges the gradients itself - push over NVLink, rank-ordered reduce + SGD;
    #  fused=4: persistent launch = compute cluster + reducer CTAs running the one-shot all-reduce+SGD;
    #  fused=5 / -1: global-batch mode, covered in depth by mp_gx_worker.py)
    for overlap, algo, fused in ((False, "auto", -1), (True, "auto", -1), (False, "nvls" if nvls_ok else "twoshot", -1),
                                 (False, "auto", 3), (False, "auto", 4), (False, "auto", 1), (False, "auto", 5)):
        cfg = TrainerConfig(dims=(784, 128, 64, 10), batch=64, lr=0.1, overlap=overlap, seed=7, allreduce=algo, fused=fused)
        t = Trainer(cfg, device=local, process_group=True)
        x, y = synthetic_mnist(64 * world * 6, seed=3)
        params = [(w.astype(np.float64), b.astype(np.float64)) for w, b in t.weights()]
        steps = 6
        st = t.fit(x, y, steps)
        if fused in (3, 4, 5) or (fused == -1 and algo == "auto"):  # (-1 picks the persistent data-parallel launch)
            assert t.kernels_per_step == 1 and t.launches_last_run == 1, (t.kernels_per_step, t.launches_last_run)
        if fused == 1:
            assert t.kernels_per_step == 2, t.kernels_per_step
        for s in range(steps):
            shards = [(x[(s * world + r) * 64:(s * world + r + 1) * 64].astype(np.float64),
                       y[(s * world + r) * 64:(s * world + r + 1) * 64]) for r in range(world)]
            params, _, _ = oracle.dp_train_step(params, shards, 0.1)
        flat = torch.from_numpy(t.get_params()).cuda()
        for (w, b), (rw, rb) in zip(t.weights(), params):
            assert np.abs(w - rw).max() < 3e-3 and np.abs(b - rb).max() < 3e-3, (overlap, np.abs(w - rw).max())
        gathered = [torch.empty_like(flat) for _ in range(world)]
        dist.all_gather(gathered, flat)
        for g in gathered:
            assert torch.equal(g, gathered[0]), "replicas diverged"
        assert np.isfinite(st["loss"])
        t.close()
        dist.barrier()
    if rank == 0:
        print("MP_WORKER_OK")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
