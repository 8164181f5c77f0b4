"""Timeline of the wgrad reduce-scatter pipeline on the wide model (torch.distributed.run, >= 2 GPUs; DMLP_RS_PROF=1):
CTA 0 of the layer's wgrad GEMM (per-tile stamps) against the consumer kernel's CTAs (start / first tile arrived / last
chunk written / end), all on rank 0's globaltimer."""
import ctypes as C
import os
import sys

os.environ["DMLP_RS_PROF"] = "1"
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
    dims, batch = (784, 8192, 8192, 10), 8192
    x, y = synthetic_mnist(batch * 4, seed=3)
    for use_graph in (True, False):
        t = Trainer(TrainerConfig(dims=dims, batch=batch, lr=0.01, precision="bf16", overlap=True, fused=0,
                                  use_graph=use_graph), device=local, process_group=True)
        assert t.pipelined_layers >= 1
        t.fit(x, y, 12)
        torch.cuda.synchronize()
        if rank == 0:
            print(f"---- use_graph={use_graph}")
        report(rank, native)
        t.close()
        dist.barrier()
    dist.destroy_process_group()
    os._exit(0)


def report(rank, native):
    lib = native.lib()
    g = (C.c_ulonglong * (16 * 16))()
    lib.dmlp_gemm_prof_read.argtypes = [C.POINTER(C.c_ulonglong), C.c_int]
    native.check(lib.dmlp_gemm_prof_read(g, 16 * 16))
    r = (C.c_ulonglong * (148 * 4))()
    lib.dmlp_wgrad_rs_prof_read.argtypes = [C.POINTER(C.c_ulonglong), C.c_int]
    native.check(lib.dmlp_wgrad_rs_prof_read(r, 148 * 4))
    if rank == 0:
        g = np.array(g, dtype=np.uint64).reshape(16, 16).astype(np.int64)
        r = np.array(r, dtype=np.uint64).reshape(148, 4).astype(np.int64)
        t0 = g[0, 0]
        names = ["prod:slot", "prod:last_tma", "mma:acc_free", "mma:first_full", "mma:last_commit", "epi:acc_ready",
                 "epi:drained", "epi:done"]
        print("wgrad GEMM, CTA 0, us since its first slot acquisition")
        print("tile " + " ".join(f"{n:>10s}" for n in names))
        for ti in range(16):
            if g[ti, 0] < t0:
                break
            print(f"{ti:4d} " + " ".join(f"{(g[ti, i] - t0) / 1e3:10.2f}" for i in range(8)))
        rel = (r - t0) / 1e3
        for i, n in enumerate(["start", "first tile arrived", "last chunk written", "end"]):
            c = rel[:, i]
            print(f"consumer {n:>20s}: min {c.min():9.2f}  median {np.median(c):9.2f}  max {c.max():9.2f} us")


if __name__ == "__main__":
    main()
