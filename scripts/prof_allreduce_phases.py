"""Per-phase timing of the one-shot all-reduce kernel from in-kernel globaltimer stamps (torchrun)."""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from dmlp_b200.runtime import native
    from dmlp_b200.runtime.comm import Communicator
    maxb = 4 << 20
    comm = Communicator.from_process_group(local, arena_bytes=2 * maxb + 4096)
    off_in, off_out = comm.arena.alloc(maxb), comm.arena.alloc(maxb)
    ins = [comm.arena_ptr(0, r, off_in) for r in range(world)]
    outs = [comm.arena_ptr(0, r, off_out) for r in range(world)]
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    sizes = [int(v) for v in os.environ.get('PROF_SIZES', '1024,65536,262144,437544,1048576').split(',')]
    for nbytes in sizes:
        count = nbytes // 4
        iters = 50
        prof = torch.zeros(iters, 128, 4, dtype=torch.int64, device="cuda")
        for skip in (False, True):
            for _ in range(5):
                comm.allreduce(ins, outs, count, algo=native.ALGO_ONESHOT, local_idx=0, stream=stream.cuda_stream,
                               skip_end_barrier=skip)
            torch.cuda.synchronize()
            dist.barrier()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=stream):
                for i in range(iters):
                    comm.allreduce(ins, outs, count, algo=native.ALGO_ONESHOT, local_idx=0, stream=stream.cuda_stream,
                                   skip_end_barrier=skip, prof=prof[i].data_ptr())
            g.replay()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            g.replay()
            e1.record(stream)
            torch.cuda.synchronize()
            p = prof.cpu().numpy().astype("float64")
            used = (p[:, :, 0] > 0)
            nb = int(used[0].sum())
            p = p[:, :nb, :]
            entry = p[:, :, 0].min(axis=1)
            row = {"debug": os.environ.get("DMLP_COLL_DEBUG", "0"), "maxblocks": os.environ.get("DMLP_COLL_MAXBLOCKS", "0"), "bytes": nbytes, "skip_end_barrier": skip, "blocks": nb, "n_gpus": world,
                   "us_per_iter_events": round(e0.elapsed_time(e1) * 1e3 / iters, 2),
                   "barrier_in_us": round(float((p[:, :, 1] - p[:, :, 0]).mean()) / 1e3, 2),
                   "data_us": round(float((p[:, :, 2] - p[:, :, 1]).mean()) / 1e3, 2),
                   "barrier_out_us": round(float((p[:, :, 3] - p[:, :, 2]).mean()) / 1e3, 2),
                   "kernel_span_us": round(float((p[:, :, 3].max(axis=1) - entry).mean()) / 1e3, 2),
                   "block_entry_spread_us": round(float((p[:, :, 0].max(axis=1) - entry).mean()) / 1e3, 2),
                   "gap_between_kernels_us": round(float((entry[1:] - p[:-1, :, 3].max(axis=1)).mean()) / 1e3, 2)}
            if rank == 0:
                print(json.dumps(row), flush=True)
            dist.barrier()
    sys.stdout.flush()
    os._exit(0)


if __name__ == "__main__":
    main()
