"""All-reduce latency / bandwidth sweep: our peer-memory kernels vs torch.distributed NCCL.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29520 scripts/bench_allreduce.py [--max-bytes 1073741824] [--out gpurun_out/allreduce_N.jsonl]

One rank per GPU. For every size (1 KiB .. max, x4 steps) and every implementation the rank launches
`iters` all-reduces back to back between two CUDA events on its stream, after warm-up and a barrier;
the reported time is the MAX over ranks divided by iters. fp32 SUM, out-of-place.
busbw = algbw * 2(n-1)/n (NCCL-tests convention). BASELINE.json config 4.
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--min-bytes", type=int, default=1024)
    ap.add_argument("--max-bytes", type=int, default=1 << 30)
    ap.add_argument("--factor", type=int, default=4)
    ap.add_argument("--out", default="")
    ap.add_argument("--algos", default="auto,oneshot,twoshot,ring,nvls,nccl")
    ap.add_argument("--graph", type=int, default=1, help="time CUDA-graph replays of `iters` launches (default) or eager launches")
    a = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from dmlp_b200.runtime import native
    from dmlp_b200.runtime.comm import Communicator

    maxb = a.max_bytes
    comm = Communicator.from_process_group(local, staging_bytes=32 << 20, arena_bytes=2 * maxb + 4096)
    off_in, off_out = comm.arena.alloc(maxb), comm.arena.alloc(maxb)
    ins = [comm.arena_ptr(0, r, off_in) for r in range(world)]
    outs = [comm.arena_ptr(0, r, off_out) for r in range(world)]
    from dmlp_b200.parallel.data_parallel import arena_tensor
    my_in = arena_tensor(ins[rank], maxb // 4, local)
    my_in.normal_()
    nccl_buf = torch.empty(maxb // 4, device="cuda")
    nvls = None
    if "nvls" in a.algos.split(",") or "auto" in a.algos.split(","):
        from dmlp_b200.runtime.nvls import NvlsArena, NvlsUnavailable
        try:
            nvls = NvlsArena.from_process_group(local, 2 * maxb + 4096)
            nv_in, nv_out = nvls.alloc.alloc(maxb), nvls.alloc.alloc(maxb)
        except NvlsUnavailable as e:
            if rank == 0:
                print(json.dumps({"nvls": "unavailable", "why": str(e)}), flush=True)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    rows = []
    size = a.min_bytes
    sizes = []
    while size <= maxb:
        sizes.append(size)
        size *= a.factor
    for nbytes in sizes:
        count = nbytes // 4
        iters = 100 if nbytes <= (1 << 20) else 40 if nbytes <= (64 << 20) else 10
        for algo in a.algos.split(","):
            if algo in ("oneshot",) and nbytes > (64 << 20):
                continue  # (n-1) x bytes pulled per rank: pointless at these sizes
            if algo == "nvls" and (nvls is None or count % 4):
                continue

            use = algo
            if algo == "auto":  # the dispatcher's choice, NVLS included when the multicast arena exists
                use = native.ALGO_NAMES.get(comm.pick_algo(nbytes, nvls is not None and count % 4 == 0), "twoshot")

            def launch():
                if algo == "nccl":
                    dist.all_reduce(nccl_buf[:count])
                elif use == "nvls":  # multimem.ld_reduce -> multimem.st through the NVSwitch
                    nvls.allreduce(comm, 0, nv_in, nv_out, count, local, stream=stream.cuda_stream)
                else:
                    comm.allreduce(ins, outs, count, algo=native.ALGO_BY_NAME[use], local_idx=0,
                                   stream=stream.cuda_stream)

            for _ in range(5):
                launch()
            torch.cuda.synchronize()
            graph = None
            if a.graph:
                # `iters` launches captured into one CUDA graph: removes the host launch cost, which
                # otherwise dominates every implementation below ~1 MiB
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph, stream=stream):
                    for _ in range(iters):
                        launch()
                graph.replay()
                torch.cuda.synchronize()
            dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            if graph is not None:
                graph.replay()
            else:
                for _ in range(iters):
                    launch()
            e1.record(stream)
            torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            if algo != "nccl":
                comm.check()
            us = t.item() * 1e3
            algbw = nbytes / (us * 1e-6) / 1e9
            picked = use
            row = {"bytes": nbytes, "impl": algo, "picked": picked, "n_gpus": world, "latency_us": round(us, 2),
                   "algbw_GBs": round(algbw, 2), "busbw_GBs": round(algbw * 2 * (world - 1) / world, 2), "iters": iters,
                   "timing": "cuda-graph replay" if a.graph else "eager launches"}
            rows.append(row)
            if rank == 0:
                print(json.dumps(row), flush=True)
    if rank == 0 and a.out:
        with open(a.out, "w") as f:
            for r in rows:
                f.write(json.dumps(r) + "\n")
    sys.stdout.flush()
    torch.cuda.synchronize()
    os._exit(0)  # graph-captured NCCL collectives can stall process-group teardown; results are out


if __name__ == "__main__":
    main()
