"""Competitive baseline (NOT the product path): the same MLP trained with stock PyTorch —
nn.Linear/cuBLAS, torch.nn.functional.cross_entropy, torch.optim.SGD and DistributedDataParallel
over NCCL — eager and CUDA-graph captured. Same data shape, batch 64 per GPU, fp32 with TF32 matmul
allowed (what our tensor-core path uses). Device-timed with CUDA events, max over ranks.

    python baseline/nccl_ddp_mlp.py --steps 2000                         # 1 GPU
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port 29530 \
        baseline/nccl_ddp_mlp.py --steps 2000
"""
import argparse
import json
import os

import torch
import torch.distributed as dist
from torch import nn


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=200)
    ap.add_argument("--dims", default="784,128,64,10")
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--bf16", action="store_true")
    a = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.backends.cuda.matmul.allow_tf32 = True
    torch.backends.cudnn.allow_tf32 = True
    dims = [int(d) for d in a.dims.split(",")]
    layers = []
    for i, (fi, fo) in enumerate(zip(dims[:-1], dims[1:])):
        layers.append(nn.Linear(fi, fo))
        if i < len(dims) - 2:
            layers.append(nn.ReLU())
    model = nn.Sequential(*layers).cuda()
    ddp = nn.parallel.DistributedDataParallel(model, device_ids=[local]) if world > 1 else model
    opt = torch.optim.SGD(ddp.parameters(), lr=0.01)
    nb = 1024 if a.batch <= 256 else 8
    xs = torch.rand(nb, a.batch, dims[0], device="cuda")
    ys = torch.randint(0, dims[-1], (nb, a.batch), device="cuda")
    results = {}

    def step(i):
        opt.zero_grad(set_to_none=True)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=a.bf16):
            loss = nn.functional.cross_entropy(ddp(xs[i % nb]), ys[i % nb])
        loss.backward()
        opt.step()
        return loss

    def timed(fn, steps):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            fn(i)
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device="cuda", dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.item()

    for i in range(max(a.warmup, 11)):
        step(i)
    ms = timed(step, a.steps)
    results["eager"] = {"ms_per_step": ms / a.steps, "samples_per_s": a.steps * a.batch * world / (ms / 1e3)}

    # CUDA-graph captured step (static input buffers, whole fwd+bwd+allreduce+optimizer in the graph)
    try:
        sx, sy = xs[0].clone(), ys[0].clone()
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(11):
                opt.zero_grad(set_to_none=True)
                nn.functional.cross_entropy(ddp(sx), sy).backward()
                opt.step()
        torch.cuda.current_stream().wait_stream(s)
        g = torch.cuda.CUDAGraph()
        opt.zero_grad(set_to_none=True)
        with torch.cuda.graph(g):
            nn.functional.cross_entropy(ddp(sx), sy).backward()
            opt.step()

        def gstep(i):
            sx.copy_(xs[i % nb])
            sy.copy_(ys[i % nb])
            g.replay()

        for i in range(20):
            gstep(i)
        ms = timed(gstep, a.steps)
        results["cuda_graph"] = {"ms_per_step": ms / a.steps, "samples_per_s": a.steps * a.batch * world / (ms / 1e3)}
    except Exception as e:  # graph capture of DDP+NCCL can be refused by some stacks
        results["cuda_graph"] = {"error": str(e)[:200]}
    if rank == 0:
        print(json.dumps({"impl": "torch DDP + NCCL + cuBLAS (baseline)", "n_gpus": world, "dims": dims,
                          "per_gpu_batch": a.batch, "bf16_autocast": a.bf16, **results}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
