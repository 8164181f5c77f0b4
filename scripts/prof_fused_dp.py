"""Phase stamps of the fused step kernel in data-parallel mode (in-kernel gradient exchange), under torchrun."""
import json
import os
import sys

os.environ["DMLP_FUSED_PROF"] = "1"
import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dmlp_b200.parallel.trainer import Trainer, TrainerConfig  # noqa: E402
from dmlp_b200.utils.data import synthetic_mnist  # noqa: E402

NAMES_GX = {15: "stage loop entered", 0: "peers' dZ1 staged", 1: "small tensors reduced", 11: "wgrad1 (all ranks) done",
            12: "dW1 applied in place", 14: "fence + end-of-step barrier"}
NAMES = {13: "step start", 2: "F1 mma+scatter", 3: "RS1 arrived", 4: "H1+F2+scatter", 5: "RS2 arrived", 6: "H2+F3+allreduce",
         7: "logits arrived", 8: "CE+dgL+gather dZ2", 9: "dgrad2 done", 10: "gather dZ1", 11: "wgrads drained + wg1 done",
         12: "gradients pushed", 15: "peer flags arrived", 0: "push drained, loads issued", 1: "layer-1 slice reduced + stored", 14: "small ranges + end-of-step barrier"}
ORDER = [13, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 15, 0, 1, 14]


def main():
    multi = "RANK" in os.environ
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if multi:
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    rank = dist.get_rank() if multi else 0
    t = Trainer(TrainerConfig(dims=(784, 128, 64, 10), batch=64, lr=0.01, fused=(int(os.environ.get('DMLP_PROF_FUSED_MODE', '3')) if multi else -1)),
                device=local, process_group=multi)
    x, y = synthetic_mnist(64 * 64, seed=rank)
    acc = np.zeros((8, 16))
    n = 0
    for it in range(20):
        t.fit(x, y, 16)
        rawfull = t.debug_buffer(9, 0, 8 * 64 * 2).view(np.uint64).reshape(8, 64)
        raw = rawfull[:, :16].astype(np.float64)
        if it == 19 and multi and int(os.environ.get('DMLP_PROF_FUSED_MODE', '3')) == 5:
            base = rawfull[:, 10:11].astype(np.int64)  # 'gather dZ1' stamp = dz1 known
            fine = {}
            for name, sl in (("t32", list(range(16, 25))), ("t128", list(range(32, 45)))):
                fine[name] = [[int(v) for v in (rawfull[c, sl].astype(np.int64) - rawfull[c, 9].astype(np.int64))] for c in (0, 3, 7)]
            fine["spins"] = rawfull[:, 48:51].astype(np.int64).tolist()
            fine["coarse_vs_dgrad2"] = {k: int(rawfull[0, k].astype(np.int64) - rawfull[0, 9].astype(np.int64)) for k in (10, 15, 0, 1, 11, 12, 14)}
            print("FINE", rank, json.dumps(fine), flush=True)
        if it >= 18 and int(os.environ.get('DMLP_PROF_FUSED_MODE', '3')) == 5 and multi:
            pass
        if it >= 5:
            acc += raw - raw[:, 13:14].min()
            n += 1
    acc /= n
    mean = acc.mean(axis=0)
    mode = int(os.environ.get('DMLP_PROF_FUSED_MODE', '3'))
    order = [i for i in ORDER if ((multi and mode == 3) or i not in (15, 0, 1))]
    if multi and mode == 5:
        NAMES.update(NAMES_GX)
        order = [13, 2, 3, 4, 5, 6, 7, 8, 9, 10, 15, 0, 1, 11, 12, 14]
    out = {"rank": rank, "world": dist.get_world_size() if multi else 1,
           "phase_end_us": {NAMES[i]: round(mean[i] / 1e3, 2) for i in order},
           "phase_us": {NAMES[order[k]]: round((mean[order[k]] - mean[order[k - 1]]) / 1e3, 2) for k in range(1, len(order))}}
    for r in range(out['world']):
        if multi:
            dist.barrier()
        if rank == r:
            print(json.dumps({'rank': rank, 'phase_us': out['phase_us']}), flush=True)
    t.close()
    if multi:
        dist.barrier()
        dist.destroy_process_group()
    os._exit(0)


if __name__ == "__main__":
    main()
