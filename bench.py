"""Headline benchmark: data-parallel MLP training throughput (samples/s) on N B200s of one node.

    python bench.py --gpus 1 --steps 2000 --warmup 200
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29500 bench.py --gpus 8 --steps 2000 --warmup 200
    python bench.py --impl reference ...        # the unmodified reference (Go) — unavailable here

Metric / config follow BASELINE.json: "MLP samples/s at 1/2/4/8 B200", MLP 784-128-64-10, SGD,
batch 64 per GPU (weak scaling), synthetic 28x28 MNIST-shaped data, random-init weights, the
all-reduce fused into the SGD update. `--model wide` runs config 5 (784-8192-8192-10, global batch
65536 at 8 GPUs, bf16).

Two numbers are measured, each bracketed by barrier + synchronize, timed on the device with CUDA
events on the compute stream, max over ranks:
  value : kernel-side step (full step incl. optimizer + collective); each step's inputs are copied
          from a DEVICE-resident dataset larger than L2 (so inputs never hit in L2)
  e2e   : the same through the public API (`Trainer.fit`) with every step's inputs copied from
          PINNED HOST memory and every step's {loss, correct} read back to the host
Rank 0 prints ONE JSON line.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# the reference's published run (README.md:199-203: 10 epochs x 937 batches x 64 in 732 s on 3 simulated
# devices; batches/epoch from DSML/client/client.go:581) = 819 samples/s
BASELINE_SAMPLES_PER_S = 819.0
MODELS = {
    "small": dict(dims=(784, 128, 64, 10), batch=64, precision="tf32", name="MLP 784-128-64-10"),
    "ref2": dict(dims=(784, 128, 10), batch=64, precision="tf32", name="MLP 784-128-10"),
    "wide": dict(dims=(784, 8192, 8192, 10), batch=8192, precision="bf16", name="MLP 784-8192-8192-10"),
}


class ClockSampler:
    """Samples nvidia-smi SM clocks and throttle reasons while the timed regions run."""

    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu, self.rows, self.proc, self._nvml = gpu_index, [], None, None
        self._t = None

    def start(self):
        # NVML in-process (10 ms period) when available: short runs still get samples under load
        try:
            import pynvml
            pynvml.nvmlInit()
            self._nvml = pynvml
            self._h = pynvml.nvmlDeviceGetHandleByIndex(self.gpu)
            self._stop = False
            self._t = threading.Thread(target=self._pump_nvml, daemon=True)
            self._t.start()
            return
        except Exception:
            self._nvml = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.gpu)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        self._t = threading.Thread(target=self._pump, daemon=True)
        self._t.start()

    def _pump_nvml(self):
        n = self._nvml
        get_reasons = getattr(n, "nvmlDeviceGetCurrentClocksEventReasons", None) or n.nvmlDeviceGetCurrentClocksThrottleReasons
        while not self._stop:
            try:
                sm = n.nvmlDeviceGetClockInfo(self._h, n.NVML_CLOCK_SM)
                mx = n.nvmlDeviceGetMaxClockInfo(self._h, n.NVML_CLOCK_SM)
                bits = get_reasons(self._h)
                act = lambda b: "Active" if bits & b else "Not Active"  # noqa: E731
                # same column layout as the nvidia-smi query
                self.rows.append((time.time(), [str(self.gpu), str(sm), str(mx), "", act(0x8), act(0x40), act(0x20), act(0x4)]))
            except Exception:
                pass
            time.sleep(0.01)

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [c.strip() for c in line.split(",")]))

    def stop(self, t0: float, t1: float) -> dict:
        if getattr(self, "_nvml", None) is not None:
            self._stop = True
            self._t.join(timeout=1.0)
        elif self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        else:
            time.sleep(0.15)
            self.proc.terminate()
        rows = [r for ts, r in self.rows if t0 <= ts <= t1] or [r for ts, r in self.rows if t0 - 0.1 <= ts <= t1 + 0.1] or [r for _, r in self.rows]
        sm, mx, reasons = [], None, set()
        for r in rows:
            try:
                sm.append(float(r[1]))
                mx = float(r[2])
            except (ValueError, IndexError):
                continue
            for name, col in (("hw_slowdown", 4), ("hw_thermal_slowdown", 5), ("sw_thermal_slowdown", 6), ("sw_power_cap", 7)):
                if len(r) > col and r[col].lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def run_reference(args) -> int:
    """The reference is a Go module (no setup.py/pyproject, no Go toolchain on this image, no network
    for its go.sum dependencies), so the one offline install attempt fails; see DESIGN.md."""
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    why = ("reference is a Go gRPC project: `pip install --target baseline/_ref /root/reference` fails "
           "(\"Neither 'setup.py' nor 'pyproject.toml' found\"), and no Go toolchain / vendored go.sum deps exist offline")
    if os.path.isdir(ref_dir) and os.listdir(ref_dir):
        why = "baseline/_ref exists but holds no runnable build of the Go reference (no Go toolchain on this image)"
    if int(os.environ.get("RANK", "0")) == 0:
        print(json.dumps({"impl": "reference", "unavailable": why}))
    return 0


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=200)
    ap.add_argument("--impl", default="ours", choices=("ours", "reference"))
    ap.add_argument("--model", default="small", choices=tuple(MODELS))
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch (default: model's)")
    ap.add_argument("--allreduce", default="auto")
    ap.add_argument("--overlap", type=int, default=-1, help="layer-by-layer all-reduce overlap (default: on for multi-GPU wide)")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--fused", type=int, default=-1, help="TrainerConfig.fused (-1 auto; 2/3/4 select variants of the fused path)")
    ap.add_argument("--host-batches", type=int, default=0, help="host dataset size in batches per rank")
    ap.add_argument("--fault", default="", help="after the timed regions: failure drill 'dead-rank:<r>@step:<s>' (s counted "
                    "from the drill's start); reports how long the survivors need to notice (BASELINE config 5)")
    ap.add_argument("--watchdog-ms", type=int, default=10000)
    a = ap.parse_args()
    if a.impl == "reference":
        return run_reference(a)

    import numpy as np
    import torch
    import torch.distributed as dist
    from dmlp_b200.parallel.trainer import DeviceDataset, Trainer, TrainerConfig, pin
    from dmlp_b200.utils.data import synthetic_mnist

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != a.gpus and rank == 0:
        print(f"warning: --gpus {a.gpus} but WORLD_SIZE={world}; using WORLD_SIZE", file=sys.stderr)
    torch.cuda.set_device(local_rank)
    multi = world > 1
    # Native libraries (NCCL's version banner, for one) write to the C-level stdout. The contract is ONE
    # JSON line on stdout, so fd 1 points at stderr until the result is ready to be printed.
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    if multi:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group(backend="nccl", device_id=torch.device("cuda", local_rank))

    m = MODELS[a.model]
    batch = a.batch or m["batch"]
    overlap = (a.model == "wide" and multi) if a.overlap < 0 else bool(a.overlap)
    cfg = TrainerConfig(dims=m["dims"], batch=batch, lr=0.01, precision=m["precision"], allreduce=a.allreduce,
                        overlap=overlap, use_graph=not a.no_graph, fused=a.fused, watchdog_ms=a.watchdog_ms)
    trainer = Trainer(cfg, device=local_rank, process_group=multi)

    # ---- synthetic data: a pinned host dataset (e2e) and a device dataset larger than L2 (kernel timing)
    feat = m["dims"][0]
    l2_bytes = 126 << 20
    dev_batches = max(8, (l2_bytes * 3 // 2) // (batch * feat * 4) + 1)
    if a.model == "wide":
        dev_batches = max(8, min(dev_batches, 16))
    host_batches = a.host_batches or min(max(dev_batches, 256), 2048)
    gen = np.random.default_rng(1234 + rank)
    base_x, base_y = synthetic_mnist(min(host_batches, 64) * batch, seed=100 + rank, features=feat, num_classes=m["dims"][-1])
    reps = -(-host_batches * batch // base_x.shape[0])
    hx = np.tile(base_x, (reps, 1))[:host_batches * batch]
    hy = np.tile(base_y, reps)[:host_batches * batch]
    hx += (0.01 * gen.standard_normal(hx.shape)).astype(np.float32)  # tiles are not bit-identical
    px, py = pin(hx), pin(hy)
    dset = DeviceDataset(local_rank, hx[:dev_batches * batch], hy[:dev_batches * batch])
    del hx, hy

    def barrier():
        torch.cuda.synchronize()
        if multi:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v: float) -> float:
        if not multi:
            return v
        t = torch.tensor([v], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # Each rank owns a private synthetic dataset and walks it densely (batch_stride=1).
    def timed(source, steps):
        return trainer.fit(source, None if isinstance(source, DeviceDataset) else py, steps,
                           return_metrics=True, device_timing=True, batch_stride=1)

    # ---- warm-up (graph capture, module load, clocks). The clock sampler starts first: the W warm-up steps run
    # directly before the timed region, so the timed steps do not begin on a GPU that idled down while we slept.
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    timed(px, max(a.warmup // 4, 3))
    timed(dset, max(a.warmup, 3))
    # ---- kernel-side timing: inputs from the device-resident dataset (> L2)
    barrier()
    t_start = time.time()
    st_dev = timed(dset, a.steps)
    launches_dev = trainer.launches_last_run or trainer.kernels_per_step * a.steps
    barrier()
    ms_dev = max_over_ranks(st_dev["device_ms"])
    # ---- end to end: pinned host -> device every step, loss/correct read back every step
    barrier()
    st_e2e = timed(px, a.steps)
    barrier()
    t_stop = time.time()
    ms_e2e = max_over_ranks(st_e2e["device_ms"])
    wall_e2e = max_over_ranks(st_e2e["seconds"] * 1e3)
    clocks = sampler.stop(t_start, t_stop)

    # ---- correctness of what was just timed, outside the timed regions: (1) every replica holds bit-identical
    # parameters after all those steps, (2) the loss is finite and fell, (3) a fresh trainer of the same configuration
    # follows the NumPy fp64 oracle (mean-gradient SGD over the ranks' shards) for a few steps.
    import hashlib
    flat = trainer.get_params()
    digest = hashlib.sha256(flat.tobytes()).hexdigest()[:16]
    digests = [digest]
    if multi:
        digests = [None] * world
        dist.all_gather_object(digests, digest)
    losses = st_e2e.get("per_step_loss")
    verified = {"replicas_equal": all(d == digests[0] for d in digests), "param_sha256_16": digests[0],
                "params_finite": bool(np.isfinite(flat).all()),
                "loss_finite": bool(np.isfinite(st_dev["loss"]) and np.isfinite(st_e2e["loss"])),
                "loss_first_last": [float(st_dev["per_step_loss"][0]), float(losses[-1])] if losses is not None else None}
    if a.model != "wide":
        from dmlp_b200.models import oracle
        k = 4
        vcfg = TrainerConfig(dims=m["dims"], batch=batch, lr=0.05, precision=m["precision"], allreduce=a.allreduce,
                             overlap=overlap, use_graph=not a.no_graph, fused=a.fused, seed=7)
        vt = Trainer(vcfg, device=local_rank, process_group=multi)
        vx, vy = synthetic_mnist(batch * world * k, seed=3, features=feat, num_classes=m["dims"][-1])
        ref = [(w.astype(np.float64), b.astype(np.float64)) for w, b in vt.weights()]
        vt.fit(vx, vy, k)  # rank r consumes batch s * world + r of the shared set
        for s_ in range(k):
            shards = [(vx[(s_ * world + r) * batch:(s_ * world + r + 1) * batch].astype(np.float64),
                       vy[(s_ * world + r) * batch:(s_ * world + r + 1) * batch]) for r in range(world)]
            ref, _, _ = oracle.dp_train_step(ref, shards, 0.05)
        err = max(max(np.abs(w - rw).max(), np.abs(b - rb).max()) for (w, b), (rw, rb) in zip(vt.weights(), ref))
        errs = [float(err)]
        if multi:
            errs = [None] * world
            dist.all_gather_object(errs, float(err))
        verified.update({"oracle_steps": k, "oracle_max_abs_err": max(errs), "oracle_tol": 2e-3,
                         "oracle_ok": bool(max(errs) < 2e-3), "dp_mode": vt.dp_mode})
        vt.close()
        if multi:
            dist.barrier()

    # ---- failure drill (optional, last: it leaves the communicator FAILED): one rank goes silent mid-run
    drill = None
    if a.fault and multi:
        from dmlp_b200.parallel.schedule import parse_fault
        from dmlp_b200.runtime import native
        dead_rank, dead_step = parse_fault(a.fault)
        drill_steps = dead_step + 32
        barrier()
        if rank == dead_rank:
            trainer.inject_fault(trainer.global_step + dead_step)
        t0 = time.perf_counter()
        err = None
        try:
            trainer.fit(px, py, drill_steps, return_metrics=False, batch_stride=1)
        except native.NativeError as e:
            err = e
        dt_ms = (time.perf_counter() - t0) * 1e3
        rec = {"rank": rank, "error": (err.message if err else None), "code": (err.code if err else 0),
               "returned_after_ms": dt_ms, "failed_state": trainer.failed}
        recs = [None] * world
        dist.all_gather_object(recs, rec)
        surv = [r_ for r_ in recs if r_["rank"] != dead_rank]
        drill = {"spec": a.fault, "watchdog_ms": a.watchdog_ms, "dead_rank": dead_rank,
                 "survivors_detected": all(r_["code"] == native.ERR_TIMEOUT and r_["failed_state"] for r_ in surv),
                 "survivors_naming_dead_rank": sum(1 for r_ in surv if f"waited for rank {dead_rank}" in (r_["error"] or "")),
                 "survivor_return_ms_max": max(r_["returned_after_ms"] for r_ in surv),
                 "detect_overhead_ms": max(r_["returned_after_ms"] for r_ in surv) - a.watchdog_ms,
                 "example": surv[0]["error"]}

    gbatch = batch * world
    value = a.steps * gbatch / (ms_dev / 1e3)
    e2e_value = a.steps * gbatch / (max(ms_e2e, wall_e2e) / 1e3)
    algo = a.allreduce
    out = {
        "metric": "MLP training samples/s (data-parallel, fused all-reduce+SGD)",
        "value": value, "unit": "samples/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": ms_dev / a.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": value / BASELINE_SAMPLES_PER_S,
        "dtype": "tf32 (fp32 storage+accumulate)" if m["precision"] == "tf32" else "bf16 (fp32 master weights+accumulate)",
        "data": "synthetic MNIST-shaped (28x28 float32, 10 classes), random-init weights U(-0.05,0.05)",
        "config": {"model": m["name"], "global_batch": gbatch, "per_gpu_batch": batch, "seq_len": None,
                   "parallelism": f"dp{world}", "optimizer": "SGD lr=0.01 fused into all-reduce",
                   "allreduce": algo, "fused_mode": a.fused, "overlap": overlap, "cuda_graph": not a.no_graph,
                   "l2_policy": f"inputs streamed from a {dset.nbytes >> 20} MiB device-resident dataset (> 126 MiB L2), "
                                f"{dev_batches} distinct batches; weights stay resident by design"},
        "e2e": {"value": e2e_value, "unit": "samples/s", "ms_per_step": max(ms_e2e, wall_e2e) / a.steps,
                "device_ms_per_step": ms_e2e / a.steps,
                "h2d_bytes_per_step": batch * feat * 4 + batch * 4, "d2h_bytes_per_step": 8,
                "api": "dmlp_b200.parallel.trainer.Trainer.fit (pinned host dataset, per-step loss/correct read back)",
                "final_loss": st_e2e.get("last_loss")},
        "gpu_launches": launches_dev,  # launches inside the device-timed region (fused path: one launch runs a whole
                                        # chunk of steps, weights updated in place between them inside the kernel)
        "kernels_per_step": trainer.kernels_per_step,
        "steps_per_launch": round(a.steps / max(launches_dev, 1), 2),
        "clocks": clocks,
        "verified": verified,
        "failure_drill": drill,
        "impl": "ours",
    }
    sys.stdout.flush()
    os.dup2(real_stdout, 1)
    if rank == 0:
        print(json.dumps(out), flush=True)
    os.dup2(2, 1)
    trainer.close()
    if multi:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
