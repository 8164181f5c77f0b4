"""Worker for the global-batch data-parallel mode of the fused step (torch.distributed.run, one rank per GPU).

For several shapes / step counts: N ranks x per-rank batch must follow the NumPy oracle's mean-gradient SGD
(`models/oracle.py`, fp64) and keep replicas bit-identical; a long run checks replica equality only. Then a
rank is made to "die" mid-run (fault injection): the survivors' in-kernel watchdogs must fire, `fit` must raise
TIMEOUT naming the dead peer, and the trainer must refuse further steps with ABORTED (the training-path
analogue of the reference's TestCoordinatorDeviceFailure, gpu_coordinator_server_test.go:370-429).
Prints 'MP_GX_OK' on rank 0.
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
    from dmlp_b200.models import oracle
    from dmlp_b200.parallel.trainer import Trainer, TrainerConfig
    from dmlp_b200.runtime import native
    from dmlp_b200.utils.data import synthetic_mnist

    long_steps = int(os.environ.get("DMLP_GX_SOAK_STEPS", "3000"))

    def replicas_equal(t):
        flat = torch.from_numpy(t.get_params()).cuda()
        gathered = [torch.empty_like(flat) for _ in range(world)]
        dist.all_gather(gathered, flat)
        return all(torch.equal(g, gathered[0]) for g in gathered)

    # ---- 1. oracle + replica equality
    for dims, batch, steps_list in (((784, 128, 64, 10), 64, (6, 11, 20)), ((784, 128, 10), 64, (9,)),
                                    ((784, 128, 64, 10), 32, (10,)), ((512, 128, 64, 10), 64, (7,))):
        cfg = TrainerConfig(dims=dims, batch=batch, lr=0.1, seed=7, fused=5)
        t = Trainer(cfg, device=local, process_group=True)
        assert t.dp_mode == 5, t.dp_mode
        total = sum(steps_list)
        x, y = synthetic_mnist(batch * world * total, seed=3, features=dims[0])
        params = [(w.astype(np.float64), b.astype(np.float64)) for w, b in t.weights()]
        s0 = 0
        for steps in steps_list:  # several fit calls: the exchange state carries over (odd / partial chunks included)
            st = t.fit(x, y, steps)
            assert np.isfinite(st["loss"])
            for s in range(s0, s0 + steps):
                shards = [(x[(s * world + r) * batch:(s * world + r + 1) * batch].astype(np.float64),
                           y[(s * world + r) * batch:(s * world + r + 1) * batch]) for r in range(world)]
                params, _, _ = oracle.dp_train_step(params, shards, 0.1)
            s0 += steps
            err = max(max(np.abs(w - rw).max(), np.abs(b - rb).max()) for (w, b), (rw, rb) in zip(t.weights(), params))
            assert err < 3e-3, (dims, batch, steps, err)
            assert replicas_equal(t), "replicas diverged"
        if rank == 0:
            print(f"gx oracle ok: dims={dims} batch={batch} steps={steps_list} max err {err:.2e}", flush=True)
        t.close()
        dist.barrier()

    # ---- 2. soak: replicas stay bit-identical, loss finite and falling
    cfg = TrainerConfig(dims=(784, 128, 64, 10), batch=64, lr=0.05, seed=11, fused=5)
    t = Trainer(cfg, device=local, process_group=True)
    assert t.dp_mode == 5, t.dp_mode
    x, y = synthetic_mnist(64 * world * 64, seed=5)
    t0 = time.time()
    st = t.fit(x, y, long_steps)
    dt = time.time() - t0
    assert np.isfinite(st["loss"]) and st["per_step_loss"][-50:].mean() < st["per_step_loss"][:50].mean(), st["loss"]
    assert replicas_equal(t), "replicas diverged in the soak run"
    if rank == 0:
        print(f"gx soak ok: {long_steps} steps, {dt * 1e6 / long_steps:.1f} us/step wall, loss "
              f"{st['per_step_loss'][:50].mean():.3f} -> {st['per_step_loss'][-50:].mean():.3f}", flush=True)
    t.close()
    dist.barrier()

    # ---- 2b. soak of the DEFAULT path (persistent compute + reducer launch): 10^5 steps, replicas bit-identical, and the
    # default flag protocol (gpu-scope fence before a flag, acquire load on the poll; "0" = round 1's relaxed poll) against
    # the conservative one (DMLP_COLL_STRICT=1: system-scope fences both sides) - all runs must end with the same bits
    import hashlib
    pdp_steps = int(os.environ.get("DMLP_PDP_SOAK_STEPS", "100000"))
    digests = {}
    for strict in ("2", "0", "1"):
        os.environ["DMLP_COLL_STRICT"] = strict  # read when the trainer builds its fused-step plan
        t = Trainer(TrainerConfig(dims=(784, 128, 64, 10), batch=64, lr=0.02, seed=13), device=local, process_group=True)
        assert t.dp_mode == 4, t.dp_mode
        x, y = synthetic_mnist(64 * world * 64, seed=9)
        t0 = time.time()
        st = t.fit(x, y, pdp_steps, return_metrics=False)
        dt = time.time() - t0
        assert replicas_equal(t), f"replicas diverged in the {pdp_steps}-step soak (strict={strict})"
        flat = t.get_params()
        assert np.isfinite(flat).all()
        digests[strict] = hashlib.sha256(flat.tobytes()).hexdigest()
        if rank == 0:
            print(f"soak ok (mode 4, strict={strict}): {pdp_steps} steps, {dt * 1e6 / pdp_steps:.1f} us/step wall, "
                  f"sha256 {digests[strict][:16]}", flush=True)
        t.close()
        dist.barrier()
    os.environ.pop("DMLP_COLL_STRICT", None)
    assert digests["2"] == digests["1"] == digests["0"], "the flag protocols produced different parameters"

    # ---- 3. a rank dies mid-run: the default path (persistent compute + reducer launch, mode 4) and the global-batch mode
    # ... and the per-layer path with its bucketed all-reduce overlapped with backward (mode 0: what the wide model runs)
    for fused, want_mode in ((-1, 4), (5, 5), (0, 0)):
        cfg = TrainerConfig(dims=(784, 128, 64, 10), batch=64, lr=0.05, seed=11, watchdog_ms=400, fused=fused,
                            overlap=(fused == 0))
        t = Trainer(cfg, device=local, process_group=True)
        assert t.dp_mode == want_mode, (t.dp_mode, want_mode)
        x, y = synthetic_mnist(64 * world * 16, seed=5)
        t.fit(x, y, 8)
        dead = world - 1
        if rank == dead:
            t.inject_fault(8 + 5)  # goes silent at the 6th step of the next call
        t0 = time.time()
        err = None
        try:
            t.fit(x, y, 16)
        except native.NativeError as e:
            err = e
        dt = time.time() - t0
        if rank != dead:
            assert err is not None and err.code == native.ERR_TIMEOUT, err
            assert f"waited for rank {dead}" in str(err), str(err)
            assert dt < 0.4 + 1.0, f"survivor returned after {dt:.2f} s"
            assert t.failed
            try:
                t.fit(x, y, 1)
                raise AssertionError("a FAILED trainer accepted more steps")
            except native.NativeError as e:
                assert e.code == native.ERR_ABORTED, e
            print(f"failure ok (mode {want_mode}) on rank {rank}: {err} after {dt * 1e3:.0f} ms", flush=True)
        t.close()
        dist.barrier()
    if rank == 0:
        print("MP_GX_OK", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
