"""BASELINE.json config 1 analogue: MLP 784-128-64-10, batch 64, SGD, 3 devices behind localhost gRPC — with the
devices' buffers and compute on a GPU (`backend=cuda`, all logical devices on GPU 0 unless --gpus lists more).

The reference's client needs ~78 ms per batch on this configuration (732 s for 9370 batches); it computes
on the CPU and ships one gradient to 3 simulated devices. Here every device trains its own shard (3 x 64 samples per
step) and only RPCs cross the process boundary.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--backend", default="cuda")
    ap.add_argument("--gpus", default="0,0,0")
    ap.add_argument("--steps", type=int, default=60)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    from dmlp_b200.grpc_client import GrpcTrainer
    from dmlp_b200.services import server as srv
    from dmlp_b200.services.backend import make_backend
    from dmlp_b200.utils.data import synthetic_mnist
    gpus = [int(g) for g in a.gpus.split(",")]
    devs = [srv.start_device_server(make_backend(a.backend, i + 1, 1 << 20, g)) for i, g in enumerate(gpus)]
    coord = srv.start_coordinator_server()
    try:
        t = GrpcTrainer(coord.addr, [d.addr for d in devs], batch=64, lr=0.01)
        x, y = synthetic_mnist(64 * len(gpus) * 20, seed=0)
        t.fit(x, y, 5)  # warm-up
        t0 = time.perf_counter()
        st = t.fit(x, y, a.steps, first_batch=5 * len(gpus))
        dt = time.perf_counter() - t0
        out = {"config": "MLP 784-128-64-10, batch 64 per device, %d devices via localhost gRPC, backend=%s" % (len(gpus), a.backend),
               "steps": a.steps, "ms_per_step": round(dt / a.steps * 1e3, 2),
               "samples_per_s": round(a.steps * 64 * len(gpus) / dt, 1), "first_loss": round(st[0]["loss"], 4),
               "last_loss": round(st[-1]["loss"], 4), "native_peer_memory_path": coord.service.communicators[t.comm_id].native is not None,
               "reference_ms_per_batch": 78.1, "reference_samples_per_s": 819}
        t.close()
    finally:
        coord.stop()
        for d in devs:
            d.stop()
            d.service.backend.close() if hasattr(d.service.backend, "close") else None
    print(json.dumps(out))
    if a.out:
        open(a.out, "w").write(json.dumps(out, indent=1) + "\n")


if __name__ == "__main__":
    main()
