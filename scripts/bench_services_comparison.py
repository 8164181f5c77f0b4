"""BASELINE.json config 2 on real GPUs: the reference's TestAllReduceComparison (1 MiB, 3 devices, naive vs ring)
through this repo's gRPC coordinator + device services with backend=cuda, one logical device per GPU.

    python scripts/bench_services_comparison.py [--gpus 0,1,2] [--out gpurun_out/services_comparison.json]

Reference numbers (macOS CPU, localhost gRPC, `latencyMs=10`): naive 83 ms, ring 8 ms.
What is timed: the wall time of the RPC as the client sees it (the reference's test does the same), plus the
coordinator's own `totalTimeMs` for the naive mode. Results are checked byte for byte.
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
    ap.add_argument("--gpus", default="0,1,2")
    ap.add_argument("--out", default="")
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--backend", default="cuda", help="cuda | sim (sim: CPU dry run of this script)")
    a = ap.parse_args()
    gpus = [int(g) for g in a.gpus.split(",")]
    from dmlp_b200.proto import gpu_sim as pb
    from dmlp_b200.proto import rpc
    from dmlp_b200.services import server as srv
    from dmlp_b200.services.backend import make_backend

    size = 1 << 20
    devs = [srv.start_device_server(make_backend(a.backend, i + 1, 0x2000, g)) for i, g in enumerate(gpus)]
    coord = srv.start_coordinator_server()
    client = rpc.CoordinatorStub(rpc.insecure_channel(coord.addr))
    out = {"config": "AllReduce 1 MiB comparison (naive vs ring) across %d GPUs, gRPC coordinator, backend=%s" % (len(gpus), a.backend),
           "gpus": gpus, "bytes": size}
    try:
        init = client.CommInit(pb.CommInitRequest(numDevices=len(gpus), device_addresses=[d.addr for d in devs]))
        native_path = coord.service.communicators[init.commId].native is not None
        out["native_peer_memory_path"] = native_path

        def h2d(dev, addr, data):
            return client.Memcpy(pb.MemcpyRequest(hostToDevice=pb.MemcpyHostToDeviceRequest(
                hostSrcData=data, dstDeviceId=pb.DeviceId(value=dev), dstMemAddr=pb.MemAddr(value=addr))))

        def d2h(dev, addr):
            return client.Memcpy(pb.MemcpyRequest(deviceToHost=pb.MemcpyDeviceToHostRequest(
                srcDeviceId=pb.DeviceId(value=dev), srcMemAddr=pb.MemAddr(value=addr), numBytes=0))).deviceToHost.dstData

        n = len(gpus)
        for lat in (10, 0):
            ts, inner = [], []
            for _ in range(3 if lat else a.reps):
                t0 = time.perf_counter()
                r = client.NaiveAllReduce(pb.NaiveAllReduceRequest(commId=init.commId, dataSize=size, latencyMs=lat))
                ts.append((time.perf_counter() - t0) * 1e3)
                inner.append(r.totalTimeMs)
                assert r.success and r.totalDataTransferred == 2 * n * size
            for md in init.devices:
                assert d2h(md.deviceId.value, 0x2000) == bytes([n]) * size
            out[f"naive_latencyMs{lat}"] = {"rpc_ms_min": round(min(ts), 3), "rpc_ms_median": round(sorted(ts)[len(ts) // 2], 3),
                                           "coordinator_totalTimeMs_min": min(inner)}
        for dtype, count in (("u8", size), ("f32", size)):  # count is BYTES for every dtype
            for algo in ("auto", "ring"):
                ts = []
                for rep in range(a.reps):
                    for md in init.devices:
                        h2d(md.deviceId.value, 0x1000, b"\x01" * size if dtype == "u8" else
                            __import__("numpy").ones(size // 4, "<f4").tobytes())
                    t0 = time.perf_counter()
                    r = client.AllReduceRing(pb.AllReduceRingRequest(commId=init.commId, count=count),
                                             metadata=(("dmlp-dtype", dtype), ("dmlp-algo", algo)))
                    ts.append((time.perf_counter() - t0) * 1e3)
                    assert r.success
                got = d2h(init.devices[0].deviceId.value, 0x1000)
                if dtype == "u8":
                    assert got == bytes([n]) * size, "ring result"
                else:
                    import numpy as np
                    assert np.all(np.frombuffer(got, "<f4") == float(n)), "ring result"
                out[f"ring_{dtype}_{algo}"] = {"rpc_ms_min": round(min(ts), 3), "rpc_ms_median": round(sorted(ts)[len(ts) // 2], 3)}
        out["reference_published_ms"] = {"naive_latencyMs10": 83, "ring": 8}
    finally:
        coord.stop()
        for d in devs:
            d.stop()
            d.service.backend.close()
    print(json.dumps(out))
    if a.out:
        with open(a.out, "w") as f:
            f.write(json.dumps(out, indent=1) + "\n")


if __name__ == "__main__":
    main()
