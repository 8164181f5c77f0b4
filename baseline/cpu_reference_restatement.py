"""Timed CPU restatement of the reference (it cannot be rebuilt here: no Go toolchain, no network).

Measures on the host CPU (a) the reference's MLP step math (NumPy restatement of
DSML/client/client.go:112-267, batch 64) and (b) the reference's "3 simulated devices via localhost
gRPC" configuration through this repo's own `backend=sim` services: per step the client uploads the
gradient to every device, calls AllReduceRing, reads it back and re-uploads the weights — the same
RPC pattern as client.go:596-647. Prints one JSON line.
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main(steps=50):
    from dmlp_b200.models import oracle
    from dmlp_b200.models.layout import ParamLayout
    from dmlp_b200.proto import gpu_sim as pb
    from dmlp_b200.proto import rpc
    from dmlp_b200.services import server as srv
    from dmlp_b200.services.backend import SimBackend
    from dmlp_b200.utils.data import synthetic_mnist

    out = {}
    x, y = synthetic_mnist(64 * steps, seed=0)
    for dims in ((784, 128, 10), (784, 128, 64, 10)):
        lay = ParamLayout(dims)
        params = [(w.copy(), b.copy()) for w, b in lay.views(lay.init_reference(0))]
        t0 = time.perf_counter()
        for s in range(steps):
            params, _, _ = oracle.train_step(params, x[s * 64:(s + 1) * 64], y[s * 64:(s + 1) * 64], 0.01)
        dt = time.perf_counter() - t0
        out["numpy_step_%s" % "-".join(map(str, dims))] = {"ms_per_step": 1e3 * dt / steps, "samples_per_s": 64 * steps / dt}

    devs = [srv.start_device_server(SimBackend(i + 1, 0x3000), register_local=False) for i in range(3)]
    coord = srv.start_coordinator_server(health_interval=3600)
    client = rpc.CoordinatorStub(rpc.insecure_channel(coord.addr))
    init = client.CommInit(pb.CommInitRequest(numDevices=3, device_addresses=[d.addr for d in devs]))
    lay = ParamLayout((784, 128, 64, 10))
    params = [(w.copy(), b.copy()) for w, b in lay.views(lay.init_reference(0))]
    n = 10
    t0 = time.perf_counter()
    for s in range(n):
        _, _, grads = oracle.loss_and_grads(params, x[s * 64:(s + 1) * 64], y[s * 64:(s + 1) * 64])
        blob = b"".join(np.ascontiguousarray(g, "<f4").tobytes() for pair in grads for g in pair)
        for md in init.devices:
            client.Memcpy(pb.MemcpyRequest(hostToDevice=pb.MemcpyHostToDeviceRequest(
                hostSrcData=blob, dstDeviceId=md.deviceId, dstMemAddr=pb.MemAddr(value=0x1000))))
        client.AllReduceRing(pb.AllReduceRingRequest(commId=init.commId, count=len(blob)))
        agg = client.Memcpy(pb.MemcpyRequest(deviceToHost=pb.MemcpyDeviceToHostRequest(
            srcDeviceId=init.devices[0].deviceId, srcMemAddr=pb.MemAddr(value=0x1000)))).deviceToHost.dstData
        vals = np.frombuffer(agg, "<f4") / 3.0
        cur = 0
        new = []
        for (w, b) in params:
            gw = vals[cur:cur + w.size].reshape(w.shape); cur += w.size
            gb = vals[cur:cur + b.size]; cur += b.size
            new.append((w - 0.01 * gw, b - 0.01 * gb))
        params = new
        wblob = b"".join(np.ascontiguousarray(t, "<f4").tobytes() for pair in params for t in pair)
        for md in init.devices:
            client.Memcpy(pb.MemcpyRequest(hostToDevice=pb.MemcpyHostToDeviceRequest(
                hostSrcData=wblob, dstDeviceId=md.deviceId, dstMemAddr=pb.MemAddr(value=0x2000))))
    dt = time.perf_counter() - t0
    out["sim_grpc_3dev_step"] = {"ms_per_step": 1e3 * dt / n, "samples_per_s": 64 * n / dt}
    # naive vs ring, 1 MiB x 3 devices (TestAllReduceComparison)
    for md in init.devices:
        client.Memcpy(pb.MemcpyRequest(hostToDevice=pb.MemcpyHostToDeviceRequest(
            hostSrcData=b"\x01" * (1 << 20), dstDeviceId=md.deviceId, dstMemAddr=pb.MemAddr(value=0x1000))))
    t0 = time.perf_counter()
    r = client.NaiveAllReduce(pb.NaiveAllReduceRequest(commId=init.commId, dataSize=1 << 20, latencyMs=10))
    out["sim_naive_1MiB_3dev_ms"] = {"reported": r.totalTimeMs, "wall": 1e3 * (time.perf_counter() - t0)}
    t0 = time.perf_counter()
    client.AllReduceRing(pb.AllReduceRingRequest(commId=init.commId, count=1 << 20), metadata=(("dmlp-dtype", "u8"),))
    out["sim_ring_1MiB_3dev_ms"] = 1e3 * (time.perf_counter() - t0)
    coord.stop()
    for d in devs:
        d.stop()
    out["host"] = os.uname().nodename
    print(json.dumps(out))


if __name__ == "__main__":
    main()
