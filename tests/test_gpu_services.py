"""The gRPC contract on ``backend=cuda``: logical devices served from real GPU memory, collectives run
by the native peer-memory kernels (all logical devices of this test live on GPU 0 → emulated ranks)."""
import time

import grpc
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture
def cuda_cluster():
    from dmlp_b200.proto import rpc
    from dmlp_b200.services import server as srv
    from dmlp_b200.services.backend import make_backend
    started = {"devices": [], "coord": None}

    def make(n, mem_size=0x2000, **kw):
        for i in range(n):
            started["devices"].append(srv.start_device_server(make_backend("cuda", i + 1, mem_size, 0)))
        started["coord"] = srv.start_coordinator_server(**kw)
        return rpc.CoordinatorStub(rpc.insecure_channel(started["coord"].addr)), started["devices"], started["coord"]

    yield make
    if started["coord"]:
        started["coord"].stop()
    for d in started["devices"]:
        d.stop()
        d.service.backend.close()


def _h2d(pb, client, dev, addr, data):
    return client.Memcpy(pb.MemcpyRequest(hostToDevice=pb.MemcpyHostToDeviceRequest(
        hostSrcData=data, dstDeviceId=pb.DeviceId(value=dev), dstMemAddr=pb.MemAddr(value=addr))))


def _d2h(pb, client, dev, addr, n=0):
    return client.Memcpy(pb.MemcpyRequest(deviceToHost=pb.MemcpyDeviceToHostRequest(
        srcDeviceId=pb.DeviceId(value=dev), srcMemAddr=pb.MemAddr(value=addr), numBytes=n))).deviceToHost.dstData


def test_memcpy_roundtrip_through_gpu_memory(cuda_cluster):
    from dmlp_b200.proto import gpu_sim as pb
    client, devs, _ = cuda_cluster(1, mem_size=0x1000)
    init = client.CommInit(pb.CommInitRequest(numDevices=1, device_addresses=[devs[0].addr]))
    did = init.devices[0].deviceId.value
    assert _h2d(pb, client, did, 0x1000, b"Hello GPU!").hostToDevice.success
    assert _d2h(pb, client, did, 0x1000) == b"Hello GPU!"
    assert devs[0].service.backend.pointer(0x1000) != 0
    with pytest.raises(grpc.RpcError) as e:
        _d2h(pb, client, did, 0x1800)
    assert e.value.code() == grpc.StatusCode.INTERNAL


@pytest.mark.parametrize("algo", ["auto", "ring", "oneshot", "twoshot"])
def test_allreduce_comparison_on_gpu(cuda_cluster, algo):
    """TestAllReduceComparison analogue: 3 devices, 1 MiB of 0x01, naive then AllReduceRing, with values checked."""
    from dmlp_b200.proto import gpu_sim as pb
    client, devs, coord = cuda_cluster(3)
    init = client.CommInit(pb.CommInitRequest(numDevices=3, device_addresses=[d.addr for d in devs]))
    assert len(init.devices) == 3
    assert coord.service.communicators[init.commId].native is not None, "native peer-memory path must be active"
    size = 1 << 20
    for md in init.devices:
        _h2d(pb, client, md.deviceId.value, 0x1000, b"\x01" * size)
    naive = client.NaiveAllReduce(pb.NaiveAllReduceRequest(commId=init.commId, dataSize=size, latencyMs=1))
    assert naive.success and naive.totalDataTransferred == 6 * size
    for md in init.devices:
        assert _d2h(pb, client, md.deviceId.value, 0x2000) == b"\x03" * size
    ring = client.AllReduceRing(pb.AllReduceRingRequest(commId=init.commId, count=size // 4),
                                metadata=(("dmlp-dtype", "u8"), ("dmlp-algo", algo)))
    assert ring.success
    for md in init.devices:
        got = _d2h(pb, client, md.deviceId.value, 0x1000)
        assert got[:size // 4] == b"\x03" * (size // 4) and got[size // 4:] == b"\x01" * (size - size // 4)
    assert client.GetCommStatus(pb.GetCommStatusRequest(commId=init.commId)).status == pb.SUCCESS


def test_ring_f32_padding_case_and_fused_sgd(cuda_cluster):
    from dmlp_b200.proto import gpu_sim as pb
    client, devs, _ = cuda_cluster(3)
    init = client.CommInit(pb.CommInitRequest(numDevices=3, device_addresses=[d.addr for d in devs]))
    nelem = 101770  # 407 080 bytes: not divisible by 3 (the reference's padding branch)
    rng = np.random.default_rng(0)
    grads = [rng.standard_normal(nelem).astype(np.float32) for _ in range(3)]
    w0 = rng.standard_normal(nelem).astype(np.float32)
    for md, g in zip(init.devices, grads):
        _h2d(pb, client, md.deviceId.value, 0x1000, g.tobytes())
        _h2d(pb, client, md.deviceId.value, 0x2000, w0.tobytes())
    # gradient average + SGD fused into the collective: weights at 0x2000 are updated in the same kernel
    ok = client.AllReduceRing(pb.AllReduceRingRequest(commId=init.commId, count=nelem * 4),
                              metadata=(("dmlp-sgd-lr", "0.01"), ("dmlp-sgd-weights", "0x2000")))
    assert ok.success
    ref = w0.astype(np.float64) - 0.01 * np.mean(np.stack(grads).astype(np.float64), axis=0)
    ws = [np.frombuffer(_d2h(pb, client, md.deviceId.value, 0x2000), np.float32) for md in init.devices]
    for w in ws:
        assert np.abs(w - ref).max() < 1e-5 and np.array_equal(w, ws[0])
    # plain all-reduce (sum) in place
    assert client.AllReduceRing(pb.AllReduceRingRequest(commId=init.commId, count=nelem * 4)).success
    s = np.sum(np.stack(grads).astype(np.float64), axis=0)
    for md in init.devices:
        assert np.abs(np.frombuffer(_d2h(pb, client, md.deviceId.value, 0x1000), np.float32) - s).max() < 1e-4


def test_concurrent_allreduce_on_two_communicators_over_the_same_gpus(cuda_cluster):
    """Two communicators over the same logical devices, hammered from two threads at once (the gRPC server has 16
    workers): native collectives are serialised per process, so neither mixes buffers nor starves the other into
    its watchdog (coordinator.py native collectives were not serialised)."""
    import threading
    from dmlp_b200.proto import gpu_sim as pb
    client, devs, coord = cuda_cluster(3)
    addrs = [d.addr for d in devs]
    c1 = client.CommInit(pb.CommInitRequest(numDevices=3, device_addresses=addrs))
    c2 = client.CommInit(pb.CommInitRequest(numDevices=3, device_addresses=addrs))
    n = 64 * 1024
    errs = []

    def worker(comm, addr, val, lr):
        try:
            for it in range(12):
                for k, md in enumerate(comm.devices):
                    _h2d(pb, client, md.deviceId.value, addr, np.full(n, val * (k + 1), np.float32).tobytes())
                    _h2d(pb, client, md.deviceId.value, addr + 0x100, np.ones(n, np.float32).tobytes())
                memaddrs = {k: pb.MemAddr(value=addr) for k in range(3)}
                r = client.AllReduceRing(pb.AllReduceRingRequest(commId=comm.commId, count=4 * n, memAddrs=memaddrs),
                                         metadata=(("dmlp-dtype", "f32"), ("dmlp-sgd-lr", str(lr)),
                                                   ("dmlp-sgd-weights", hex(addr + 0x100))))
                assert r.success
                want = 1.0 - lr * (val * 6) / 3
                for md in comm.devices:
                    got = np.frombuffer(_d2h(pb, client, md.deviceId.value, addr + 0x100), np.float32)
                    assert np.allclose(got, want, rtol=1e-5, atol=1e-6), (it, got[:4], want)
        except Exception as e:  # noqa: BLE001
            errs.append(e)

    t1 = threading.Thread(target=worker, args=(c1, 0x1000, 1.0, 0.5))
    t2 = threading.Thread(target=worker, args=(c2, 0x1400, 3.0, 0.125))
    t1.start(); t2.start(); t1.join(); t2.join()
    assert not errs, errs
    for c in (c1, c2):
        assert client.GetCommStatus(pb.GetCommStatusRequest(commId=c.commId)).status == pb.SUCCESS


def test_group_end_submits_the_queue_with_one_wait(cuda_cluster):
    """GroupStart .. GroupEnd: the queued collectives are launched back to back and waited for once; each
    keeps its own learning-rate scalar."""
    from dmlp_b200.proto import gpu_sim as pb
    client, devs, coord = cuda_cluster(2)
    init = client.CommInit(pb.CommInitRequest(numDevices=2, device_addresses=[d.addr for d in devs]))
    n = 4096
    for k, md in enumerate(init.devices):
        _h2d(pb, client, md.deviceId.value, 0x1000, np.full(n, float(k + 1), np.float32).tobytes())
        _h2d(pb, client, md.deviceId.value, 0x1400, np.ones(n, np.float32).tobytes())
    assert client.GroupStart(pb.GroupStartRequest(commId=init.commId)).success
    for lr in (0.5, 0.25):  # two fused updates of the same weights, queued
        memaddrs = {k: pb.MemAddr(value=0x1000) for k in range(2)}
        assert client.AllReduceRing(pb.AllReduceRingRequest(commId=init.commId, count=4 * n, memAddrs=memaddrs),
                                    metadata=(("dmlp-dtype", "f32"), ("dmlp-sgd-lr", str(lr)),
                                              ("dmlp-sgd-weights", hex(0x1400)))).success
    assert client.GroupEnd(pb.GroupEndRequest(commId=init.commId)).success
    for md in init.devices:
        got = np.frombuffer(_d2h(pb, client, md.deviceId.value, 0x1400), np.float32)
        assert np.allclose(got, 1.0 - (0.5 + 0.25) * 1.5), got[:4]
    assert client.GetCommStatus(pb.GetCommStatusRequest(commId=init.commId)).status == pb.SUCCESS


def test_p2p_stream_is_a_device_copy_kernel(cuda_cluster):
    """BeginSend/BeginReceive + payload-free StreamSend → P2P copy kernel; status follows its CUDA event."""
    from dmlp_b200.proto import gpu_sim as pb
    from dmlp_b200.proto import rpc
    client, devs, _ = cuda_cluster(2)
    client.CommInit(pb.CommInitRequest(numDevices=2, device_addresses=[d.addr for d in devs]))
    d0 = rpc.DeviceStub(rpc.insecure_channel(devs[0].addr))
    payload = bytes(range(256)) * 4096  # 1 MiB
    d0.Memcpy(pb.MemcpyRequest(hostToDevice=pb.MemcpyHostToDeviceRequest(hostSrcData=payload, dstMemAddr=pb.MemAddr(value=0x1000))))
    s = d0.BeginSend(pb.BeginSendRequest(sendBuffAddr=pb.MemAddr(value=0x1000), numBytes=len(payload), dstRank=pb.Rank(value=1)))
    d0.BeginReceive(pb.BeginReceiveRequest(streamId=s.streamId, recvBuffAddr=pb.MemAddr(value=0x2000), numBytes=len(payload),
                                           srcRank=pb.Rank(value=0)))
    assert d0.StreamSend(iter([pb.DataChunk(streamId=s.streamId.value)])).success
    deadline = time.time() + 5
    st = pb.IN_PROGRESS
    while st == pb.IN_PROGRESS and time.time() < deadline:
        st = d0.GetStreamStatus(pb.GetStreamStatusRequest(streamId=s.streamId)).status
    assert st == pb.SUCCESS
    assert devs[1].service.backend.read(0x2000) == payload


def test_gpu_failure_mid_allreduce_is_surfaced(cuda_cluster):
    """A rank that stops participating (GPU fault injection) → in-kernel watchdog → the AllReduceRing RPC
    fails, the communicator is FAILED and refuses further collectives (reference :281-285,341-344)."""
    from dmlp_b200.proto import gpu_sim as pb
    client, devs, coord = cuda_cluster(3, watchdog_ms=300, health_interval=30)
    init = client.CommInit(pb.CommInitRequest(numDevices=3, device_addresses=[d.addr for d in devs]))
    for md in init.devices:
        _h2d(pb, client, md.deviceId.value, 0x1000, np.ones(4096, np.float32).tobytes())
    assert client.AllReduceRing(pb.AllReduceRingRequest(commId=init.commId, count=16384)).success
    coord.service.communicators[init.commId].native.inject_dead_ranks(1 << 2)
    t0 = time.time()
    with pytest.raises(grpc.RpcError) as e:
        client.AllReduceRing(pb.AllReduceRingRequest(commId=init.commId, count=16384))
    assert e.value.code() == grpc.StatusCode.UNKNOWN and "waited for rank" in e.value.details()
    assert time.time() - t0 < 5, "the watchdog, not a hang, must end the collective"
    assert client.GetCommStatus(pb.GetCommStatusRequest(commId=init.commId)).status == pb.FAILED
    with pytest.raises(grpc.RpcError) as e:
        client.AllReduceRing(pb.AllReduceRingRequest(commId=init.commId, count=16384))
    assert e.value.code() == grpc.StatusCode.FAILED_PRECONDITION


@pytest.mark.parametrize("dims", [(784, 128, 10), (784, 128, 64, 10)])
def test_run_forward_backward_on_the_gpu_match_the_oracle(dims):
    """RunForward / RunBackward (the reference's never-implemented device RPCs) on ``backend=cuda``: buffers stay in
    GPU memory, the math runs on the tcgen05 layer; logits and the gradient blob match the fp64 oracle."""
    from dmlp_b200.models import oracle
    from dmlp_b200.models.layout import ParamLayout
    from dmlp_b200.proto import gpu_sim as pb
    from dmlp_b200.services.backend import make_backend
    from dmlp_b200.services.device import GPUDeviceService
    from dmlp_b200.services.testing import FakeContext
    from dmlp_b200.utils.data import synthetic_mnist
    backend = make_backend("cuda", 7, 1 << 20, 0)
    svc, ctx = GPUDeviceService(backend), FakeContext()
    try:
        lay = ParamLayout(dims)
        flat = lay.init_reference(11)
        B = 40
        x, y = synthetic_mnist(B, seed=4)

        def h2d(addr, data):
            svc.Memcpy(pb.MemcpyRequest(hostToDevice=pb.MemcpyHostToDeviceRequest(
                hostSrcData=data, dstDeviceId=pb.DeviceId(value=7), dstMemAddr=pb.MemAddr(value=addr))), ctx)

        def d2h(addr):
            return svc.Memcpy(pb.MemcpyRequest(deviceToHost=pb.MemcpyDeviceToHostRequest(
                srcDeviceId=pb.DeviceId(value=7), srcMemAddr=pb.MemAddr(value=addr))), ctx).deviceToHost.dstData

        h2d(0x2000, lay.to_reference_bytes(flat))
        h2d(0x3000, x.astype("<f4").tobytes())
        h2d(0x4000, np.eye(10, dtype="<f4")[y].tobytes())
        assert svc.RunForward(pb.RunForwardRequest(deviceId=7, inputAddr=0x3000, outputAddr=0x5000), ctx).success
        params = [(w.astype(np.float64), b.astype(np.float64)) for w, b in lay.views(flat)]
        logits, _ = oracle.forward(params, x.astype(np.float64))
        got = np.frombuffer(d2h(0x5000), "<f4").reshape(B, 10)
        assert np.abs(got - logits).max() < 5e-3 * max(1.0, np.abs(logits).max())  # tf32 operands
        assert svc.RunBackward(pb.RunBackwardRequest(deviceId=7, gradientAddr=0x1000), ctx).success
        _, _, grads = oracle.loss_and_grads(params, x.astype(np.float64), y)
        blob = d2h(0x1000)
        assert len(blob) == lay.reference_count * 4
        for (gw, gb), (rw, rb) in zip(lay.views(lay.from_reference_bytes(blob)), grads):
            assert np.abs(gw - rw).max() < 2e-2 * max(1e-3, np.abs(rw).max())
            assert np.abs(gb - rb).max() < 2e-2 * max(1e-3, np.abs(rb).max())
        # weights updated through Memcpy are picked up by the next call (no stale copy on the device side)
        h2d(0x2000, lay.to_reference_bytes(np.zeros_like(flat)))
        assert svc.RunForward(pb.RunForwardRequest(deviceId=7, inputAddr=0x3000, outputAddr=0x5000), ctx).success
        assert np.abs(np.frombuffer(d2h(0x5000), "<f4")).max() == 0.0
    finally:
        backend.close()


def test_grpc_client_trains_on_cuda_devices(cuda_cluster):
    """dmlp_b200.grpc_client against 3 logical devices in GPU memory: shards by Memcpy, RunForward/RunBackward on the
    tcgen05 layer, AllReduceRing with the fused average+SGD kernel. Must follow the oracle's data-parallel steps."""
    from dmlp_b200.grpc_client import GrpcTrainer
    from dmlp_b200.models import oracle
    from dmlp_b200.models.layout import ParamLayout
    from dmlp_b200.utils.data import synthetic_mnist
    dims, batch, lr, steps, n = (784, 128, 64, 10), 32, 0.1, 3, 3
    _, devs, coord = cuda_cluster(n, mem_size=1 << 20)
    t = GrpcTrainer(coord.addr, [d.addr for d in devs], dims=dims, batch=batch, lr=lr, seed=5)
    assert coord.service.communicators[t.comm_id].native is not None
    x, y = synthetic_mnist(batch * n * steps, seed=2)
    lay = ParamLayout(dims)
    ref = [(w.astype(np.float64), b.astype(np.float64)) for w, b in lay.views(lay.init_reference(5))]
    stats = t.fit(x, y, steps)
    for s in range(steps):
        shards = [(x[(s * n + d) * batch:(s * n + d + 1) * batch].astype(np.float64),
                   y[(s * n + d) * batch:(s * n + d + 1) * batch]) for d in range(n)]
        ref, _, _ = oracle.dp_train_step(ref, shards, lr)
    blobs = [t.weights_blob(d) for d in range(n)]
    assert blobs[0] == blobs[1] == blobs[2], "replicas diverged"
    for (w, b), (rw, rb) in zip(lay.views(lay.from_reference_bytes(blobs[0])), ref):
        assert np.abs(w - rw).max() < 3e-3 and np.abs(b - rb).max() < 3e-3
    assert all(np.isfinite(s["loss"]) for s in stats)
    t.close()
