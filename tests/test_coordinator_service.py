"""Coordinator integration tests over real gRPC on loopback — the reference's black-box suite
(DSML/gpu_coordinator_service/gpu_coordinator_server_test.go:67-429 and
allreduce_comparison_test.go:32-133) re-expressed against ``backend=sim``, plus the numerical
checks the reference never had."""
import time

import grpc
import numpy as np
import pytest

from dmlp_b200.proto import gpu_sim as pb
from dmlp_b200.proto import rpc
from dmlp_b200.services import server as srv
from dmlp_b200.services.backend import SimBackend


@pytest.fixture
def cluster():
    """Starts servers in-process on ephemeral ports (same fixture idea as the reference, :20-64)."""
    started = {"devices": [], "coord": None}

    def make(n_devices, mem_size=0x1000, first_id=1, register_local=True, **coord_kwargs):
        for i in range(n_devices):
            started["devices"].append(srv.start_device_server(SimBackend(first_id + i, mem_size),
                                                              register_local=register_local))
        started["coord"] = srv.start_coordinator_server(**coord_kwargs)
        ch = rpc.insecure_channel(started["coord"].addr)
        return rpc.CoordinatorStub(ch), started["devices"], started["coord"]

    yield make
    if started["coord"]:
        started["coord"].stop()
    for d in started["devices"]:
        d.stop()


def h2d(client, dev, addr, data):
    return client.Memcpy(pb.MemcpyRequest(hostToDevice=pb.MemcpyHostToDeviceRequest(
        hostSrcData=data, dstDeviceId=pb.DeviceId(value=dev), dstMemAddr=pb.MemAddr(value=addr))))


def d2h(client, dev, addr, n=0):
    return client.Memcpy(pb.MemcpyRequest(deviceToHost=pb.MemcpyDeviceToHostRequest(
        srcDeviceId=pb.DeviceId(value=dev), srcMemAddr=pb.MemAddr(value=addr), numBytes=n))).deviceToHost.dstData


def test_comm_init_with_invalid_devices(cluster):
    client, devs, _ = cluster(1, probe_timeout=0.5)
    with pytest.raises(grpc.RpcError) as e:
        client.CommInit(pb.CommInitRequest(numDevices=3, device_addresses=[devs[0].addr, "localhost:99999", "wrongformat"]))
    assert e.value.code() == grpc.StatusCode.INTERNAL
    # the failed call still consumed a communicator id (reference :125-126,167-169)
    ok = client.CommInit(pb.CommInitRequest(numDevices=1, device_addresses=[devs[0].addr]))
    assert ok.success and ok.commId == 1


def test_memcpy_host_to_device_and_back(cluster):
    client, devs, _ = cluster(1)
    init = client.CommInit(pb.CommInitRequest(numDevices=1, device_addresses=[devs[0].addr]))
    assert init.success and len(init.devices) == 1
    dev_id = init.devices[0].deviceId.value
    assert (init.devices[0].minMemAddr.value, init.devices[0].maxMemAddr.value) == (0x1000, 0x2000)
    assert h2d(client, dev_id, 0x1000, b"Hello GPU!").hostToDevice.success
    assert d2h(client, dev_id, 0x1000, 10) == b"Hello GPU!"
    # one source of truth: the bytes are in the *device's* memory, not a coordinator-side map
    assert devs[0].service.backend.read(0x1000) == b"Hello GPU!"
    assert client.CommDestroy(pb.CommDestroyRequest(commId=init.commId)).success


def test_memcpy_error_codes(cluster):
    client, devs, _ = cluster(1)
    with pytest.raises(grpc.RpcError) as e:
        h2d(client, 1, 0x3000, b"x")
    assert e.value.code() == grpc.StatusCode.INVALID_ARGUMENT
    with pytest.raises(grpc.RpcError) as e:
        d2h(client, 4242, 0x1000)
    assert e.value.code() == grpc.StatusCode.INTERNAL
    with pytest.raises(grpc.RpcError) as e:
        client.Memcpy(pb.MemcpyRequest())
    assert e.value.code() == grpc.StatusCode.INVALID_ARGUMENT
    # unknown device: staged on the coordinator like the reference, readable back
    h2d(client, 555, 0x1000, b"staged")
    assert d2h(client, 555, 0x1000) == b"staged"


def test_group_operations_without_comm(cluster):
    client, _, _ = cluster(0)
    for call, req in ((client.GroupStart, pb.GroupStartRequest(commId=9999)), (client.GroupEnd, pb.GroupEndRequest(commId=9999))):
        with pytest.raises(grpc.RpcError) as e:
            call(req)
        assert e.value.code() == grpc.StatusCode.NOT_FOUND


def test_unknown_comm_is_not_found_everywhere(cluster):
    client, _, _ = cluster(0)
    calls = [(client.CommDestroy, pb.CommDestroyRequest(commId=8888)), (client.GetCommStatus, pb.GetCommStatusRequest(commId=8888)),
             (client.AllReduceRing, pb.AllReduceRingRequest(commId=8888, count=4)),
             (client.NaiveAllReduce, pb.NaiveAllReduceRequest(commId=8888, dataSize=4)),
             (client.CommFinalize, pb.CommFinalizeRequest(commId=8888))]
    for call, req in calls:
        with pytest.raises(grpc.RpcError) as e:
            call(req)
        assert e.value.code() == grpc.StatusCode.NOT_FOUND


def test_coordinator_allreduce_ring_zero_devices(cluster):
    """The reference's TestCoordinatorAllReduceRing sends NO addresses (:259) → a valid 0-device comm."""
    client, _, _ = cluster(3)
    init = client.CommInit(pb.CommInitRequest(numDevices=3, device_addresses=[]))
    assert init.success and len(init.devices) == 0
    assert client.GroupStart(pb.GroupStartRequest(commId=init.commId)).success
    assert client.AllReduceRing(pb.AllReduceRingRequest(commId=init.commId, count=16)).success
    assert client.GetCommStatus(pb.GetCommStatusRequest(commId=init.commId)).status == pb.SUCCESS
    assert client.GroupEnd(pb.GroupEndRequest(commId=init.commId)).success
    assert client.CommDestroy(pb.CommDestroyRequest(commId=init.commId)).success


def test_allreduce_ring_single_device(cluster):
    client, devs, _ = cluster(1)
    init = client.CommInit(pb.CommInitRequest(numDevices=1, device_addresses=[devs[0].addr]))
    assert client.GetCommStatus(pb.GetCommStatusRequest(commId=init.commId)).status == pb.IN_PROGRESS
    assert client.AllReduceRing(pb.AllReduceRingRequest(commId=init.commId, count=0)).success
    assert client.GetCommStatus(pb.GetCommStatusRequest(commId=init.commId)).status == pb.SUCCESS
    assert client.CommDestroy(pb.CommDestroyRequest(commId=init.commId)).success


def test_coordinator_device_failure(cluster):
    """Stop a device server → FAILED within one health interval (reference waits 6 s for a 5 s interval)."""
    client, devs, _ = cluster(2, health_interval=0.5, probe_timeout=0.5)
    init = client.CommInit(pb.CommInitRequest(numDevices=2, device_addresses=[d.addr for d in devs]))
    assert client.GetCommStatus(pb.GetCommStatusRequest(commId=init.commId)).status == pb.IN_PROGRESS
    devs[0].stop()
    deadline = time.time() + 6.0
    status = pb.IN_PROGRESS
    while time.time() < deadline and status != pb.FAILED:
        time.sleep(0.2)
        status = client.GetCommStatus(pb.GetCommStatusRequest(commId=init.commId)).status
    assert status == pb.FAILED
    for call, req in ((client.AllReduceRing, pb.AllReduceRingRequest(commId=init.commId, count=4)),
                      (client.NaiveAllReduce, pb.NaiveAllReduceRequest(commId=init.commId, dataSize=4))):
        with pytest.raises(grpc.RpcError) as e:
            call(req)
        assert e.value.code() == grpc.StatusCode.FAILED_PRECONDITION


@pytest.mark.parametrize("local", [True, False], ids=["in-process-devices", "remote-devices"])
def test_allreduce_comparison(cluster, local):
    """TestAllReduceComparison (:32-133): 3 devices, memSize 0x2000, 1 MiB of 0x01, naive then ring —
    here with the numerical check the reference lacks."""
    client, devs, _ = cluster(3, mem_size=0x2000, register_local=local)
    init = client.CommInit(pb.CommInitRequest(numDevices=3, device_addresses=[d.addr for d in devs]))
    assert len(init.devices) == 3
    size = 1 << 20
    for md in init.devices:
        assert h2d(client, md.deviceId.value, 0x1000, b"\x01" * size).hostToDevice.success
    naive = client.NaiveAllReduce(pb.NaiveAllReduceRequest(commId=init.commId, dataSize=size, latencyMs=10))
    assert naive.success and naive.totalDataTransferred == 2 * 3 * size and naive.totalTimeMs >= 60
    for d in devs:
        assert d.service.backend.read(0x2000) == b"\x03" * size
    ring = client.AllReduceRing(pb.AllReduceRingRequest(commId=init.commId, count=size // 4),
                                metadata=(("dmlp-dtype", "u8"),))
    assert ring.success
    for md in init.devices:  # first count bytes are the byte-wise sum, the rest is untouched
        got = d2h(client, md.deviceId.value, 0x1000)
        assert got[:size // 4] == b"\x03" * (size // 4) and got[size // 4:] == b"\x01" * (size - size // 4)
    assert client.GetCommStatus(pb.GetCommStatusRequest(commId=init.commId)).status == pb.SUCCESS


@pytest.mark.parametrize("n,nelem", [(2, 10), (3, 101770), (4, 7), (5, 1001)])
def test_ring_matches_numpy_oracle_f32(cluster, n, nelem):
    """True cross-device reduction incl. counts not divisible by n (the reference's padding case,
    407 080 B over 3 ranks, :297-334)."""
    client, devs, _ = cluster(n, mem_size=0x2000)
    init = client.CommInit(pb.CommInitRequest(numDevices=n, device_addresses=[d.addr for d in devs]))
    rng = np.random.default_rng(n)
    grads = [rng.standard_normal(nelem).astype(np.float32) for _ in range(n)]
    for md, g in zip(init.devices, grads):
        h2d(client, md.deviceId.value, 0x1000, g.tobytes())
    assert client.AllReduceRing(pb.AllReduceRingRequest(commId=init.commId, count=nelem * 4)).success
    ref = np.sum(np.stack(grads).astype(np.float64), axis=0)
    outs = [np.frombuffer(d2h(client, md.deviceId.value, 0x1000), dtype=np.float32) for md in init.devices]
    for o in outs:
        assert np.allclose(o, ref, rtol=1e-5, atol=1e-5)
        assert np.array_equal(o, outs[0])


def test_ring_honours_op_and_mem_addrs(cluster):
    client, devs, _ = cluster(3, mem_size=0x2000)
    init = client.CommInit(pb.CommInitRequest(numDevices=3, device_addresses=[d.addr for d in devs]))
    vals = [np.array([1, 9, -3, 4], np.int32), np.array([5, 2, 7, -8], np.int32), np.array([0, 3, 3, 3], np.int32)]
    for md, v in zip(init.devices, vals):
        h2d(client, md.deviceId.value, 0x1800, v.tobytes())
    req = pb.AllReduceRingRequest(commId=init.commId, count=16, op=pb.MAX)
    for r in range(3):
        req.memAddrs[r].value = 0x1800
    assert client.AllReduceRing(req, metadata=(("dmlp-dtype", "i32"),)).success
    for md in init.devices:
        assert np.array_equal(np.frombuffer(d2h(client, md.deviceId.value, 0x1800), np.int32), [5, 9, 7, 4])


def test_group_queues_collectives_until_group_end(cluster):
    client, devs, _ = cluster(2, mem_size=0x2000)
    init = client.CommInit(pb.CommInitRequest(numDevices=2, device_addresses=[d.addr for d in devs]))
    for md in init.devices:
        h2d(client, md.deviceId.value, 0x1000, np.ones(4, np.float32).tobytes())
    client.GroupStart(pb.GroupStartRequest(commId=init.commId))
    assert client.AllReduceRing(pb.AllReduceRingRequest(commId=init.commId, count=16)).success
    assert np.array_equal(np.frombuffer(d2h(client, init.devices[0].deviceId.value, 0x1000), np.float32), np.ones(4))
    client.GroupEnd(pb.GroupEndRequest(commId=init.commId))  # launches the queued all-reduce
    assert np.array_equal(np.frombuffer(d2h(client, init.devices[0].deviceId.value, 0x1000), np.float32), 2 * np.ones(4))
    assert client.CommFinalize(pb.CommFinalizeRequest(commId=init.commId)).success


def test_failure_mid_allreduce_is_surfaced(cluster):
    """A device that stops answering in the middle of a collective → the RPC fails and the
    communicator is FAILED (reference :341-344); later collectives get FailedPrecondition."""
    client, devs, _ = cluster(3, mem_size=0x2000, register_local=False, health_interval=30)
    init = client.CommInit(pb.CommInitRequest(numDevices=3, device_addresses=[d.addr for d in devs]))
    for md in init.devices:
        h2d(client, md.deviceId.value, 0x1000, np.ones(1024, np.float32).tobytes())
    devs[1].service.backend.dead = True  # fault injector: device stops responding
    with pytest.raises(grpc.RpcError) as e:
        client.AllReduceRing(pb.AllReduceRingRequest(commId=init.commId, count=4096))
    assert e.value.code() == grpc.StatusCode.UNKNOWN
    assert client.GetCommStatus(pb.GetCommStatusRequest(commId=init.commId)).status == pb.FAILED
    with pytest.raises(grpc.RpcError) as e:
        client.AllReduceRing(pb.AllReduceRingRequest(commId=init.commId, count=4096))
    assert e.value.code() == grpc.StatusCode.FAILED_PRECONDITION


def test_concurrent_rpcs_do_not_corrupt_state(cluster):
    """Stress: many clients create/destroy communicators and copy memory concurrently."""
    import threading
    client, devs, coord = cluster(2, mem_size=0x2000)
    errors = []

    def worker(k):
        try:
            ch = rpc.insecure_channel(coord.addr)
            c = rpc.CoordinatorStub(ch)
            for i in range(10):
                init = c.CommInit(pb.CommInitRequest(numDevices=2, device_addresses=[d.addr for d in devs]))
                payload = bytes([k]) * 64
                h2d(c, init.devices[k % 2].deviceId.value, 0x1000 + 16 * (k + 1), payload)
                assert d2h(c, init.devices[k % 2].deviceId.value, 0x1000 + 16 * (k + 1)) == payload
                c.CommDestroy(pb.CommDestroyRequest(commId=init.commId))
        except Exception as e:  # pragma: no cover
            errors.append(e)

    ts = [threading.Thread(target=worker, args=(k,)) for k in range(6)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert not errors


def test_grpc_client_trains_data_parallel_over_the_contract(cluster):
    """The reference client's loop with every number produced on the devices: Memcpy shards, RunForward/RunBackward,
    AllReduceRing with the fused average+SGD metadata. 3 sim devices; must equal the oracle's DP steps."""
    from dmlp_b200.grpc_client import GrpcTrainer
    from dmlp_b200.models import oracle
    from dmlp_b200.models.layout import ParamLayout
    from dmlp_b200.utils.data import synthetic_mnist
    dims, batch, lr, steps, n = (784, 128, 64, 10), 16, 0.1, 3, 3
    _, devs, coord = cluster(n, mem_size=1 << 20)
    t = GrpcTrainer(coord.addr, [d.addr for d in devs], dims=dims, batch=batch, lr=lr, seed=5)
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
        assert np.abs(w - rw).max() < 1e-5 and np.abs(b - rb).max() < 1e-5
    assert all(np.isfinite(s["loss"]) for s in stats) and 0.0 <= stats[-1]["accuracy"] <= 1.0
    t.close()


def test_symmetric_heap_placement_logic():
    """Host logic of the multicast heap behind AllReduceRing (`_heap_place` / `_allreduce_switch`) against stand-ins for
    the arena and the CUDA backends: one slot per address tuple at the same offset on every device, sized for the
    longest buffer, a new slot after growth, None when the heap is full, NVLS only for f32 SUM."""
    from types import SimpleNamespace
    from dmlp_b200.runtime.comm import ArenaAllocator
    from dmlp_b200.services import coordinator as co

    class FakeBackend:
        def __init__(self, size):
            self.sizes, self.placed = {0x1000: size}, {}

        def size(self, addr):
            return self.sizes[addr]

        def adopt(self, addr, ptr, cap):
            assert self.sizes[addr] <= cap
            self.placed[addr] = (ptr, cap)

    class FakeHeap:
        def __init__(self, nbytes):
            self.uc = {0: 0x10000000, 1: 0x20000000}
            self.alloc = ArenaAllocator(nbytes)
            self.calls = []

        def allreduce(self, comm, li, off_in, off_out, count, dev, **kw):
            self.calls.append((li, off_in, off_out, count, dev, kw))

    bes = [FakeBackend(300_000), FakeBackend(100_000)]
    devices = [SimpleNamespace(rank=r, local=SimpleNamespace(backend=b)) for r, b in enumerate(bes)]
    svc = co.GPUCoordinatorService(start_monitor=False)
    comm = co.Communicator(7, devices)
    comm.native = SimpleNamespace(local_devs=[0, 1], emulated=False, pick_algo=lambda nbytes, nvls: 4 if nbytes >= 65536 else 1)
    comm.heap = heap = FakeHeap(1 << 20)
    addrs = {0: 0x1000, 1: 0x1000}
    off = svc._heap_place(comm, heap, devices, addrs, 100_000)
    assert off == 0 and comm.heap_slots[(0x1000, 0x1000)][1] >= 300_000  # the longest buffer decides
    assert bes[0].placed[0x1000][0] - heap.uc[0] == bes[1].placed[0x1000][0] - heap.uc[1] == off
    assert svc._heap_place(comm, heap, devices, addrs, 100_000) == off  # same slot again
    bes[1].sizes[0x1000] = 500_000  # rewritten with more data: a new slot
    off2 = svc._heap_place(comm, heap, devices, addrs, 100_000)
    assert off2 > off and comm.heap_slots[(0x1000, 0x1000)][1] >= 500_000
    bes[0].sizes[0x1000] = 2 << 20
    assert svc._heap_place(comm, heap, devices, addrs, 100_000) is None  # heap exhausted: caller takes the unicast kernels
    bes[0].sizes[0x1000], bes[1].sizes[0x1000] = 300_000, 300_000
    from dmlp_b200.runtime import native
    # dispatch: f32 SUM + table says NVLS -> one launch per local device on the heap, in place
    assert svc._allreduce_switch(comm, devices, addrs, 100_000, "f32", pb.SUM, "auto", {}, native, None, None)
    assert [c[0] for c in heap.calls] == [0, 1] and all(c[1] == c[2] and c[3] == 25_000 for c in heap.calls)
    assert not svc._allreduce_switch(comm, devices, addrs, 1024, "f32", pb.SUM, "auto", {}, native, None, None)  # table: unicast
    assert not svc._allreduce_switch(comm, devices, addrs, 100_000, "i32", pb.SUM, "nvls", {}, native, None, None)
    assert not svc._allreduce_switch(comm, devices, addrs, 100_000, "f32", pb.MAX, "nvls", {}, native, None, None)
    comm.heap = False
    with pytest.raises(RuntimeError):
        svc._allreduce_switch(comm, devices, addrs, 100_000, "f32", pb.SUM, "nvls", {}, native, None, None)
    assert not svc._allreduce_switch(comm, devices, addrs, 100_000, "f32", pb.SUM, "auto", {}, native, None, None)
