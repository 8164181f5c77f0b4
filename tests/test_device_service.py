"""Device-service unit tests: the reference's white-box suite
(DSML/gpu_device_service/gpu_device_server_test.go:46-164) re-expressed, plus bounds/error cases
the reference only has in code. Direct method calls, no network; backend = sim (CPU)."""
import grpc
import pytest

from dmlp_b200.proto import gpu_sim as pb
from dmlp_b200.services.backend import SimBackend
from dmlp_b200.services.device import GPUDeviceService, StreamState
from dmlp_b200.services.testing import FakeContext, RpcAbort


def make(device_id=1001, mem=1 << 20):
    return GPUDeviceService(SimBackend(device_id, mem))


def test_get_device_metadata():
    md = make().GetDeviceMetadata(pb.GetDeviceMetadataRequest(), FakeContext()).metadata
    assert (md.deviceId.value, md.minMemAddr.value, md.maxMemAddr.value) == (1001, 0x1000, 0x101000)


def test_begin_send_allocates_nonzero_stream_ids():
    svc = make()
    r1 = svc.BeginSend(pb.BeginSendRequest(sendBuffAddr=pb.MemAddr(value=0x1000), numBytes=1024,
                                           dstRank=pb.Rank(value=2)), FakeContext())
    r2 = svc.BeginSend(pb.BeginSendRequest(numBytes=1), FakeContext())
    assert r1.initiated and r1.streamId.value == 1 and r2.streamId.value == 2


def test_begin_receive_on_seeded_stream():
    svc = make()
    svc.streams[1] = StreamState(status=pb.IN_PROGRESS)
    r = svc.BeginReceive(pb.BeginReceiveRequest(streamId=pb.StreamId(value=1), recvBuffAddr=pb.MemAddr(value=0x2000),
                                                numBytes=1024, srcRank=pb.Rank(value=1)), FakeContext())
    assert r.initiated and svc.streams[1].recv_addr == 0x2000 and svc.streams[1].initiated_recv


def test_begin_receive_errors():
    svc = make(mem=0x1000)
    with pytest.raises(RpcAbort) as e:
        svc.BeginReceive(pb.BeginReceiveRequest(streamId=pb.StreamId(value=9)), FakeContext())
    assert e.value.code == grpc.StatusCode.NOT_FOUND
    svc.streams[1] = StreamState()
    with pytest.raises(RpcAbort) as e:  # 0x2000 == maxAddr is out of bounds for memSize 0x1000
        svc.BeginReceive(pb.BeginReceiveRequest(streamId=pb.StreamId(value=1), recvBuffAddr=pb.MemAddr(value=0x2000)),
                         FakeContext())
    assert e.value.code == grpc.StatusCode.INVALID_ARGUMENT


def test_stream_send_concatenates_chunks_in_order():
    svc = make()
    svc.streams[1] = StreamState(send_addr=0x1000, recv_addr=0x2000, count=12, status=pb.IN_PROGRESS,
                                 initiated_send=True, initiated_recv=True)
    chunks = [pb.DataChunk(data=b"chunk1", streamId=1), pb.DataChunk(data=b"chunk2", streamId=1)]
    assert svc.StreamSend(iter(chunks), FakeContext()).success
    assert svc.backend.read(0x2000) == b"chunk1chunk2"
    assert svc.streams[1].status == pb.SUCCESS


def test_stream_send_failure_modes():
    svc = make()
    assert not svc.StreamSend(iter([]), FakeContext()).success  # no chunks at all
    with pytest.raises(RpcAbort) as e:
        svc.StreamSend(iter([pb.DataChunk(data=b"x")]), FakeContext())  # no stream id
    assert e.value.code == grpc.StatusCode.INVALID_ARGUMENT
    assert not svc.StreamSend(iter([pb.DataChunk(data=b"x", streamId=77)]), FakeContext()).success  # unknown
    svc.streams[2] = StreamState(count=5, initiated_send=True, initiated_recv=True, recv_addr=0x2000)
    assert not svc.StreamSend(iter([pb.DataChunk(data=b"abc", streamId=2)]), FakeContext()).success  # size mismatch
    assert svc.streams[2].status == pb.FAILED
    svc.streams[3] = StreamState(count=3, initiated_send=True, initiated_recv=False)
    assert not svc.StreamSend(iter([pb.DataChunk(data=b"abc", streamId=3)]), FakeContext()).success  # no receive
    svc.streams[4] = StreamState(count=3, initiated_send=False)
    assert not svc.StreamSend(iter([pb.DataChunk(data=b"abc", streamId=4)]), FakeContext()).success  # no send


def test_get_stream_status():
    svc = make()
    svc.streams[1] = StreamState(status=pb.IN_PROGRESS)
    assert svc.GetStreamStatus(pb.GetStreamStatusRequest(streamId=pb.StreamId(value=1)), FakeContext()).status == pb.IN_PROGRESS
    assert svc.GetStreamStatus(pb.GetStreamStatusRequest(streamId=pb.StreamId(value=5)), FakeContext()).status == pb.FAILED


def test_memcpy_roundtrip_and_errors():
    svc = make(mem=0x1000)
    h2d = pb.MemcpyRequest(hostToDevice=pb.MemcpyHostToDeviceRequest(hostSrcData=b"Hello GPU!", dstMemAddr=pb.MemAddr(value=0x1000)))
    assert svc.Memcpy(h2d, FakeContext()).hostToDevice.success
    d2h = pb.MemcpyRequest(deviceToHost=pb.MemcpyDeviceToHostRequest(srcMemAddr=pb.MemAddr(value=0x1000)))
    assert svc.Memcpy(d2h, FakeContext()).deviceToHost.dstData == b"Hello GPU!"
    part = pb.MemcpyRequest(deviceToHost=pb.MemcpyDeviceToHostRequest(srcMemAddr=pb.MemAddr(value=0x1000), numBytes=5))
    assert svc.Memcpy(part, FakeContext()).deviceToHost.dstData == b"Hello"
    for req, code in [
        (pb.MemcpyRequest(hostToDevice=pb.MemcpyHostToDeviceRequest(dstMemAddr=pb.MemAddr(value=0x2000))), grpc.StatusCode.INVALID_ARGUMENT),
        (pb.MemcpyRequest(deviceToHost=pb.MemcpyDeviceToHostRequest(srcMemAddr=pb.MemAddr(value=0x800))), grpc.StatusCode.INVALID_ARGUMENT),
        (pb.MemcpyRequest(deviceToHost=pb.MemcpyDeviceToHostRequest(srcMemAddr=pb.MemAddr(value=0x1800))), grpc.StatusCode.INTERNAL),
        (pb.MemcpyRequest(), grpc.StatusCode.INVALID_ARGUMENT),
    ]:
        with pytest.raises(RpcAbort) as e:
            svc.Memcpy(req, FakeContext())
        assert e.value.code == code


def test_payload_free_stream_send_is_a_device_to_device_copy():
    a, b = make(1), make(2)
    peers = {0: a, 1: b}
    a.set_peers(0, peers)
    b.set_peers(1, peers)
    a.backend.write(0x1000, b"0123456789")
    s = a.BeginSend(pb.BeginSendRequest(sendBuffAddr=pb.MemAddr(value=0x1000), numBytes=6, dstRank=pb.Rank(value=1)), FakeContext())
    a.BeginReceive(pb.BeginReceiveRequest(streamId=s.streamId, recvBuffAddr=pb.MemAddr(value=0x2000), numBytes=6,
                                          srcRank=pb.Rank(value=0)), FakeContext())
    assert a.StreamSend(iter([pb.DataChunk(streamId=s.streamId.value)]), FakeContext()).success
    assert a.GetStreamStatus(pb.GetStreamStatusRequest(streamId=s.streamId), FakeContext()).status == pb.SUCCESS
    assert b.backend.read(0x2000) == b"012345"


def test_backend_without_compute_hooks_answers_unimplemented_like_the_reference_stub():
    from dmlp_b200.services.backend import DeviceBackend

    class Bare(DeviceBackend):
        kind = "bare"

    with pytest.raises(RpcAbort) as e:
        GPUDeviceService(Bare(1, 1 << 20)).RunForward(pb.RunForwardRequest(deviceId=1), FakeContext())
    assert e.value.code == grpc.StatusCode.UNIMPLEMENTED


@pytest.mark.parametrize("dims,onehot", [((784, 128, 10), True), ((784, 128, 64, 10), False)])
def test_run_forward_and_backward_compute_on_the_device(dims, onehot):
    """RunForward / RunBackward (stale stub RPCs of the reference) with the client's memory map: weights blob at
    0x2000, data at 0x3000, labels at 0x4000, gradients to 0x1000 (services/compute.py)."""
    import numpy as np
    from dmlp_b200.models import oracle
    from dmlp_b200.models.layout import ParamLayout
    from dmlp_b200.utils.data import synthetic_mnist
    svc = make(mem=1 << 20)
    ctx = FakeContext()
    lay = ParamLayout(dims)
    flat = lay.init_reference(11)
    x, y = synthetic_mnist(24, seed=4)

    def h2d(addr, data):
        svc.Memcpy(pb.MemcpyRequest(hostToDevice=pb.MemcpyHostToDeviceRequest(
            hostSrcData=data, dstDeviceId=pb.DeviceId(value=1001), dstMemAddr=pb.MemAddr(value=addr))), ctx)

    def d2h(addr):
        return svc.Memcpy(pb.MemcpyRequest(deviceToHost=pb.MemcpyDeviceToHostRequest(
            srcDeviceId=pb.DeviceId(value=1001), srcMemAddr=pb.MemAddr(value=addr))), ctx).deviceToHost.dstData

    h2d(0x2000, lay.to_reference_bytes(flat))
    h2d(0x3000, x.astype("<f4").tobytes())
    h2d(0x4000, np.eye(10, dtype="<f4")[y].tobytes() if onehot else y.astype("<i4").tobytes())
    assert svc.RunForward(pb.RunForwardRequest(deviceId=1001, inputAddr=0x3000, outputAddr=0x5000), ctx).success
    params = [(w.astype(np.float64), b.astype(np.float64)) for w, b in lay.views(flat)]
    logits, _ = oracle.forward(params, x.astype(np.float64))
    got = np.frombuffer(d2h(0x5000), "<f4").reshape(24, 10)
    assert np.abs(got - logits).max() < 1e-5
    assert svc.RunBackward(pb.RunBackwardRequest(deviceId=1001, gradientAddr=0x1000), ctx).success
    _, _, grads = oracle.loss_and_grads(params, x.astype(np.float64), y)
    blob = d2h(0x1000)
    assert len(blob) == lay.reference_count * 4  # same size as the weight blob: ready for AllReduceRing
    gflat = lay.from_reference_bytes(blob)
    for (gw, gb), (rw, rb) in zip(lay.views(gflat), grads):
        assert np.abs(gw - rw).max() < 1e-6 and np.abs(gb - rb).max() < 1e-6


def test_run_forward_error_codes():
    svc = make(mem=1 << 20)
    ctx = FakeContext()
    with pytest.raises(RpcAbort) as e:  # no weights yet
        svc.RunForward(pb.RunForwardRequest(deviceId=1001, inputAddr=0x3000, outputAddr=0x5000), ctx)
    assert e.value.code == grpc.StatusCode.INTERNAL
    svc.backend.write(0x2000, b"\x00" * 100)  # not a known model shape
    svc.backend.write(0x3000, b"\x00" * 784 * 4)
    with pytest.raises(RpcAbort) as e:
        svc.RunForward(pb.RunForwardRequest(deviceId=1001, inputAddr=0x3000, outputAddr=0x5000), ctx)
    assert e.value.code == grpc.StatusCode.INVALID_ARGUMENT
    with pytest.raises(RpcAbort) as e:  # backward before any forward
        svc.RunBackward(pb.RunBackwardRequest(deviceId=1001, gradientAddr=0x1000), ctx)
    assert e.value.code == grpc.StatusCode.INVALID_ARGUMENT
