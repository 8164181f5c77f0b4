"""Service metrics (services/observability.py): per-RPC counters / latencies by status code, communicator states, the
Prometheus exposition, over real gRPC on loopback with sim devices."""
import urllib.request

import grpc
import pytest

from dmlp_b200.proto import gpu_sim as pb
from dmlp_b200.proto import rpc
from dmlp_b200.services import observability as obs
from dmlp_b200.services import server as srv
from dmlp_b200.services.backend import make_backend


def test_rpc_metrics_and_communicator_states_are_exported():
    registry = obs.new_registry()
    ci, di = obs.MetricsInterceptor("GPUCoordinator", registry), obs.MetricsInterceptor("GPUDevice", registry)
    devs = [srv.start_device_server(make_backend("sim", i + 1, 0x2000, 0), interceptors=[di]) for i in range(2)]
    coord = srv.start_coordinator_server(interceptors=[ci], start_monitor=False)
    obs.watch_communicators(registry, coord.service)
    try:
        client = rpc.CoordinatorStub(rpc.insecure_channel(coord.addr))
        init = client.CommInit(pb.CommInitRequest(numDevices=2, device_addresses=[d.addr for d in devs]))
        for md in init.devices:
            client.Memcpy(pb.MemcpyRequest(hostToDevice=pb.MemcpyHostToDeviceRequest(
                hostSrcData=b"\x01" * 64, dstDeviceId=md.deviceId, dstMemAddr=pb.MemAddr(value=0x1000))))
        assert client.AllReduceRing(pb.AllReduceRingRequest(commId=init.commId, count=64),
                                    metadata=(("dmlp-dtype", "u8"),)).success
        with pytest.raises(grpc.RpcError) as e:
            client.GetCommStatus(pb.GetCommStatusRequest(commId=4242))
        assert e.value.code() == grpc.StatusCode.NOT_FOUND
        snap = ci.snapshot()
        assert snap["counts"]["CommInit:OK"] == 1 and snap["counts"]["Memcpy:OK"] == 2
        assert snap["counts"]["AllReduceRing:OK"] == 1 and snap["counts"]["GetCommStatus:NOT_FOUND"] == 1
        assert snap["in_flight"] == 0 and snap["seconds"]["AllReduceRing"] > 0
        dsnap = di.snapshot()
        assert dsnap["counts"].get("GetDeviceMetadata:OK", 0) >= 2  # CommInit asked every device who it is
        text = obs.render(registry)
        if text:  # prometheus_client present
            assert 'dmlp_rpc_total{code="OK",method="AllReduceRing",service="GPUCoordinator"} 1.0' in text
            assert 'dmlp_rpc_total{code="NOT_FOUND",method="GetCommStatus",service="GPUCoordinator"} 1.0' in text
            assert 'dmlp_rpc_seconds_count{method="Memcpy",service="GPUCoordinator"} 2.0' in text
            assert 'dmlp_communicators{status="SUCCESS"} 1.0' in text and 'dmlp_communicators{status="FAILED"} 0.0' in text
            assert 'service="GPUDevice"' in text
            started = obs.start_http_exporter(registry, 0)  # an ephemeral port: scrape it once
            if started is not None:
                httpd = started[0]
                body = urllib.request.urlopen(f"http://127.0.0.1:{httpd.server_port}/metrics", timeout=10).read().decode()
                assert "dmlp_rpc_total" in body
                httpd.shutdown()
    finally:
        coord.stop()
        for d in devs:
            d.stop()


def test_standard_grpc_health_protocol():
    """grpc.health.v1.Health/Check on both servers: SERVING while alive, NOT_SERVING for a device marked dead, NOT_FOUND
    for a service the server does not host, UNAVAILABLE once the server is gone."""
    from dmlp_b200.proto import health
    dev = srv.start_device_server(make_backend("sim", 1, 0x2000, 0))
    coord = srv.start_coordinator_server(start_monitor=False)
    try:
        dch, cch = rpc.insecure_channel(dev.addr), rpc.insecure_channel(coord.addr)
        assert health.check(dch) == health.SERVING and health.check(dch, "gpu_sim.GPUDevice") == health.SERVING
        assert health.check(cch) == health.SERVING and health.check(cch, "gpu_sim.GPUCoordinator") == health.SERVING
        with pytest.raises(grpc.RpcError) as e:
            health.check(cch, "gpu_sim.GPUDevice")
        assert e.value.code() == grpc.StatusCode.NOT_FOUND
        dev.service.backend.dead = True
        assert health.check(dch) == health.NOT_SERVING
        dev.service.backend.dead = False
        # the contract's own RPCs are untouched by the extra service
        assert rpc.DeviceStub(dch).GetDeviceMetadata(pb.GetDeviceMetadataRequest()).metadata.deviceId.value == 1
    finally:
        coord.stop()
        dev.stop()
    with pytest.raises(grpc.RpcError) as e:
        health.check(rpc.insecure_channel(dev.addr), timeout=1.0)
    assert e.value.code() in (grpc.StatusCode.UNAVAILABLE, grpc.StatusCode.DEADLINE_EXCEEDED)
