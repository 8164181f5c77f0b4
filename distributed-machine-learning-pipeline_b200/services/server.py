"""Server bring-up helpers and the process launchers' shared code.

Reference launchers: ``DSML/cmd/gpu_device_server/main.go`` (three devices, ids 1/2/3 on ports
5003/5004/5005, memSize 0x3000) and ``DSML/cmd/gpu_coordinator_server/main.go`` (port 50051).
The same defaults are kept, but everything is a flag (the reference has no config).
"""
from __future__ import annotations

import logging
from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import grpc

from ..proto import health, rpc
from .backend import DeviceBackend, make_backend
from .coordinator import GPUCoordinatorService, register_local_device, unregister_local_device
from .device import GPUDeviceService

log = logging.getLogger("dmlp.server")

DEFAULT_COORDINATOR_PORT = 50051
DEFAULT_DEVICE_PORTS = (5003, 5004, 5005)
DEFAULT_DEVICE_IDS = (1, 2, 3)
DEFAULT_MEM_SIZE = 0x3000


@dataclass
class RunningDevice:
    server: grpc.Server
    service: GPUDeviceService
    addr: str

    def stop(self, grace: Optional[float] = 0) -> None:
        """Stops answering RPCs — what the reference's failure test does to a device
        (``grpcServer.Stop()``, gpu_coordinator_server_test.go:415)."""
        self.server.stop(grace)
        unregister_local_device(self.addr)


@dataclass
class RunningCoordinator:
    server: grpc.Server
    service: GPUCoordinatorService
    addr: str

    def stop(self, grace: Optional[float] = 0) -> None:
        self.server.stop(grace)
        self.service.Stop()


def start_device_server(backend: DeviceBackend, addr: str = "127.0.0.1:0", register_local: bool = True,
                        max_workers: int = 8, interceptors: Optional[list] = None) -> RunningDevice:
    svc = GPUDeviceService(backend)
    server = rpc.make_server(max_workers=max_workers, interceptors=interceptors)
    rpc.add_service(server, "GPUDevice", svc)
    # standard health protocol: a device that was marked dead (fault injection / lost GPU) reports NOT_SERVING
    health.add_health_service(server, ["gpu_sim.GPUDevice"], lambda name: not svc.backend.dead)
    host = addr.rsplit(":", 1)[0]
    port = server.add_insecure_port(addr)
    if port == 0:
        raise RuntimeError(f"could not bind device server to {addr}")
    server.start()
    bound = f"{host}:{port}"
    if register_local:
        register_local_device(bound, svc)
    log.info("GPU device %d (%s) listening on %s", backend.device_id, backend.kind, bound)
    return RunningDevice(server, svc, bound)


def start_coordinator_server(addr: str = "127.0.0.1:0", service: Optional[GPUCoordinatorService] = None,
                             max_workers: int = 16, interceptors: Optional[list] = None,
                             **service_kwargs) -> RunningCoordinator:
    svc = service or GPUCoordinatorService(**service_kwargs)
    server = rpc.make_server(max_workers=max_workers, interceptors=interceptors)
    rpc.add_service(server, "GPUCoordinator", svc)
    health.add_health_service(server, ["gpu_sim.GPUCoordinator"], lambda name: not svc._stop.is_set())
    host = addr.rsplit(":", 1)[0]
    port = server.add_insecure_port(addr)
    if port == 0:
        raise RuntimeError(f"could not bind coordinator to {addr}")
    server.start()
    log.info("GPU coordinator listening on %s:%d", host, port)
    return RunningCoordinator(server, svc, f"{host}:{port}")


def start_devices(backend: str, ports: Sequence[int], device_ids: Sequence[int], mem_size: int,
                  cuda_ordinals: Optional[Sequence[int]] = None, host: str = "127.0.0.1",
                  interceptors: Optional[list] = None) -> List[RunningDevice]:
    out = []
    for i, (port, did) in enumerate(zip(ports, device_ids)):
        ordinal = None
        if backend == "cuda":
            ordinal = cuda_ordinals[i] if cuda_ordinals else i
        out.append(start_device_server(make_backend(backend, did, mem_size, ordinal), f"{host}:{port}",
                                       interceptors=interceptors))
    return out


def parse_int_list(text: str) -> Tuple[int, ...]:
    return tuple(int(t, 0) for t in text.split(",") if t.strip())
