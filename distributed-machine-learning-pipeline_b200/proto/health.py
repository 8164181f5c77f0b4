"""The standard gRPC health-checking protocol (``grpc.health.v1.Health/Check``), built at run time like ``gpu_sim``.

The reference has its own liveness probe (the coordinator calls ``GetDeviceMetadata`` every 5 s,
``gpu_coordinator_server.go:69-119``) and nothing a load balancer or ``grpc_health_probe`` could ask. Both servers here
also answer the standard protocol: ``Check{service: "" | "gpu_sim.GPUDevice" | "gpu_sim.GPUCoordinator"}`` ->
``SERVING`` / ``NOT_SERVING`` / ``NOT_FOUND`` for an unknown service name. ``Watch`` (server streaming) is not served.
"""
from __future__ import annotations

from typing import Callable, Iterable

import grpc
from google.protobuf import descriptor_pb2, descriptor_pool, message_factory

PACKAGE = "grpc.health.v1"
UNKNOWN, SERVING, NOT_SERVING, SERVICE_UNKNOWN = 0, 1, 2, 3


def _build() -> descriptor_pb2.FileDescriptorProto:
    f = descriptor_pb2.FileDescriptorProto()
    f.name, f.package, f.syntax = "dmlp_b200/health.proto", PACKAGE, "proto3"
    req = f.message_type.add()
    req.name = "HealthCheckRequest"
    fd = req.field.add()
    fd.name, fd.number, fd.label, fd.type, fd.json_name = "service", 1, fd.LABEL_OPTIONAL, fd.TYPE_STRING, "service"
    resp = f.message_type.add()
    resp.name = "HealthCheckResponse"
    e = resp.enum_type.add()
    e.name = "ServingStatus"
    for name, num in (("UNKNOWN", 0), ("SERVING", 1), ("NOT_SERVING", 2), ("SERVICE_UNKNOWN", 3)):
        v = e.value.add()
        v.name, v.number = name, num
    fd = resp.field.add()
    fd.name, fd.number, fd.label, fd.type, fd.json_name = "status", 1, fd.LABEL_OPTIONAL, fd.TYPE_ENUM, "status"
    fd.type_name = f".{PACKAGE}.HealthCheckResponse.ServingStatus"
    s = f.service.add()
    s.name = "Health"
    m = s.method.add()
    m.name, m.input_type, m.output_type = "Check", f".{PACKAGE}.HealthCheckRequest", f".{PACKAGE}.HealthCheckResponse"
    return f


_pool = descriptor_pool.DescriptorPool()
_fd = _build()
_pool.Add(_fd)
HealthCheckRequest = message_factory.GetMessageClass(_pool.FindMessageTypeByName(f"{PACKAGE}.HealthCheckRequest"))
HealthCheckResponse = message_factory.GetMessageClass(_pool.FindMessageTypeByName(f"{PACKAGE}.HealthCheckResponse"))
CHECK = f"/{PACKAGE}.Health/Check"


def add_health_service(server: grpc.Server, services: Iterable[str], serving: Callable[[str], bool]) -> None:
    """Registers ``Health/Check``. ``services``: fully qualified names this server hosts ("" = the server as a whole is
    always accepted); ``serving(name)`` says whether that service should receive traffic right now."""
    known = set(services) | {""}

    def check(request, context):
        if request.service not in known:
            context.abort(grpc.StatusCode.NOT_FOUND, f"unknown service {request.service!r}")
        return HealthCheckResponse(status=SERVING if serving(request.service) else NOT_SERVING)

    handler = grpc.unary_unary_rpc_method_handler(check, request_deserializer=HealthCheckRequest.FromString,
                                                  response_serializer=HealthCheckResponse.SerializeToString)
    server.add_generic_rpc_handlers((grpc.method_handlers_generic_handler(f"{PACKAGE}.Health", {"Check": handler}),))


def check(channel: grpc.Channel, service: str = "", timeout: float = 2.0) -> int:
    """Client side: the serving status of ``service`` on the server behind ``channel``."""
    call = channel.unary_unary(CHECK, request_serializer=HealthCheckRequest.SerializeToString,
                               response_deserializer=HealthCheckResponse.FromString)
    return call(HealthCheckRequest(service=service), timeout=timeout).status
