"""gRPC plumbing for the runtime-built ``gpu_sim`` contract: generic server handlers and client stubs.

Server side: ``add_service(server, "GPUDevice", servicer)`` registers a handler for every method of
the service; a method the servicer does not define answers ``UNIMPLEMENTED`` (what the reference's
generated ``Unimplemented*Server`` embedding does, ``DSML/proto/gpu_sim_grpc.pb.go:187-192,560``).
Client side: ``DeviceStub(channel)`` / ``CoordinatorStub(channel)`` expose one callable per method.
"""
from __future__ import annotations

from typing import Optional

import grpc

from . import gpu_sim as pb

# Wide-model weights exceed gRPC's 4 MiB default; both ends raise the cap.
MAX_MESSAGE_BYTES = 1 << 30
CHANNEL_OPTIONS = [
    ("grpc.max_send_message_length", MAX_MESSAGE_BYTES),
    ("grpc.max_receive_message_length", MAX_MESSAGE_BYTES),
]


def _unimplemented(name):
    def handler(request, context):
        context.abort(grpc.StatusCode.UNIMPLEMENTED, f"method {name} not implemented")
    return handler


def add_service(server: grpc.Server, service: str, servicer) -> None:
    handlers = {}
    for method, req, resp, client_streaming in pb.SERVICES[service]:
        fn = getattr(servicer, method, None) or _unimplemented(method)
        req_cls, resp_cls = pb.message_class(req), pb.message_class(resp)
        maker = grpc.stream_unary_rpc_method_handler if client_streaming else grpc.unary_unary_rpc_method_handler
        handlers[method] = maker(fn, request_deserializer=req_cls.FromString,
                                 response_serializer=resp_cls.SerializeToString)
    server.add_generic_rpc_handlers((grpc.method_handlers_generic_handler(f"{pb.PACKAGE}.{service}", handlers),))


class _Stub:
    _service = ""

    def __init__(self, channel: grpc.Channel):
        self.channel = channel
        for method, req, resp, client_streaming in pb.SERVICES[self._service]:
            req_cls, resp_cls = pb.message_class(req), pb.message_class(resp)
            maker = channel.stream_unary if client_streaming else channel.unary_unary
            setattr(self, method, maker(pb.full_method(self._service, method),
                                        request_serializer=req_cls.SerializeToString,
                                        response_deserializer=resp_cls.FromString))


class DeviceStub(_Stub):
    _service = "GPUDevice"


class CoordinatorStub(_Stub):
    _service = "GPUCoordinator"


def insecure_channel(addr: str) -> grpc.Channel:
    """Insecure channel like the reference's ``grpc.Dial(..., insecure.NewCredentials())``
    (``DSML/gpu_coordinator_service/gpu_coordinator_server.go:137``)."""
    return grpc.insecure_channel(addr, options=CHANNEL_OPTIONS)


def make_server(max_workers: int = 16, interceptors: Optional[list] = None) -> grpc.Server:
    from concurrent import futures
    return grpc.server(futures.ThreadPoolExecutor(max_workers=max_workers), options=CHANNEL_OPTIONS,
                       interceptors=interceptors or None)
