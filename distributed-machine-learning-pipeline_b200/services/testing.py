"""Test doubles for calling servicers directly (no network), like the reference's
``mockStreamSendServer`` (DSML/gpu_device_service/gpu_device_server_test.go:14-44)."""
import grpc


class RpcAbort(Exception):
    def __init__(self, code: grpc.StatusCode, details: str):
        super().__init__(f"{code.name}: {details}")
        self.code, self.details = code, details


class FakeContext:
    """Minimal grpc.ServicerContext: ``abort`` raises RpcAbort, metadata is settable."""

    def __init__(self, metadata=()):
        self._md = tuple(metadata)
        self.trailing = ()

    def abort(self, code, details):
        raise RpcAbort(code, details)

    def invocation_metadata(self):
        return self._md

    def set_trailing_metadata(self, md):
        self.trailing = tuple(md)
