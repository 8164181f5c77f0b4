"""GPUDevice service: metadata, host<->device Memcpy and the point-to-point stream state machine.

Behavioural contract (reference ``DSML/gpu_device_service/gpu_device_server.go``):
  * ``GetDeviceMetadata``  → {deviceId, 0x1000, 0x1000+memSize}                      (:51-62)
  * ``BeginSend``          → new streamId (counter from 1), state IN_PROGRESS         (:64-88)
  * ``BeginReceive``       → attaches recv buffer to a known stream; NotFound for an unknown
                             id, InvalidArgument when the address is out of bounds    (:90-110)
  * ``StreamSend``         → concatenates chunks; length must equal numBytes and both Begin*
                             must have happened, else the stream FAILS                (:112-181)
  * ``GetStreamStatus``    → stream status; unknown id reports FAILED without error   (:183-193)
  * ``Memcpy``             → H2D stores a copy, D2H returns the buffer; OOB →
                             InvalidArgument, missing → Internal                      (:195-230)

What is new relative to the reference: with the ``cuda`` backend the bytes live in GPU memory, and a
``StreamSend`` that carries *no payload* executes the registered transfer device-to-device with the
native P2P copy kernel on the communicator's stream (the destination is the peer registered for
``dstRank``); ``GetStreamStatus`` then reflects the CUDA event of that copy. ``D2H numBytes`` is
honoured when non-zero (the reference ignores it).
"""
from __future__ import annotations

import logging
import threading
from dataclasses import dataclass, field
from typing import Callable, Dict, Iterable, Optional

import grpc

from ..proto import gpu_sim as pb
from ..utils.tracing import traced
from .backend import AddressError, DeviceBackend, MissingBufferError

log = logging.getLogger("dmlp.device")


@dataclass
class StreamState:
    send_addr: int = 0
    recv_addr: int = 0
    count: int = 0
    src_rank: int = 0
    dst_rank: int = 0
    status: int = pb.IN_PROGRESS
    initiated_send: bool = False
    initiated_recv: bool = False
    # native transfers: callable returning True once the copy's CUDA event has completed
    poll: Optional[Callable[[], bool]] = None
    data: bytearray = field(default_factory=bytearray)


class GPUDeviceService:
    """Servicer for ``gpu_sim.GPUDevice``. One instance per (simulated or real) GPU."""

    def __init__(self, backend: DeviceBackend):
        self.backend = backend
        self._mu = threading.Lock()
        self.streams: Dict[int, StreamState] = {}
        self._next_stream = 1
        # rank -> service of the communicator this device was last initialised into (in-process only)
        self.peers: Dict[int, "GPUDeviceService"] = {}
        self.rank: Optional[int] = None

    # ------------------------------------------------------------------ helpers
    @property
    def device_id(self) -> int:
        return self.backend.device_id

    def _guard_alive(self, context) -> None:
        if self.backend.dead:
            context.abort(grpc.StatusCode.UNAVAILABLE, f"device {self.device_id} is not responding (fault injected)")

    def set_peers(self, rank: int, peers: Dict[int, "GPUDeviceService"]) -> None:
        with self._mu:
            self.rank, self.peers = rank, dict(peers)

    # ------------------------------------------------------------------ RPCs
    def GetDeviceMetadata(self, request, context):
        self._guard_alive(context)
        md = pb.DeviceMetadata(deviceId=pb.DeviceId(value=self.backend.device_id),
                               minMemAddr=pb.MemAddr(value=self.backend.min_addr),
                               maxMemAddr=pb.MemAddr(value=self.backend.max_addr))
        try:  # extra facts travel as trailing metadata, not as new proto fields
            context.set_trailing_metadata(tuple((f"dmlp-{k}", v) for k, v in self.backend.info().items()))
        except Exception:  # direct (non-gRPC) callers may pass a minimal context
            pass
        return pb.GetDeviceMetadataResponse(metadata=md)

    def BeginSend(self, request, context):
        self._guard_alive(context)
        with self._mu:
            sid = self._next_stream
            self._next_stream += 1
            self.streams[sid] = StreamState(send_addr=request.sendBuffAddr.value, count=request.numBytes,
                                            dst_rank=request.dstRank.value, initiated_send=True)
        return pb.BeginSendResponse(initiated=True, streamId=pb.StreamId(value=sid))

    def BeginReceive(self, request, context):
        self._guard_alive(context)
        with self._mu:
            st = self.streams.get(request.streamId.value)
            if st is None:
                context.abort(grpc.StatusCode.NOT_FOUND, f"unknown streamId {request.streamId.value}")
            addr = request.recvBuffAddr.value
            try:
                self.backend.check_addr(addr)
            except AddressError as e:
                context.abort(grpc.StatusCode.INVALID_ARGUMENT, f"RecvBuffAddr {e}")
            st.recv_addr, st.src_rank, st.initiated_recv = addr, request.srcRank.value, True
        return pb.BeginReceiveResponse(initiated=True)

    @traced("dmlp.rpc.device.StreamSend")
    def StreamSend(self, request_iterator: Iterable, context):
        self._guard_alive(context)
        it = iter(request_iterator)
        try:
            first = next(it)
        except StopIteration:
            return pb.StreamSendResponse(success=False)
        sid = first.streamId
        if sid == 0:
            context.abort(grpc.StatusCode.INVALID_ARGUMENT, "no streamId in the first chunk")
        with self._mu:
            st = self.streams.get(sid)
            if st is None:
                return pb.StreamSendResponse(success=False)
            if not st.initiated_send:
                st.status = pb.FAILED
                log.warning("Stream %d: send not initiated", sid)
                return pb.StreamSendResponse(success=False)
            st.data.extend(first.data)
        try:
            for chunk in it:
                with self._mu:
                    st.data.extend(chunk.data)
        except Exception:
            with self._mu:
                st.status = pb.FAILED
            raise
        with self._mu:
            if len(st.data) == 0 and st.count > 0 and st.initiated_recv and self.peers:
                return self._native_transfer(sid, st)
            if len(st.data) != st.count:
                st.status = pb.FAILED
                log.warning("Stream %d: data size mismatch. Expected %d, got %d", sid, st.count, len(st.data))
                return pb.StreamSendResponse(success=False)
            if not st.initiated_recv:
                st.status = pb.FAILED
                log.warning("Stream %d: receive not initiated before data arrived", sid)
                return pb.StreamSendResponse(success=False)
            self.backend.write(st.recv_addr, bytes(st.data))
            st.data = bytearray()
            st.status = pb.SUCCESS
        return pb.StreamSendResponse(success=True)

    def _native_transfer(self, sid: int, st: StreamState):
        """Payload-free StreamSend: copy send buffer → peer[dstRank]'s recv buffer on the device(s)."""
        dst = self.peers.get(st.dst_rank)
        if dst is None:
            st.status = pb.FAILED
            return pb.StreamSendResponse(success=False)
        try:
            if hasattr(self.backend, "p2p_copy_to") and hasattr(dst.backend, "p2p_copy_to"):
                st.poll = self.backend.p2p_copy_to(dst.backend, st.send_addr, st.recv_addr, st.count)
            else:  # sim devices: a host copy is the whole transfer
                dst.backend.write(st.recv_addr, self.backend.read(st.send_addr, st.count))
                st.status = pb.SUCCESS
        except (MissingBufferError, AddressError) as e:
            log.warning("Stream %d: native transfer failed: %s", sid, e)
            st.status = pb.FAILED
            return pb.StreamSendResponse(success=False)
        return pb.StreamSendResponse(success=True)

    def GetStreamStatus(self, request, context):
        self._guard_alive(context)
        with self._mu:
            st = self.streams.get(request.streamId.value)
            if st is None:
                return pb.GetStreamStatusResponse(status=pb.FAILED)
            if st.poll is not None and st.status == pb.IN_PROGRESS:
                try:
                    if st.poll():
                        st.status, st.poll = pb.SUCCESS, None
                except Exception as e:  # CUDA error on the copy stream
                    log.error("stream %d failed: %s", request.streamId.value, e)
                    st.status, st.poll = pb.FAILED, None
            return pb.GetStreamStatusResponse(status=st.status)

    @traced("dmlp.rpc.device.Memcpy")
    def Memcpy(self, request, context):
        self._guard_alive(context)
        which = request.WhichOneof("either")
        if which == "hostToDevice":
            h2d = request.hostToDevice
            addr = h2d.dstMemAddr.value
            try:
                self.backend.check_addr(addr)
            except AddressError as e:
                context.abort(grpc.StatusCode.INVALID_ARGUMENT, f"HostToDevice: {e}")
            self.backend.write(addr, h2d.hostSrcData)
            return pb.MemcpyResponse(hostToDevice=pb.MemcpyHostToDeviceResponse(success=True))
        if which == "deviceToHost":
            d2h = request.deviceToHost
            addr = d2h.srcMemAddr.value
            try:
                self.backend.check_addr(addr)
            except AddressError as e:
                context.abort(grpc.StatusCode.INVALID_ARGUMENT, f"DeviceToHost: {e}")
            try:
                data = self.backend.read(addr, d2h.numBytes)
            except MissingBufferError:
                context.abort(grpc.StatusCode.INTERNAL, f"DeviceToHost: no data at address 0x{addr:x}")
            return pb.MemcpyResponse(deviceToHost=pb.MemcpyDeviceToHostResponse(dstData=data))
        context.abort(grpc.StatusCode.INVALID_ARGUMENT, "Invalid Memcpy request: specify hostToDevice or deviceToHost")

    # The reference's stale stubs declare these and no server implements them. Here they run the MLP on the
    # device that holds the buffers (semantics: services/compute.py); a backend without compute hooks
    # answers UNIMPLEMENTED like the reference.
    @traced("dmlp.rpc.device.RunForward")
    def RunForward(self, request, context):
        hook = getattr(self.backend, "run_forward", None)
        if hook is None:
            context.abort(grpc.StatusCode.UNIMPLEMENTED, "method RunForward not implemented")
        self._guard_alive(context)
        self._run_compute(context, "RunForward", hook, request.inputAddr, request.outputAddr)
        return pb.RunForwardResponse(success=True)

    @traced("dmlp.rpc.device.RunBackward")
    def RunBackward(self, request, context):
        hook = getattr(self.backend, "run_backward", None)
        if hook is None:
            context.abort(grpc.StatusCode.UNIMPLEMENTED, "method RunBackward not implemented")
        self._guard_alive(context)
        self._run_compute(context, "RunBackward", hook, request.gradientAddr)
        return pb.RunBackwardResponse(success=True)

    @staticmethod
    def _run_compute(context, what, hook, *addrs) -> None:
        """Same status codes as Memcpy: address outside the device range / malformed buffers -> InvalidArgument,
        a buffer that was never written -> Internal."""
        try:
            hook(*addrs)
        except (AddressError, ValueError) as e:
            context.abort(grpc.StatusCode.INVALID_ARGUMENT, f"{what}: {e}")
        except MissingBufferError as e:
            context.abort(grpc.StatusCode.INTERNAL, f"{what}: {e}")
