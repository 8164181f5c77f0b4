"""Service metrics: per-RPC counters and latency histograms, communicator states, a Prometheus endpoint.

The reference logs with ``log.Printf`` and exposes nothing. For a deployment that serves many clients
the two servers can export

    dmlp_rpc_total{service, method, code}          RPCs handled, by gRPC status code
    dmlp_rpc_seconds{service, method}              handler latency histogram (50 us .. 10 s)
    dmlp_rpc_in_flight{service}                    RPCs being handled now
    dmlp_communicators{status}                     communicators by status (IN_PROGRESS / SUCCESS / FAILED)

through ``prometheus_client`` (``--metrics-port`` on ``serve_coordinator`` / ``serve_devices``, or
``MetricsInterceptor`` + ``start_http_exporter`` from code). Without ``prometheus_client`` the interceptor still counts
into plain dictionaries (``MetricsInterceptor.snapshot()``), so tests and ``--log-metrics`` work anywhere.
"""
from __future__ import annotations

import threading
import time
from collections import defaultdict
from typing import Callable, Dict, Optional, Tuple

import grpc

try:  # optional
    import prometheus_client as _prom
except Exception:  # pragma: no cover
    _prom = None

_BUCKETS = (50e-6, 100e-6, 250e-6, 500e-6, 1e-3, 2.5e-3, 5e-3, 10e-3, 25e-3, 50e-3, 0.1, 0.25, 0.5, 1.0, 2.5, 10.0)


_COLLECTORS: Dict[int, tuple] = {}
_COLLECTORS_MU = threading.Lock()


def _collectors(registry):
    """One set of series per registry, shared by every interceptor on it (the ``service`` label tells them apart)."""
    with _COLLECTORS_MU:
        c = _COLLECTORS.get(id(registry))
        if c is None:
            c = (_prom.Counter("dmlp_rpc_total", "RPCs handled", ["service", "method", "code"], registry=registry),
                 _prom.Histogram("dmlp_rpc_seconds", "RPC handler latency", ["service", "method"], buckets=_BUCKETS,
                                 registry=registry),
                 _prom.Gauge("dmlp_rpc_in_flight", "RPCs being handled", ["service"], registry=registry),
                 registry)  # (keeps the registry alive as long as its id is a key)
            _COLLECTORS[id(registry)] = c
        return c


class MetricsInterceptor(grpc.ServerInterceptor):
    """Counts and times every RPC of the server it is installed on (``rpc.make_server(interceptors=[...])``)."""

    def __init__(self, service: str, registry=None):
        self.service = service
        self._mu = threading.Lock()
        self.counts: Dict[Tuple[str, str], int] = defaultdict(int)      # (method, code) -> n
        self.seconds: Dict[str, float] = defaultdict(float)             # method -> total handler time
        self.in_flight = 0
        self.registry = registry
        self._p_total = self._p_secs = self._p_inflight = None
        if _prom is not None and registry is not None:
            self._p_total, self._p_secs, self._p_inflight, _ = _collectors(registry)

    # -- bookkeeping
    def _begin(self) -> None:
        with self._mu:
            self.in_flight += 1
        if self._p_inflight is not None:
            self._p_inflight.labels(self.service).inc()

    def _end(self, method: str, code: str, dt: float) -> None:
        with self._mu:
            self.in_flight -= 1
            self.counts[(method, code)] += 1
            self.seconds[method] += dt
        if self._p_total is not None:
            self._p_inflight.labels(self.service).dec()
            self._p_total.labels(self.service, method, code).inc()
            self._p_secs.labels(self.service, method).observe(dt)

    def snapshot(self) -> dict:
        with self._mu:
            return {"service": self.service, "in_flight": self.in_flight,
                    "counts": {f"{m}:{c}": n for (m, c), n in sorted(self.counts.items())},
                    "seconds": dict(sorted(self.seconds.items()))}

    # -- grpc.ServerInterceptor
    def intercept_service(self, continuation, handler_call_details):
        handler = continuation(handler_call_details)
        if handler is None:
            return None
        method = handler_call_details.method.rsplit("/", 1)[-1]

        def wrap(behavior: Callable) -> Callable:
            def timed(request_or_iterator, context):
                self._begin()
                t0 = time.perf_counter()
                code = "OK"
                try:
                    return behavior(request_or_iterator, context)
                except BaseException:  # context.abort raises after recording the status on the context
                    try:
                        c = context.code()
                        code = c.name if c is not None else "UNKNOWN"
                    except Exception:
                        code = "UNKNOWN"
                    raise
                finally:
                    self._end(method, code, time.perf_counter() - t0)
            return timed

        if handler.unary_unary is not None:
            return grpc.unary_unary_rpc_method_handler(wrap(handler.unary_unary),
                                                       request_deserializer=handler.request_deserializer,
                                                       response_serializer=handler.response_serializer)
        if handler.stream_unary is not None:
            return grpc.stream_unary_rpc_method_handler(wrap(handler.stream_unary),
                                                        request_deserializer=handler.request_deserializer,
                                                        response_serializer=handler.response_serializer)
        return handler  # (the contract has no server-streaming method)


def new_registry():
    """A private Prometheus registry (None without ``prometheus_client``)."""
    return _prom.CollectorRegistry() if _prom is not None else None


def watch_communicators(registry, coordinator_service) -> None:
    """Exports ``dmlp_communicators{status}`` from a GPUCoordinatorService, evaluated at scrape time."""
    if _prom is None or registry is None:
        return
    from ..proto import gpu_sim as pb
    names = {pb.IN_PROGRESS: "IN_PROGRESS", pb.SUCCESS: "SUCCESS", pb.FAILED: "FAILED"}
    g = _prom.Gauge("dmlp_communicators", "communicators by status", ["status"], registry=registry)

    def count(status):
        def f():
            with coordinator_service.mu:
                comms = list(coordinator_service.communicators.values())
            return float(sum(1 for c in comms if c.status == status))
        return f

    for st, name in names.items():
        g.labels(name).set_function(count(st))


def start_http_exporter(registry, port: int, addr: str = "127.0.0.1"):
    """Serves ``/metrics`` for ``registry``; returns (server, thread) of prometheus_client or None."""
    if _prom is None or registry is None:
        return None
    return _prom.start_http_server(port, addr=addr, registry=registry)


def render(registry) -> str:
    """The exposition text a scrape would return."""
    return _prom.generate_latest(registry).decode() if (_prom is not None and registry is not None) else ""
