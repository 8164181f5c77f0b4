"""``python -m dmlp_b200.serve_coordinator`` — launches the GPUCoordinator server.

Default port 50051 like ``DSML/cmd/gpu_coordinator_server/main.go:13``. With ``--with-devices`` the
device servers run in the same process (the B200 deployment: one daemon owns the box's GPUs), which
lets the coordinator drive the GPUs through the native peer-memory kernels instead of RPC payloads.
"""
import argparse
import logging
import signal
import threading

from .services import server as srv


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("--port", type=int, default=srv.DEFAULT_COORDINATOR_PORT)
    ap.add_argument("--host", default="127.0.0.1")
    ap.add_argument("--health-interval", type=float, default=5.0, help="seconds between liveness probes (reference: 5)")
    ap.add_argument("--probe-timeout", type=float, default=2.0, help="per-probe timeout in seconds (reference: 2)")
    ap.add_argument("--watchdog-ms", type=int, default=10000, help="in-kernel collective watchdog")
    ap.add_argument("--with-devices", action="store_true", help="also serve devices in this process")
    ap.add_argument("--backend", choices=("sim", "cuda"), default="sim")
    ap.add_argument("--device-ports", default=",".join(map(str, srv.DEFAULT_DEVICE_PORTS)))
    ap.add_argument("--device-ids", default=",".join(map(str, srv.DEFAULT_DEVICE_IDS)))
    ap.add_argument("--mem-size", default=hex(srv.DEFAULT_MEM_SIZE))
    ap.add_argument("--cuda-devices", default="")
    ap.add_argument("--metrics-port", type=int, default=0,
                    help="serve Prometheus metrics (RPC counts / latencies, communicator states) on this port; 0 = off")
    a = ap.parse_args(argv)
    logging.basicConfig(level=logging.INFO, format="%(asctime)s %(message)s")
    dev_icpt = coord_icpt = registry = None
    if a.metrics_port:
        from .services import observability as obs
        registry = obs.new_registry()
        coord_icpt = [obs.MetricsInterceptor("GPUCoordinator", registry)]
        dev_icpt = [obs.MetricsInterceptor("GPUDevice", registry)]
    devs = []
    if a.with_devices:
        devs = srv.start_devices(a.backend, srv.parse_int_list(a.device_ports), srv.parse_int_list(a.device_ids),
                                 int(a.mem_size, 0), srv.parse_int_list(a.cuda_devices) or None, a.host,
                                 interceptors=dev_icpt)
    coord = srv.start_coordinator_server(f"{a.host}:{a.port}", health_interval=a.health_interval,
                                         probe_timeout=a.probe_timeout, watchdog_ms=a.watchdog_ms,
                                         interceptors=coord_icpt)
    if a.metrics_port:
        obs.watch_communicators(registry, coord.service)
        obs.start_http_exporter(registry, a.metrics_port, a.host)
        logging.getLogger("dmlp.server").info("metrics on http://%s:%d/metrics", a.host, a.metrics_port)
    stop = threading.Event()
    for sig in (signal.SIGINT, signal.SIGTERM):
        signal.signal(sig, lambda *_: stop.set())
    stop.wait()
    coord.stop()
    for d in devs:
        d.stop()
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
