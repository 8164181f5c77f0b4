"""``python -m dmlp_b200.serve_devices`` — launches GPUDevice servers.

Defaults reproduce the reference launcher (``DSML/cmd/gpu_device_server/main.go:13-23``): device ids
1,2,3 on ports 5003,5004,5005 with memSize 0x3000. ``--backend cuda`` binds each server to a GPU
(ordinals from ``--cuda-devices``; repeating an ordinal serves several logical devices from one GPU).
"""
import argparse
import logging
import signal
import threading

from .services import server as srv


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("--backend", choices=("sim", "cuda"), default="sim")
    ap.add_argument("--ports", default=",".join(map(str, srv.DEFAULT_DEVICE_PORTS)))
    ap.add_argument("--device-ids", default=",".join(map(str, srv.DEFAULT_DEVICE_IDS)))
    ap.add_argument("--mem-size", default=hex(srv.DEFAULT_MEM_SIZE))
    ap.add_argument("--cuda-devices", default="", help="CUDA ordinal per server (default 0,1,2,...)")
    ap.add_argument("--host", default="127.0.0.1")
    ap.add_argument("--metrics-port", type=int, default=0, help="serve Prometheus metrics (RPC counts / latencies); 0 = off")
    a = ap.parse_args(argv)
    logging.basicConfig(level=logging.INFO, format="%(asctime)s %(message)s")
    ports, ids = srv.parse_int_list(a.ports), srv.parse_int_list(a.device_ids)
    ordinals = srv.parse_int_list(a.cuda_devices) or None
    icpt = None
    if a.metrics_port:
        from .services import observability as obs
        registry = obs.new_registry()
        icpt = [obs.MetricsInterceptor("GPUDevice", registry)]
        obs.start_http_exporter(registry, a.metrics_port, a.host)
    devs = srv.start_devices(a.backend, ports, ids, int(a.mem_size, 0), ordinals, a.host, interceptors=icpt)
    stop = threading.Event()
    for sig in (signal.SIGINT, signal.SIGTERM):
        signal.signal(sig, lambda *_: stop.set())
    stop.wait()
    for d in devs:
        d.stop()
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
