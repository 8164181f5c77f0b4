"""Multi-process host-side logic on CPU (gloo, world size 2): descriptor passing and data-parallel semantics."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_gloo_two_ranks(tmp_path):
    blob = tmp_path / "handle.bin"
    blob.write_bytes(b"multicast handle stand-in")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29741",
                        os.path.join(ROOT, "tests", "gloo_worker.py"), str(blob)],
                       capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0 and "GLOO_WORKER_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-3000:]


def _free_ports(n):
    import socket
    socks, ports = [], []
    for _ in range(n):
        s = socket.socket()
        s.bind(("127.0.0.1", 0))
        socks.append(s)
        ports.append(s.getsockname()[1])
    for s in socks:
        s.close()
    return ports


def test_launchers_and_grpc_client_as_separate_processes(tmp_path):
    """The reference's deployment shape (cmd/gpu_coordinator_server, cmd/gpu_device_server, client): the coordinator +
    device daemon and the client are separate OS processes talking gRPC over loopback."""
    import time
    ports = _free_ports(4)
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    log = open(tmp_path / "coord.log", "w")
    server = subprocess.Popen([sys.executable, "-m", "dmlp_b200.serve_coordinator", "--port", str(ports[0]), "--with-devices",
                               "--backend", "sim", "--device-ports", ",".join(map(str, ports[1:])), "--device-ids", "1,2,3",
                               "--mem-size", str(1 << 20)], stdout=log, stderr=subprocess.STDOUT, env=env, cwd=ROOT)
    try:
        deadline = time.time() + 60
        while time.time() < deadline:
            if "coordinator listening" in open(tmp_path / "coord.log").read():
                break
            assert server.poll() is None, open(tmp_path / "coord.log").read()[-2000:]
            time.sleep(0.2)
        r = subprocess.run([sys.executable, "-m", "dmlp_b200.grpc_client", "--coordinator", f"127.0.0.1:{ports[0]}",
                            "--devices", ",".join(f"127.0.0.1:{p}" for p in ports[1:]), "--batch", "8", "--steps-per-epoch", "3",
                            "--epochs", "1", "--samples", "128", "--lr", "0.1"],
                           capture_output=True, text=True, timeout=120, env=env, cwd=ROOT)
        assert r.returncode == 0 and "Training complete." in r.stderr + r.stdout, (r.stdout + r.stderr)[-3000:]
        assert "Epoch 1 complete: Avg Loss:" in r.stderr + r.stdout
    finally:
        server.terminate()  # the exact process we started
        try:
            server.wait(timeout=10)
        except subprocess.TimeoutExpired:
            server.kill()
        log.close()
