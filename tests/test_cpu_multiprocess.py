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
