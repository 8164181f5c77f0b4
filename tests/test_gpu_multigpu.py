"""Real multi-GPU tests (need >= 2 CUDA devices; skipped otherwise).

* single process driving several GPUs through peer access (the coordinator's deployment)
* one process per GPU under torch.distributed.run with CUDA-IPC mapped segments (the trainer's)
"""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


needs2 = pytest.mark.skipif(_ngpu() < 2, reason="needs at least two GPUs")


@needs2
@pytest.mark.parametrize("algo", ["oneshot", "twoshot", "ring"])
def test_single_process_peer_allreduce(algo):
    from dmlp_b200.runtime import native
    from dmlp_b200.runtime.comm import Communicator, peer_matrix
    n = min(_ngpu(), 8)
    devs = list(range(n))
    topo = peer_matrix(devs)
    assert all(all(v == 1 for v in row) for row in topo["access"]), topo
    comm = Communicator.local(devs, staging_bytes=4 << 20)
    count = 1 << 20
    ins = [torch.randn(count, device=f"cuda:{d}") for d in devs]
    outs = [torch.empty(count, device=f"cuda:{d}") for d in devs]
    ref = sum(t.double().cpu() for t in ins)
    for d in devs:
        torch.cuda.synchronize(d)
    comm.allreduce([t.data_ptr() for t in ins], [t.data_ptr() for t in outs], count, algo=native.ALGO_BY_NAME[algo])
    comm.sync()
    comm.check()
    for o in outs:
        assert (o.double().cpu() - ref).abs().max().item() < 1e-4
        assert torch.equal(o.cpu(), outs[0].cpu())
    comm.destroy()


@needs2
def test_single_process_dead_gpu_rank_times_out():
    from dmlp_b200.runtime import native
    from dmlp_b200.runtime.comm import Communicator
    devs = [0, 1]
    comm = Communicator.local(devs)
    comm.set_timeout_ms(300)
    bufs = [torch.ones(4096, device=f"cuda:{d}") for d in devs]
    comm.inject_dead_ranks(0b10)
    comm.allreduce([t.data_ptr() for t in bufs], [t.data_ptr() for t in bufs], 4096)
    comm.sync()
    st = comm.status()
    assert st["code"] == native.ERR_TIMEOUT and st["rank"] == 0 and st["peer"] == 1
    comm.destroy()


@needs2
def test_torchrun_ipc_collectives_and_dp_trainer():
    n = min(_ngpu(), 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
           "--master-addr", "127.0.0.1", "--master-port", "29617", os.path.join(ROOT, "tests", "mp_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0 and "MP_WORKER_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


@needs2
def test_torchrun_global_batch_mode_oracle_soak_and_rank_failure():
    """The default data-parallel path of the fused step on every GPU of the box (2..8 ranks): oracle equality, replica
    bit-equality over a long run, and a rank that dies mid-run (watchdog -> TIMEOUT -> FAILED trainer)."""
    n = min(_ngpu(), 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
           "--master-addr", "127.0.0.1", "--master-port", "29619", os.path.join(ROOT, "tests", "mp_gx_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert r.returncode == 0 and "MP_GX_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


@needs2
def test_single_process_nvls_allreduce():
    """In-switch reduction through the multicast address (skipped when the fabric has no multicast)."""
    from dmlp_b200.runtime import native
    from dmlp_b200.runtime.comm import Communicator
    from dmlp_b200.runtime.nvls import NvlsArena, NvlsUnavailable
    n = min(_ngpu(), 8)
    devs = list(range(n))
    count = 1 << 18
    try:
        arena = NvlsArena.local(devs, 3 * count * 4)
    except NvlsUnavailable as e:
        pytest.skip(f"NVLS unavailable: {e}")
    comm = Communicator.local(devs)
    o_in, o_out = arena.alloc.alloc(count * 4), arena.alloc.alloc(count * 4)
    from dmlp_b200.parallel.data_parallel import arena_tensor
    ins = [arena_tensor(arena.uc[d] + o_in, count, d) for d in devs]
    for t in ins:
        t.normal_()
    ref = sum(t.double().cpu() for t in ins)
    for d in devs:
        torch.cuda.synchronize(d)
    for li, d in enumerate(devs):
        arena.allreduce(comm, li, o_in, o_out, count, d)
    comm.sync()
    comm.check()
    outs = [arena_tensor(arena.uc[d] + o_out, count, d) for d in devs]
    for o in outs:
        assert (o.double().cpu() - ref).abs().max().item() < 1e-4
    arena.destroy()
    comm.destroy()
