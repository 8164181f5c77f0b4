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
def test_torchrun_wgrad_reduce_scatter_pipeline():
    """wgrad GEMM with the reduce-scatter epilogue + its consumer kernel against the unfused GEMM + bucket all-reduce,
    both precisions, graph and eager; then a rank that dies mid-run."""
    n = min(_ngpu(), 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}",
           "--master-addr", "127.0.0.1", "--master-port", "29623", os.path.join(ROOT, "tests", "mp_rs_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0 and "MP_RS_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


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


@needs2
def test_grpc_allreduce_runs_on_the_symmetric_heap():
    """Devices on distinct GPUs behind the gRPC contract: an f32 SUM AllReduceRing moves the buffers into the
    communicator's multicast-bound heap (same offset on every GPU) and reduces in the switch; Memcpy, the fused SGD
    update, the unicast kernels and a larger rewrite of the same address keep working on the moved buffers."""
    import numpy as np
    from dmlp_b200.proto import gpu_sim as pb
    from dmlp_b200.proto import rpc
    from dmlp_b200.runtime.nvls import supported
    from dmlp_b200.services import server as srv
    from dmlp_b200.services.backend import make_backend
    n = min(_ngpu(), 8)
    if not all(supported(d) for d in range(n)):
        pytest.skip("multicast objects are not supported on this box")
    devs = [srv.start_device_server(make_backend("cuda", i + 1, 0x2000, i)) for i in range(n)]
    coord = srv.start_coordinator_server()
    try:
        client = rpc.CoordinatorStub(rpc.insecure_channel(coord.addr))
        init = client.CommInit(pb.CommInitRequest(numDevices=n, device_addresses=[d.addr for d in devs]))
        comm = coord.service.communicators[init.commId]
        assert comm.native is not None and not comm.native.emulated

        def h2d(dev, addr, data):
            assert client.Memcpy(pb.MemcpyRequest(hostToDevice=pb.MemcpyHostToDeviceRequest(
                hostSrcData=data, dstDeviceId=pb.DeviceId(value=dev), dstMemAddr=pb.MemAddr(value=addr)))).hostToDevice.success

        def d2h(dev, addr):
            return np.frombuffer(client.Memcpy(pb.MemcpyRequest(deviceToHost=pb.MemcpyDeviceToHostRequest(
                srcDeviceId=pb.DeviceId(value=dev), srcMemAddr=pb.MemAddr(value=addr), numBytes=0))).deviceToHost.dstData,
                np.float32)

        ids = [md.deviceId.value for md in init.devices]
        rng = np.random.default_rng(3)
        nelem = 200003  # odd: the scalar tail of the switch kernel
        grads = [rng.standard_normal(nelem).astype(np.float32) for _ in ids]
        w0 = rng.standard_normal(nelem).astype(np.float32)
        for i, g in zip(ids, grads):
            h2d(i, 0x1000, g.tobytes())
            h2d(i, 0x2000, w0.tobytes())
        before = [d.service.backend.pointer(0x1000) for d in devs]
        # 1. fused average + SGD in the switch-reduced kernel
        assert client.AllReduceRing(pb.AllReduceRingRequest(commId=init.commId, count=nelem * 4),
                                    metadata=(("dmlp-algo", "nvls"), ("dmlp-sgd-lr", "0.05"))).success
        assert comm.heap, "the symmetric heap was not created"
        after = [d.service.backend.pointer(0x1000) for d in devs]
        assert all(a != b for a, b in zip(after, before))
        offs = {a - comm.heap.uc[i] for i, a in enumerate(after)}
        assert len(offs) == 1, f"not symmetric: {offs}"
        ref = w0.astype(np.float64) - 0.05 * np.mean(np.stack(grads).astype(np.float64), axis=0)
        ws = [d2h(i, 0x2000) for i in ids]
        for w in ws:
            assert np.abs(w - ref).max() < 1e-5 and np.array_equal(w, ws[0])
        for i, g in zip(ids, grads):  # gradients untouched by the fused form
            assert np.array_equal(d2h(i, 0x1000), g)
        # 2. plain sum, in place on the heap
        assert client.AllReduceRing(pb.AllReduceRingRequest(commId=init.commId, count=nelem * 4),
                                    metadata=(("dmlp-algo", "nvls"),)).success
        s = np.sum(np.stack(grads).astype(np.float64), axis=0)
        outs = [d2h(i, 0x1000) for i in ids]
        for o in outs:
            assert np.abs(o - s).max() < 1e-4 and np.array_equal(o, outs[0])
        # 3. the unicast kernels read and write heap memory of the peers too
        for i, g in zip(ids, grads):
            h2d(i, 0x1000, g.tobytes())
        assert client.AllReduceRing(pb.AllReduceRingRequest(commId=init.commId, count=nelem * 4),
                                    metadata=(("dmlp-algo", "twoshot"),)).success
        for i in ids:
            assert np.abs(d2h(i, 0x1000) - s).max() < 1e-4
        # 4. auto: the dispatcher table decides (same answer either way), a group of two launches, one wait
        for i, g in zip(ids, grads):
            h2d(i, 0x1000, g.tobytes())
        assert client.GroupStart(pb.GroupStartRequest(commId=init.commId)).success
        for _ in range(2):
            assert client.AllReduceRing(pb.AllReduceRingRequest(commId=init.commId, count=nelem * 4)).success
        assert client.GroupEnd(pb.GroupEndRequest(commId=init.commId)).success
        for i in ids:
            assert np.abs(d2h(i, 0x1000) - n * s).max() < 1e-3 * n
        # 5. a larger rewrite of the address leaves the slot; the next collective places it again
        big = [rng.standard_normal(2 * nelem).astype(np.float32) for _ in ids]
        for i, g in zip(ids, big):
            h2d(i, 0x1000, g.tobytes())
        assert client.AllReduceRing(pb.AllReduceRingRequest(commId=init.commId, count=2 * nelem * 4),
                                    metadata=(("dmlp-algo", "nvls"),)).success
        sb = np.sum(np.stack(big).astype(np.float64), axis=0)
        for i in ids:
            assert np.abs(d2h(i, 0x1000) - sb).max() < 1e-4
        # 6. destroying the communicator moves the data back to private memory: still readable through the device
        assert client.CommDestroy(pb.CommDestroyRequest(commId=init.commId)).success
        for d in devs:
            be = d.service.backend
            got = np.frombuffer(be.read(0x1000), np.float32)
            assert np.abs(got - sb).max() < 1e-4
    finally:
        coord.stop()
        for d in devs:
            d.stop()
            d.service.backend.close()
