"""All-reduce kernels vs a NumPy/PyTorch oracle (GPU only).

Runs on ONE GPU: a communicator whose ranks all live on device 0 is "emulated" — a single
cooperative launch plays every rank (gridDim.y = world) through exactly the same flag protocol
and data paths the multi-GPU launches use. Real multi-GPU runs are in test_gpu_multigpu.py.
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

ALGOS = ["oneshot", "twoshot", "ring"]


def _comm(world, staging=1 << 20):
    from dmlp_b200.runtime.comm import Communicator
    return Communicator.local([0] * world, staging_bytes=staging)


def _native():
    from dmlp_b200.runtime import native
    return native


@pytest.mark.parametrize("algo", ALGOS)
@pytest.mark.parametrize("world", [1, 2, 3, 8])
@pytest.mark.parametrize("count", [1, 4, 101770, 262144 + 3])
def test_allreduce_f32_sum(algo, world, count):
    n = _native()
    torch.manual_seed(world * 1000 + count)
    comm = _comm(world)
    ins = [torch.randn(count, device="cuda") for _ in range(world)]
    outs = [torch.full((count,), float("nan"), device="cuda") for _ in range(world)]
    ref = torch.stack([t.double() for t in ins]).sum(0)
    comm.allreduce([t.data_ptr() for t in ins], [t.data_ptr() for t in outs], count,
                   dtype=n.F32, op=n.SUM, algo=n.ALGO_BY_NAME[algo])
    torch.cuda.synchronize()
    comm.check()
    for o in outs:
        assert (o.double() - ref).abs().max().item() < 1e-5 * max(1.0, ref.abs().max().item())
    for o in outs[1:]:
        assert torch.equal(o, outs[0]), "replicas must be bit-identical"
    comm.destroy()


@pytest.mark.parametrize("algo", ALGOS)
def test_allreduce_in_place_and_repeat(algo):
    """In-place (out == in) and back-to-back calls reuse flags/epochs/staging slots correctly."""
    n = _native()
    world, count = 3, 50001
    comm = _comm(world, staging=64 << 10)  # small staging: the ring needs several slabs
    bufs = [torch.randn(count, device="cuda") for _ in range(world)]
    ref = torch.stack([t.double() for t in bufs]).sum(0)
    for it in range(4):
        comm.allreduce([t.data_ptr() for t in bufs], [t.data_ptr() for t in bufs], count,
                       algo=n.ALGO_BY_NAME[algo])
        ref = ref * world if it > 0 else ref
    torch.cuda.synchronize()
    comm.check()
    for b in bufs:
        assert (b.double() - ref).abs().max().item() < 1e-4 * ref.abs().max().item()
    comm.destroy()


@pytest.mark.parametrize("algo", ALGOS)
@pytest.mark.parametrize("dtype,op", [("u8", "sum"), ("u8", "max"), ("i32", "sum"), ("i32", "min"), ("i32", "prod"),
                                      ("bf16", "sum"), ("bf16", "max"), ("f32", "prod"), ("f32", "min"), ("f32", "avg")])
def test_allreduce_dtypes_ops(algo, dtype, op):
    n = _native()
    world, count = 3, 4099
    comm = _comm(world)
    rng = np.random.default_rng(7)
    if dtype == "u8":
        arrs = [rng.integers(0, 256, count, dtype=np.uint8) for _ in range(world)]
        tt, nd = torch.uint8, n.U8
    elif dtype == "i32":
        arrs = [rng.integers(-5, 6, count).astype(np.int32) for _ in range(world)]
        tt, nd = torch.int32, n.I32
    else:
        arrs = [rng.standard_normal(count).astype(np.float32) for _ in range(world)]
        tt, nd = (torch.bfloat16, n.BF16) if dtype == "bf16" else (torch.float32, n.F32)
    ins = [torch.from_numpy(a).cuda().to(tt) for a in arrs]
    outs = [torch.zeros(count, device="cuda", dtype=tt) for _ in range(world)]
    nop = {"sum": n.SUM, "prod": n.PROD, "min": n.MIN, "max": n.MAX, "avg": n.AVG}[op]
    comm.allreduce([t.data_ptr() for t in ins], [t.data_ptr() for t in outs], count, dtype=nd, op=nop,
                   algo=n.ALGO_BY_NAME[algo])
    torch.cuda.synchronize()
    comm.check()
    host = [t.float().cpu().numpy().astype(np.float64) if dtype in ("bf16", "f32") else t.cpu().numpy() for t in ins]
    if dtype == "u8":
        acc = host[0].astype(np.uint8)
        for h in host[1:]:
            acc = {"sum": lambda a, b: (a.astype(np.uint16) + b).astype(np.uint8), "max": np.maximum}[op](acc, h)
        assert np.array_equal(outs[0].cpu().numpy(), acc)  # wrap-around byte add like the reference
    elif dtype == "i32":
        acc = host[0].copy()
        for h in host[1:]:
            acc = {"sum": np.add, "min": np.minimum, "prod": np.multiply}[op](acc, h)
        assert np.array_equal(outs[0].cpu().numpy(), acc)
    else:
        acc = host[0].copy()
        for h in host[1:]:
            acc = {"sum": np.add, "prod": np.multiply, "min": np.minimum, "max": np.maximum, "avg": np.add}[op](acc, h)
        if op == "avg":
            acc = acc / world
        tol = 3e-2 if dtype == "bf16" else 1e-5
        got = outs[0].float().cpu().numpy().astype(np.float64)
        assert np.abs(got - acc).max() < tol * max(1.0, np.abs(acc).max())
    for o in outs[1:]:
        assert torch.equal(o, outs[0])
    comm.destroy()


@pytest.mark.parametrize("algo", ALGOS)
@pytest.mark.parametrize("world", [1, 2, 4])
def test_fused_allreduce_avg_sgd(algo, world):
    """w -= lr * mean_r(g_r), fused into the collective; replicas stay identical; bf16 shadow refreshed."""
    n = _native()
    count = 109386  # parameters of the 784-128-64-10 MLP
    torch.manual_seed(3)
    comm = _comm(world)
    w0 = torch.randn(count, device="cuda")
    ws = [w0.clone() for _ in range(world)]
    shadows = [torch.zeros(count, device="cuda", dtype=torch.bfloat16) for _ in range(world)]
    gs = [torch.randn(count, device="cuda") for _ in range(world)]
    lrs = [torch.tensor([0.01], device="cuda") for _ in range(world)]
    ref = w0.double() - 0.01 * torch.stack([g.double() for g in gs]).mean(0)
    for _ in range(2):
        comm.allreduce([g.data_ptr() for g in gs], [w.data_ptr() for w in ws], count, algo=n.ALGO_BY_NAME[algo],
                       fuse_sgd=True, lr=[t.data_ptr() for t in lrs], scale=1.0 / world,
                       out_bf16=[s.data_ptr() for s in shadows])
    ref = ref - 0.01 * torch.stack([g.double() for g in gs]).mean(0)
    torch.cuda.synchronize()
    comm.check()
    for w, s in zip(ws, shadows):
        assert (w.double() - ref).abs().max().item() < 1e-5
        assert torch.equal(w, ws[0])
        assert torch.equal(s, w.to(torch.bfloat16))
    comm.destroy()


@pytest.mark.parametrize("algo", ALGOS)
def test_dead_rank_is_detected_not_hung(algo):
    """Fault injection: one rank never arrives -> survivors time out in-kernel, the failure is
    reported with {rank, peer, phase}, and the communicator refuses further work."""
    from dmlp_b200.runtime.comm import CommFailure
    n = _native()
    world, count = 3, 4096
    comm = _comm(world)
    comm.set_timeout_ms(200)
    bufs = [torch.ones(count, device="cuda") for _ in range(world)]
    comm.allreduce([t.data_ptr() for t in bufs], [t.data_ptr() for t in bufs], count, algo=n.ALGO_BY_NAME[algo])
    torch.cuda.synchronize()
    comm.check()
    comm.inject_dead_ranks(1 << 1)
    comm.allreduce([t.data_ptr() for t in bufs], [t.data_ptr() for t in bufs], count, algo=n.ALGO_BY_NAME[algo])
    torch.cuda.synchronize()  # returns: kernels exit on their watchdog
    st = comm.status()
    # The reporter is a survivor. It names the peer it was waiting for: the dead rank itself, or (ring)
    # a survivor that is stalled behind the dead rank.
    assert st["code"] == n.ERR_TIMEOUT and st["rank"] in (0, 2) and st["peer"] != st["rank"]
    if algo != "ring":
        assert st["peer"] == 1
    with pytest.raises(CommFailure):
        comm.check()
    with pytest.raises(n.NativeError) as ei:
        comm.allreduce([t.data_ptr() for t in bufs], [t.data_ptr() for t in bufs], count)
    assert ei.value.code == n.ERR_ABORTED
    comm.destroy()


def test_p2p_copy_and_reduce_n():
    import ctypes as C
    n = _native()
    src = torch.arange(1000003, device="cuda", dtype=torch.uint8)
    dst = torch.zeros_like(src)
    n.check(n.lib().dmlp_p2p_copy(0, dst.data_ptr(), src.data_ptr(), src.numel(), None))
    # unaligned views take the byte path
    n.check(n.lib().dmlp_p2p_copy(0, dst[1:].data_ptr(), src[1:].data_ptr(), 999, None))
    torch.cuda.synchronize()
    assert torch.equal(dst, src)
    xs = [torch.randn(5000, device="cuda") for _ in range(3)]
    out = torch.empty(5000, device="cuda")
    ptrs = (C.c_void_p * 3)(*[x.data_ptr() for x in xs])
    n.check(n.lib().dmlp_reduce_n(0, out.data_ptr(), ptrs, 3, 5000, n.F32, n.SUM, None))
    torch.cuda.synchronize()
    assert torch.allclose(out, xs[0] + xs[1] + xs[2], atol=1e-6)


def test_device_info_and_topology():
    from dmlp_b200.runtime import comm as rc
    assert rc.device_count() >= 1
    p = rc.device_props(0)
    assert p.sm_count > 0 and p.cc.startswith("10.") and p.uuid.startswith("GPU-")
    m = rc.peer_matrix([0])
    assert m["access"] == [[1]]
