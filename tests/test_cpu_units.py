"""CPU-only unit tests: layout/checkpoint, oracle identities, data readers, schedules."""
import os

import numpy as np
import pytest

from dmlp_b200.models import oracle
from dmlp_b200.models.layout import ParamLayout
from dmlp_b200.parallel import schedule
from dmlp_b200.utils import data


def test_layout_matches_reference_sizes():
    assert ParamLayout((784, 128, 10)).reference_count * 4 == 407080      # client.go gradient blob
    assert ParamLayout((784, 128, 64, 10)).reference_count == 109386      # north-star MLP
    assert ParamLayout((784, 8192, 8192, 10)).reference_count == 73629706  # wide MLP
    lay = ParamLayout((784, 128, 64, 10))
    for s in lay.layers:
        assert s.w_off % 64 == 0 and s.b_off % 64 == 0 and s.out_pad % 16 == 0
    assert lay.total % 64 == 0


def test_reference_blob_roundtrip_and_order():
    lay = ParamLayout((6, 4, 3))
    flat = lay.init_reference(0)
    blob = lay.to_reference_bytes(flat)
    assert len(blob) == (6 * 4 + 4 + 4 * 3 + 3) * 4
    vals = np.frombuffer(blob, "<f4")
    (w1, b1), (w2, b2) = lay.views(flat)
    assert np.array_equal(vals[:24].reshape(6, 4), w1) and np.array_equal(vals[24:28], b1)
    assert np.array_equal(vals[28:40].reshape(4, 3), w2) and np.array_equal(vals[40:], b2)
    assert np.array_equal(lay.from_reference_bytes(blob), flat)
    assert np.abs(flat).max() <= 0.05 + 1e-7
    with pytest.raises(ValueError):
        lay.from_reference_bytes(blob[:-4])


def test_oracle_gradients_match_finite_differences():
    rng = np.random.default_rng(0)
    params = [(rng.standard_normal((5, 4)) * 0.3, rng.standard_normal(4) * 0.1),
              (rng.standard_normal((4, 3)) * 0.3, rng.standard_normal(3) * 0.1)]
    x, y = rng.standard_normal((6, 5)), rng.integers(0, 3, 6)
    loss, _, grads = oracle.loss_and_grads(params, x, y)
    eps = 1e-6
    for l in range(2):
        for which in range(2):
            p = params[l][which]
            idx = tuple(0 for _ in p.shape)
            old = p[idx]
            p[idx] = old + eps
            lp = oracle.loss_and_grads(params, x, y)[0]
            p[idx] = old - eps
            lm = oracle.loss_and_grads(params, x, y)[0]
            p[idx] = old
            assert abs((lp - lm) / (2 * eps) - grads[l][which][idx]) < 1e-5


@pytest.mark.parametrize("n,size", [(2, 10), (3, 407080), (3, 262144), (5, 7), (8, 1001)])
def test_ring_schedule_equals_sum_for_any_size(n, size):
    rng = np.random.default_rng(n)
    bufs = [rng.integers(0, 256, size, dtype=np.uint8) for _ in range(n)]
    out = oracle.ring_allreduce(bufs)  # uint8 wrap-around like the reference's byte adds
    ref = np.sum(np.stack(bufs).astype(np.uint32), axis=0).astype(np.uint8)
    for o in out:
        assert np.array_equal(o, ref)
    assert np.array_equal(oracle.naive_allreduce(bufs)[0], ref)


def test_dp_step_equals_large_batch_step():
    rng = np.random.default_rng(1)
    params = [(rng.standard_normal((8, 6)) * 0.2, np.zeros(6)), (rng.standard_normal((6, 3)) * 0.2, np.zeros(3))]
    x, y = rng.standard_normal((12, 8)), rng.integers(0, 3, 12)
    one, _, _ = oracle.train_step(params, x, y, 0.1)
    dp, _, _ = oracle.dp_train_step(params, [(x[:6], y[:6]), (x[6:], y[6:])], 0.1)
    for (a, b), (c, d) in zip(one, dp):
        assert np.allclose(a, c) and np.allclose(b, d)


def test_idx_reader_roundtrip_and_magic(tmp_path):
    imgs = (np.arange(3 * 28 * 28) % 256).astype(np.uint8).reshape(3, 28, 28)
    labels = np.array([3, 1, 4], np.int32)
    pi, pl = os.path.join(tmp_path, "i.gz"), os.path.join(tmp_path, "l.gz")
    data.write_idx_images(pi, imgs)
    data.write_idx_labels(pl, labels)
    x = data.load_idx_images(pi)
    assert x.shape == (3, 784) and x.dtype == np.float32 and np.allclose(x, imgs.reshape(3, -1) / 255.0)
    assert np.array_equal(data.load_idx_labels(pl), labels)
    assert np.array_equal(data.one_hot(labels)[0], np.eye(10, dtype=np.float32)[3])
    with pytest.raises(ValueError):
        data.load_idx_images(pl)
    with pytest.raises(ValueError):
        data.load_idx_labels(pi)


def test_reference_t10k_files_if_mounted():
    found = data.find_reference_data()
    if not found or "test_images" not in found:
        pytest.skip("reference data not mounted")
    x = data.load_idx_images(found["test_images"])
    y = data.load_idx_labels(found["test_labels"])
    assert x.shape == (10000, 784) and y.shape == (10000,) and 0 <= y.min() and y.max() == 9


def test_synthetic_mnist_shape_and_learnability():
    x, y = data.synthetic_mnist(512, seed=3)
    assert x.shape == (512, 784) and x.dtype == np.float32 and y.dtype == np.int32
    assert 0.0 <= x.min() and x.max() <= 1.0 and set(np.unique(y)) <= set(range(10))
    lay = ParamLayout((784, 32, 10))
    params = [(w.astype(np.float64), b.astype(np.float64)) for w, b in lay.views(lay.init_reference(0))]
    first = last = None
    for it in range(30):
        params, loss, _ = oracle.train_step(params, x[:256].astype(np.float64), y[:256], 0.5)
        first = loss if first is None else first
        last = loss
    assert last < 0.5 * first


def test_schedules():
    assert schedule.constant(0.01)(123) == 0.01
    assert schedule.step_decay(1.0, 10, 0.5)(25) == 0.25
    c = schedule.cosine(1.0, 100)
    assert abs(c(0) - 1.0) < 1e-9 and abs(c(100)) < 1e-9 and c(50) == pytest.approx(0.5)
    w = schedule.warmup(schedule.constant(1.0), 4)
    assert w(0) == 0.25 and w(3) == 1.0 and w(10) == 1.0
    assert schedule.make("warmup-cosine", 0.1, 100)(0) < 0.1


def test_checkpoint_sidecar_roundtrip(tmp_path):
    """Checkpoint = reference weight blob + JSON sidecar (step counter, epoch, schedule, shape); SURVEY 5.4."""
    import numpy as np
    from dmlp_b200.models.layout import ParamLayout
    from dmlp_b200.utils import checkpoint as ck
    lay = ParamLayout((784, 128, 64, 10))
    flat = lay.init_reference(3)
    blob = lay.to_reference_bytes(flat)
    path = str(tmp_path / "w.bin")
    ck.write_checkpoint(path, blob, dims=(784, 128, 64, 10), batch=64, global_step=1874, epoch=2, lr=0.01,
                        schedule="cosine", world=8)
    got, meta = ck.read_checkpoint(path, expect_dims=(784, 128, 64, 10))
    assert got == blob and meta["global_step"] == 1874 and meta["epoch"] == 2 and meta["schedule"] == "cosine"
    assert np.array_equal(lay.from_reference_bytes(got), lay.from_reference_bytes(blob))
    import pytest
    with pytest.raises(ValueError):
        ck.read_checkpoint(path, expect_dims=(784, 128, 10))
    # a bare blob (what the reference client would produce) loads with empty metadata
    bare = str(tmp_path / "bare.bin")
    open(bare, "wb").write(blob)
    got2, meta2 = ck.read_checkpoint(bare)
    assert got2 == blob and meta2 == {}
    # truncated blob against its sidecar is refused
    open(path, "wb").write(blob[:-4])
    with pytest.raises(ValueError):
        ck.read_checkpoint(path)


def test_compute_helpers_shapes_and_labels():
    """services/compute.py: model shape from the blob size, batch from the buffer size, both label encodings."""
    import numpy as np
    import pytest
    from dmlp_b200.services import compute as cp
    assert cp.infer_dims(407080) == (784, 128, 10)          # the reference's weight blob (client.go:204-218)
    assert cp.infer_dims(437544) == (784, 128, 64, 10)      # the README's three-layer shape
    assert cp.infer_dims(4 * (8 * 16 + 16 + 16 * 4 + 4), hint=(8, 16, 4)) == (8, 16, 4)
    with pytest.raises(ValueError):
        cp.infer_dims(1000)
    assert cp.batch_of(64 * 784 * 4, 784) == 64
    with pytest.raises(ValueError):
        cp.batch_of(100, 784)
    y = np.array([3, 0, 9, 1], dtype=np.int32)
    assert np.array_equal(cp.labels_from_bytes(np.eye(10, dtype="<f4")[y].tobytes(), 4, 10), y)
    assert np.array_equal(cp.labels_from_bytes(y.astype("<i4").tobytes(), 4, 10), y)
    with pytest.raises(ValueError):
        cp.labels_from_bytes(b"\x00" * 12, 4, 10)
    with pytest.raises(ValueError):
        cp.labels_from_bytes(np.array([0, 1, 2, 10], dtype="<i4").tobytes(), 4, 10)


def test_grpc_client_cli_against_in_process_servers():
    """`python -m dmlp_b200.grpc_client` main(): coordinator + 2 sim devices in this process, two short epochs."""
    from dmlp_b200 import grpc_client
    from dmlp_b200.services import server as srv
    from dmlp_b200.services.backend import SimBackend
    devs = [srv.start_device_server(SimBackend(i + 1, 1 << 20)) for i in range(2)]
    coord = srv.start_coordinator_server()
    try:
        rc = grpc_client.main(["--coordinator", coord.addr, "--devices", ",".join(d.addr for d in devs), "--batch", "8",
                               "--steps-per-epoch", "2", "--epochs", "2", "--samples", "64", "--dims", "784,128,10"])
        assert rc == 0
        assert not coord.service.communicators  # CommDestroy at the end of the run
    finally:
        coord.stop()
        for d in devs:
            d.stop()


def test_real_mnist_t10k_convergence_with_the_oracle_if_mounted():
    """Real data sanity: the reference ships only the t10k split. Train the reference-shaped MLP
    on 8 000 of its images with the NumPy oracle (the same math the kernels are tested against: client.go:112-267,
    batch 64, SGD) and evaluate on the other 2 000."""
    import numpy as np
    import pytest
    from dmlp_b200.models import oracle
    from dmlp_b200.models.layout import ParamLayout
    from dmlp_b200.utils import data
    found = data.find_reference_data()
    if not found or "test_images" not in found or "test_labels" not in found:
        pytest.skip("reference data not mounted")
    x = data.load_idx_images(found["test_images"]).astype(np.float64)
    y = data.load_idx_labels(found["test_labels"])
    xtr, ytr, xte, yte = x[:8000], y[:8000], x[8000:], y[8000:]
    lay = ParamLayout((784, 128, 10))  # the shape the reference's code trains (client.go:22-33)
    params = [(w.astype(np.float64), b.astype(np.float64)) for w, b in lay.views(lay.init_reference(0))]
    lr, B = 0.1, 64
    first = None
    for epoch in range(3):
        for s in range(8000 // B):
            params, loss, _ = oracle.train_step(params, xtr[s * B:(s + 1) * B], ytr[s * B:(s + 1) * B], lr)
            first = loss if first is None else first
    logits, _ = oracle.forward(params, xte)
    acc = float((logits.argmax(1) == yte).mean())
    assert loss < 0.6 * first and acc > 0.85, (first, loss, acc)


def test_fault_spec_parsing():
    """``Config.fault`` / ``bench.py --fault`` syntax (BASELINE config 5: a rank dies at a given step)."""
    import pytest
    from dmlp_b200.parallel.schedule import parse_fault
    assert parse_fault("") == (-1, -1)
    assert parse_fault(None) == (-1, -1)
    assert parse_fault("dead-rank:3@step:40") == (3, 40)
    for bad in ("dead-rank:3", "rank:1@step:2", "dead-rank:x@step:2", "dead-rank:1@epoch:2"):
        with pytest.raises(ValueError):
            parse_fault(bad)


def test_nvtx_tracing_is_a_no_op_without_a_gpu(monkeypatch):
    """utils/tracing.py: ranges and the decorator work with DMLP_NVTX on or off, with or without CUDA (here: without)."""
    from dmlp_b200.utils import tracing
    calls = []

    @tracing.traced("dmlp.test")
    def f(a, b=2):
        calls.append((a, b))
        return a + b

    monkeypatch.delenv("DMLP_NVTX", raising=False)
    assert not tracing.enabled() and f(1) == 3
    monkeypatch.setenv("DMLP_NVTX", "1")
    assert tracing.enabled() and f(2, b=5) == 7 and f.__name__ == "f"
    with tracing.nvtx_range("dmlp.block"):
        calls.append("inside")
    assert calls == [(1, 2), (2, 5), "inside"]
    # with a tracer bound, push / pop bracket the call even when it raises
    pushed = []
    monkeypatch.setattr(tracing, "_PUSH", lambda n: pushed.append(("push", n)))
    monkeypatch.setattr(tracing, "_POP", lambda: pushed.append(("pop",)))

    @tracing.traced("dmlp.boom")
    def boom():
        raise ValueError("x")

    import pytest
    with pytest.raises(ValueError):
        boom()
    assert pushed == [("push", "dmlp.boom"), ("pop",)]
