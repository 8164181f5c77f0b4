"""Native trainer vs the NumPy oracle on one GPU: loss trajectory, weights, eval, checkpoints."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _setup(dims=(784, 128, 64, 10), batch=64, precision="tf32", lr=0.05, **kw):
    from dmlp_b200.parallel.trainer import Trainer, TrainerConfig
    from dmlp_b200.utils.data import synthetic_mnist
    cfg = TrainerConfig(dims=dims, batch=batch, lr=lr, precision=precision, **kw)
    t = Trainer(cfg, device=0)
    x, y = synthetic_mnist(batch * 40, seed=1, features=dims[0], num_classes=dims[-1])
    return t, x, y


@pytest.mark.parametrize("use_graph", [False, True])
@pytest.mark.parametrize("dims", [(784, 128, 64, 10), (784, 128, 10)])
def test_steps_match_oracle(dims, use_graph):
    from dmlp_b200.models import oracle
    t, x, y = _setup(dims=dims, use_graph=use_graph)
    params = [(w.astype(np.float64), b.astype(np.float64)) for w, b in t.weights()]
    steps = 12
    st = t.fit(x, y, steps)
    ref_losses = []
    for s in range(steps):
        xb, yb = x[s * 64:(s + 1) * 64].astype(np.float64), y[s * 64:(s + 1) * 64]
        params, loss, _ = oracle.train_step(params, xb, yb, 0.05)
        ref_losses.append(loss)
    got = st["per_step_loss"]
    assert np.allclose(got, ref_losses, rtol=2e-2, atol=2e-3), f"{got} vs {ref_losses}"
    for (w, b), (rw, rb) in zip(t.weights(), params):
        assert np.abs(w - rw).max() < 2e-3 and np.abs(b - rb).max() < 2e-3
    flat = t.get_params()
    for s in t.layout.layers:  # padded columns stay exactly zero
        wpad = flat[s.w_off:s.w_off + s.fan_in * s.out_pad].reshape(s.fan_in, s.out_pad)[:, s.fan_out:]
        assert not wpad.any()
    t.close()


def test_loss_decreases_and_eval_runs():
    t, x, y = _setup(lr=0.2)
    first = t.fit(x, y, 5)["loss"]
    hist = t.train_epochs(x, y, epochs=3)
    assert hist[-1]["loss"] < 0.5 * first
    ev = t.evaluate(x, y)
    assert ev["accuracy"] > 0.9 and ev["samples"] == 40 * 64
    t.close()


def test_bf16_mode_tracks_fp32():
    from dmlp_b200.models import oracle
    t, x, y = _setup(precision="bf16")
    params = [(w.astype(np.float64), b.astype(np.float64)) for w, b in t.weights()]
    st = t.fit(x, y, 8)
    ref = []
    for s in range(8):
        params, loss, _ = oracle.train_step(params, x[s * 64:(s + 1) * 64].astype(np.float64), y[s * 64:(s + 1) * 64], 0.05)
        ref.append(loss)
    assert np.allclose(st["per_step_loss"], ref, rtol=5e-2, atol=2e-2)
    t.close()


def test_lr_schedule_is_applied_without_recapture():
    from dmlp_b200.parallel import schedule
    t, x, y = _setup()
    w0 = t.get_params().copy()
    t.fit(x, y, 4, lr_schedule=schedule.constant(0.0))  # lr 0: nothing may move
    assert np.array_equal(t.get_params(), w0)
    t.fit(x, y, 4, lr_schedule=schedule.warmup(schedule.cosine(0.05, 100), 2))
    assert not np.array_equal(t.get_params(), w0)
    t.close()


def test_checkpoint_roundtrip_reference_layout(tmp_path):
    t, x, y = _setup(dims=(784, 128, 10))
    t.fit(x, y, 3)
    path = os.path.join(tmp_path, "weights.bin")
    t.save_weights(path)
    assert os.path.getsize(path) == 407080  # the reference's gradient/weight blob size
    before = t.get_params().copy()
    t.set_params(np.zeros_like(before))
    t.load_weights(path)
    assert np.array_equal(t.get_params(), before)
    blob = np.frombuffer(open(path, "rb").read(), "<f4")
    w1 = t.weights()[0][0]
    assert np.array_equal(blob[:784 * 128].reshape(784, 128), w1)  # [in][out] row-major first
    t.close()


def test_resume_from_checkpoint_continues_identically(tmp_path):
    """weights + step counter from the sidecar: a resumed run follows the uninterrupted one bit for bit."""
    from dmlp_b200.parallel import schedule
    sched = schedule.warmup(schedule.cosine(0.05, 40), 3)
    a, x, y = _setup()
    a.fit(x, y, 8, lr_schedule=sched)
    path = os.path.join(tmp_path, "ck.bin")
    a.save_checkpoint(path, epoch=1, schedule="cosine")
    b, _, _ = _setup()
    meta = b.load_checkpoint(path)
    assert meta["global_step"] == 8 and meta["epoch"] == 1 and b.global_step == a.global_step == 8
    assert np.array_equal(a.get_params(), b.get_params())
    a.fit(x, y, 8, first_batch=8, lr_schedule=sched)
    b.fit(x, y, 8, first_batch=8, lr_schedule=sched)
    assert np.array_equal(a.get_params(), b.get_params())
    a.close()
    b.close()


def test_wide_shape_smoke_bf16():
    """Scaled-down wide model (same code path as 784-8192-8192-10): multi-tile GEMMs, bf16 shadow."""
    from dmlp_b200.models import oracle
    t, x, y = _setup(dims=(784, 512, 512, 10), batch=512, precision="bf16")
    params = [(w.astype(np.float64), b.astype(np.float64)) for w, b in t.weights()]
    st = t.fit(x, y, 3)
    ref = []
    for s in range(3):
        params, loss, _ = oracle.train_step(params, x[s * 512:(s + 1) * 512].astype(np.float64), y[s * 512:(s + 1) * 512], 0.05)
        ref.append(loss)
    assert np.allclose(st["per_step_loss"], ref, rtol=5e-2, atol=2e-2)
    t.close()


def test_resident_timing_path():
    t, x, y = _setup()
    t.load_resident_inputs(x, y)
    ms = t.run_resident(20)
    assert ms > 0 and t.kernels_per_step >= 1
    t.close()
