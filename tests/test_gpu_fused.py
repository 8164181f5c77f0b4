"""Cluster-fused step kernel (one launch = forward + loss + backward) vs the NumPy oracle and vs the
per-layer kernel path."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _trainer(dims, fused, batch=64, lr=0.05, use_graph=False):
    from dmlp_b200.parallel.trainer import Trainer, TrainerConfig
    return Trainer(TrainerConfig(dims=dims, batch=batch, lr=lr, fused=fused, use_graph=use_graph), device=0)


def _tf32(a: np.ndarray) -> np.ndarray:
    """fp32 -> what kind::tf32 consumes (10-bit mantissa, low 13 bits dropped), as float64 for exact accumulation."""
    return (np.ascontiguousarray(a, dtype=np.float32).view(np.int32) & ~0x1FFF).view(np.float32).astype(np.float64)


def _tf32_chain_oracle(params, x, y):
    """The reference math (client.go:112-202) with every GEMM operand truncated to tf32 where the kernel feeds it to a
    tensor core, everything else in float64: pins the kernel's data flow (K slices, exchanges, masks) far tighter than
    the plain oracle, whose distance to a tf32 kernel is dominated by operand rounding."""
    B = x.shape[0]
    acts, h = [x.astype(np.float64)], x.astype(np.float64)
    for i, (w, b) in enumerate(params):
        z = _tf32(h) @ _tf32(w) + b
        h = np.maximum(z, 0) if i < len(params) - 1 else z
        acts.append(h)
    logits = acts[-1]
    e = np.exp(logits - logits.max(1, keepdims=True))
    p = e / e.sum(1, keepdims=True)
    onehot = np.zeros_like(p)
    onehot[np.arange(B), y] = 1
    loss = float(np.mean(-np.log((p * onehot).sum(1) + 1e-10)))
    dz = (p - onehot) / B
    grads = [None] * len(params)
    for l in range(len(params) - 1, -1, -1):
        grads[l] = (_tf32(acts[l]).T @ _tf32(dz), _tf32(dz).sum(0))  # (bias gradient = ones column of the same MMA)
        if l > 0:
            dz = (_tf32(dz) @ _tf32(params[l][0]).T) * (acts[l] > 0)
    return loss, grads


@pytest.mark.parametrize("dims", [(784, 128, 64, 10), (784, 128, 10)])
@pytest.mark.parametrize("batch", [64, 32])
def test_fused_gradients_match_oracle(dims, batch):
    """One step with lr = 0: parameters stay put, the gradient buffer holds what the kernel produced."""
    from dmlp_b200.models import oracle
    from dmlp_b200.utils.data import synthetic_mnist
    t = _trainer(dims, fused=2, batch=batch, lr=0.0)  # mode 2: gradients are written out (mode 1 applies SGD in place)
    x, y = synthetic_mnist(batch * 4, seed=5)
    flat0 = t.get_params()
    rng = np.random.default_rng(0)
    flat = flat0.copy()
    for s in t.layout.layers:  # bigger weights than the reference init: non-trivial activations / masks
        w = flat[s.w_off:s.w_off + s.fan_in * s.out_pad].reshape(s.fan_in, s.out_pad)
        w[:, :s.fan_out] = rng.standard_normal((s.fan_in, s.fan_out)).astype(np.float32) * (2.0 / np.sqrt(s.fan_in))
        flat[s.b_off:s.b_off + s.fan_out] = rng.standard_normal(s.fan_out).astype(np.float32) * 0.1
    t.set_params(flat)
    params = [(w.astype(np.float64), b.astype(np.float64)) for w, b in t.weights()]
    st = t.fit(x, y, 1)
    loss, correct, grads = oracle.loss_and_grads(params, x[:batch].astype(np.float64), y[:batch])
    assert abs(st["per_step_loss"][0] - loss) < 2e-2 * max(1.0, loss), (st["per_step_loss"][0], loss)
    g = t.debug_buffer(2, 0, t.layout.total)
    for l, ((gw, gb), s) in enumerate(zip(grads, t.layout.layers)):
        kw = g[s.w_off:s.w_off + s.fan_in * s.out_pad].reshape(s.fan_in, s.out_pad)[:, :s.fan_out]
        kb = g[s.b_off:s.b_off + s.fan_out]
        sw, sb = max(1e-4, np.abs(gw).max()), max(1e-4, np.abs(gb).max())
        assert np.abs(kw - gw).max() < 5e-2 * sw, f"dW{l + 1}: {np.abs(kw - gw).max()} vs scale {sw}"
        assert np.abs(kb - gb).max() < 5e-2 * sb, f"db{l + 1}: {np.abs(kb - gb).max()} vs scale {sb}"
    # against the oracle that truncates operands where the kernel does: an order of magnitude tighter (a wrong K-slice
    # or exchange contribution of a percent would show)
    loss_t, grads_t = _tf32_chain_oracle(params, x[:batch], y[:batch])
    assert abs(st["per_step_loss"][0] - loss_t) < 2e-3 * max(1.0, loss_t), (st["per_step_loss"][0], loss_t)
    for l, ((gw, gb), s) in enumerate(zip(grads_t, t.layout.layers)):
        kw = g[s.w_off:s.w_off + s.fan_in * s.out_pad].reshape(s.fan_in, s.out_pad)[:, :s.fan_out]
        kb = g[s.b_off:s.b_off + s.fan_out]
        sw, sb = max(1e-4, np.abs(gw).max()), max(1e-4, np.abs(gb).max())
        assert np.abs(kw - gw).max() < 5e-3 * sw, f"dW{l + 1} vs tf32 oracle: {np.abs(kw - gw).max()} vs scale {sw}"
        assert np.abs(kb - gb).max() < 5e-3 * sb, f"db{l + 1} vs tf32 oracle: {np.abs(kb - gb).max()} vs scale {sb}"
    t.close()


@pytest.mark.parametrize("dims", [(784, 128, 64, 10), (784, 128, 10)])
@pytest.mark.parametrize("use_graph", [False, True])
def test_fused_training_matches_per_layer_path(dims, use_graph):
    from dmlp_b200.utils.data import synthetic_mnist
    x, y = synthetic_mnist(64 * 20, seed=2)
    a, b = _trainer(dims, fused=1, use_graph=use_graph, lr=0.1), _trainer(dims, fused=0, use_graph=use_graph, lr=0.1)
    sa, sb = a.fit(x, y, 20), b.fit(x, y, 20)
    assert np.allclose(sa["per_step_loss"], sb["per_step_loss"], rtol=2e-2, atol=2e-3)
    assert np.abs(a.get_params() - b.get_params()).max() < 3e-3
    assert a.kernels_per_step == 1 and b.kernels_per_step >= 6
    # single-GPU fused path: a chunk of steps runs inside ONE launch (weights updated in place between steps)
    assert a.launches_last_run == 3 and b.launches_last_run == 0
    a.close()
    b.close()


def test_fused_multi_step_launch_with_lr_schedule():
    """A learning-rate change inside a chunk splits the launch; results still follow the per-layer path."""
    from dmlp_b200.utils.data import synthetic_mnist
    dims = (784, 128, 64, 10)
    x, y = synthetic_mnist(64 * 24, seed=5)
    sched = lambda s: 0.1 if s < 5 else (0.05 if s < 13 else 0.02)  # noqa: E731
    a, b = _trainer(dims, fused=1, lr=0.1), _trainer(dims, fused=0, lr=0.1)
    sa, sb = a.fit(x, y, 24, lr_schedule=sched), b.fit(x, y, 24, lr_schedule=sched)
    assert np.allclose(sa["per_step_loss"], sb["per_step_loss"], rtol=2e-2, atol=2e-3)
    assert np.abs(a.get_params() - b.get_params()).max() < 3e-3
    assert a.launches_last_run == 5  # [0,5) [5,8) [8,13) [13,16) [16,24)
    a.close()
    b.close()


def test_fused_is_auto_selected_only_for_supported_shapes():
    a = _trainer((784, 128, 64, 10), fused=-1)
    assert a.kernels_per_step == 0  # not run yet
    from dmlp_b200.utils.data import synthetic_mnist
    x, y = synthetic_mnist(64 * 4, seed=1)
    a.fit(x, y, 2)
    assert a.kernels_per_step == 1
    a.close()
    b = _trainer((784, 256, 64, 10), fused=-1)
    b.fit(x, y, 2)
    assert b.kernels_per_step >= 9
    b.close()
    from dmlp_b200.runtime.native import NativeError
    with pytest.raises(NativeError):
        _trainer((784, 256, 64, 10), fused=1)
