"""PyTorch framework layer (autograd functions / MLP module / DataParallel) vs torch autograd."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _torch_reference(module, x, y):
    """Same model in plain PyTorch fp32 ops, fed the module's current weights."""
    ws = [w.detach().clone().requires_grad_(True) for w in module.weights]
    bs = [b.detach().clone().requires_grad_(True) for b in module.biases]
    h = x
    for l, (w, b) in enumerate(zip(ws, bs)):
        h = h @ w + b
        if l < len(ws) - 1:
            h = torch.relu(h)
    C = module.num_classes
    p_true = torch.softmax(h[:, :C], 1).gather(1, y.long()[:, None])
    loss = -(torch.log(p_true + 1e-10)).mean()  # the reference's epsilon form (client.go:148-155)
    loss.backward()
    return loss.item(), ws, bs, h


@pytest.mark.parametrize("dims,B", [((784, 128, 64, 10), 64), ((784, 128, 10), 64), ((256, 512, 256, 10), 300)])
def test_mlp_loss_and_grads_match_torch(dims, B):
    from dmlp_b200.models.mlp import MLP
    torch.manual_seed(0)
    m = MLP(dims, seed=1)
    with torch.no_grad():
        m.flat.mul_(3.0)  # leave the tiny-init regime so gradients are well conditioned
        for s in m.layout.layers:  # keep the padded columns at zero
            m.flat[s.w_off:s.w_off + s.fan_in * s.out_pad].view(s.fan_in, s.out_pad)[:, s.fan_out:] = 0
            m.flat[s.b_off + s.fan_out:s.b_off + s.out_pad] = 0
    x = torch.rand(B, dims[0], device="cuda")
    y = torch.randint(0, dims[-1], (B,), device="cuda", dtype=torch.int32)
    ref_loss, ws, bs, ref_logits = _torch_reference(m, x, y)
    logits = m(x)
    assert (logits - ref_logits).abs().max().item() < 5e-3 * max(1.0, ref_logits.abs().max().item())
    loss, correct = m.loss(x, y)
    torch.cuda.synchronize()
    torch.cuda.set_sync_debug_mode("error")  # loss.backward() must not synchronise with the host anywhere
    try:
        loss.backward()
    finally:
        torch.cuda.set_sync_debug_mode("default")
    torch.cuda.synchronize()
    assert abs(loss.item() - ref_loss) < 5e-3 * max(1.0, ref_loss)
    assert abs(correct.item() - (ref_logits[:, :dims[-1]].argmax(1) == y.long()).sum().item()) <= 2
    for l, s in enumerate(m.layout.layers):
        gw = m.flat_grad[s.w_off:s.w_off + s.fan_in * s.out_pad].view(s.fan_in, s.out_pad)
        gb = m.flat_grad[s.b_off:s.b_off + s.out_pad]
        scale = max(1e-3, ws[l].grad.abs().max().item())
        assert (gw - ws[l].grad).abs().max().item() < 2e-2 * scale, f"layer {l} dW"
        assert (gb - bs[l].grad).abs().max().item() < 2e-2 * max(1e-3, bs[l].grad.abs().max().item()), f"layer {l} db"


def test_generic_autograd_path_returns_grads():
    """Without caller-owned gradient buffers the functions behave like ordinary autograd ops."""
    from dmlp_b200.ops import functional as F
    torch.manual_seed(1)
    x = torch.randn(64, 128, device="cuda", requires_grad=True)
    w = (torch.randn(128, 64, device="cuda") * 0.1).requires_grad_(True)
    b = torch.randn(64, device="cuda", requires_grad=True)
    out = F.fused_linear(x, w, b, relu=True)
    out.square().sum().backward()
    xr, wr, br = [t.detach().clone().requires_grad_(True) for t in (x, w, b)]
    torch.relu(xr @ wr + br).square().sum().backward()
    for got, ref in ((x.grad, xr.grad), (w.grad, wr.grad), (b.grad, br.grad)):
        assert (got - ref).abs().max().item() < 2e-2 * max(1.0, ref.abs().max().item())


def test_data_parallel_single_rank_training_and_blob(tmp_path):
    from dmlp_b200.parallel.data_parallel import DataParallel
    from dmlp_b200.utils.data import synthetic_mnist
    dp = DataParallel((784, 128, 64, 10), device=0, lr=0.2, overlap=True)
    x, y = synthetic_mnist(64 * 30, seed=2)
    xs, ys = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
    losses = []
    for s in range(60):
        i = s % 30
        loss, _ = dp.loss(xs[i * 64:(i + 1) * 64], ys[i * 64:(i + 1) * 64])
        loss.backward()
        dp.step()
        losses.append(loss.item())
    assert losses[-1] < 0.5 * losses[0]
    blob = dp.module.reference_blob()
    assert len(blob) == 437544
    w_before = dp.module.flat.clone()
    dp.module.load_reference_blob(blob)
    assert torch.equal(dp.module.flat, w_before)
