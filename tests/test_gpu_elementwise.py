"""softmax-CE / SGD / cast kernels vs PyTorch fp32 references (GPU only)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("B,C,ld", [(64, 10, 16), (64, 10, 10), (1000, 10, 16), (37, 100, 104), (8192, 10, 16)])
def test_softmax_ce(B, C, ld):
    from dmlp_b200.ops import kernels as k
    torch.manual_seed(0)
    buf = torch.randn(B, ld, device="cuda") * 3
    logits = buf[:, :ld]
    labels = torch.randint(0, C, (B,), device="cuda", dtype=torch.int32)
    dl = torch.full((B, ld), float("nan"), device="cuda")
    loss = torch.zeros(1, device="cuda")
    correct = torch.zeros(1, device="cuda", dtype=torch.int32)
    cs = torch.zeros(ld, device="cuda")
    k.softmax_ce(logits, labels, C, 1.0 / B, dl, loss, correct, colsum=cs)
    torch.cuda.synchronize()
    lg = logits[:, :C].double()
    p = torch.softmax(lg, dim=1)
    onehot = torch.nn.functional.one_hot(labels.long(), C).double()
    ref_d = (p - onehot) / B
    ref_loss = -(torch.log((p * onehot).sum(1) + 1e-10)).sum()
    assert (dl[:, :C].double() - ref_d).abs().max().item() < 1e-5
    assert (dl[:, C:] == 0).all()
    assert abs(loss.item() - ref_loss.item()) < 1e-3 * max(1.0, abs(ref_loss.item()))
    assert correct.item() == (lg.argmax(1) == labels.long()).sum().item()
    assert (cs[:C].double() - ref_d.sum(0)).abs().max().item() < 1e-4


@pytest.mark.parametrize("n", [1, 7, 1024, 109386, 1 << 20])
def test_sgd_and_shadow(n):
    from dmlp_b200.ops import kernels as k
    torch.manual_seed(0)
    w = torch.randn(n, device="cuda")
    g = torch.randn(n, device="cuda")
    lr = torch.tensor([0.01], device="cuda")
    w16 = torch.zeros(n, device="cuda", dtype=torch.bfloat16)
    ref = w - 0.01 * 0.5 * g
    k.sgd(w, g, lr, 0.5, w16)
    torch.cuda.synchronize()
    assert torch.allclose(w, ref, rtol=1e-6, atol=1e-7)
    assert torch.equal(w16, ref.to(torch.bfloat16)) or (w16.float() - ref).abs().max() < 1e-2


def test_build_info_and_errors():
    from dmlp_b200.runtime import native
    assert "sm_100a" in native.build_info()
    d = native.GemmDesc()
    rc = native.lib().dmlp_gemm(d, None)
    assert rc == native.ERR_INVALID
    assert b"dmlp_gemm" in native.lib().dmlp_last_error()
