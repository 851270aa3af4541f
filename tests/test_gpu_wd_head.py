"""Fused Wide&Deep head (output layer + wide part + BCE) vs plain PyTorch fp32."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("bsz,n", [(64, 400), (1000, 400), (33, 64)])
def test_head_bce_matches_fp32_reference(bsz, n):
    from paddle_operator_b200.ops.wd_head import head_bce_eligible, head_bce_loss

    torch.manual_seed(bsz + n)
    dev = "cuda"
    h = torch.relu(torch.randn(bsz, n, device=dev)).to(torch.bfloat16)
    w_out = (torch.randn(1, n, device=dev) * 0.05).to(torch.bfloat16)
    b_out = torch.tensor([0.1], device=dev, dtype=torch.bfloat16)
    wide = torch.randn(bsz, device=dev) * 0.3
    dense = torch.randn(bsz, 13, device=dev)
    w_d = (torch.randn(1, 13, device=dev) * 0.2).to(torch.bfloat16)
    b_d = torch.tensor([-0.2], device=dev, dtype=torch.bfloat16)
    labels = (torch.rand(bsz, device=dev) > 0.5).float()
    assert head_bce_eligible(h, w_out, dense)
    ha, wa, ba, wia, wda, bda = (t.clone().requires_grad_(True) for t in (h, w_out, b_out, wide, w_d, b_d))
    loss = head_bce_loss(ha, wa, ba, wia, dense, wda, bda, labels)
    loss.backward()
    # fp32 oracle on the same (bf16-rounded) values; the oracle's dh is masked by hand (h is a ReLU output)
    hb, wb, bb, wib, wdb, bdb = (t.float().clone().requires_grad_(True) for t in (h, w_out, b_out, wide, w_d, b_d))
    z = hb @ wb.t().squeeze(1) + bb + wib + dense @ wdb.t().squeeze(1) + bdb
    ref = torch.nn.functional.binary_cross_entropy_with_logits(z, labels)
    ref.backward()
    torch.testing.assert_close(loss.float(), ref, atol=2e-3, rtol=2e-3)
    dh_ref = hb.grad * (h.float() > 0)
    torch.testing.assert_close(ha.grad.float(), dh_ref, atol=2e-5, rtol=3e-2)
    torch.testing.assert_close(wa.grad.float(), wb.grad, atol=2e-3, rtol=3e-2)
    torch.testing.assert_close(wda.grad.float(), wdb.grad, atol=2e-3, rtol=3e-2)
    torch.testing.assert_close(wia.grad, wib.grad, atol=1e-6, rtol=1e-3)
    torch.testing.assert_close(ba.grad.float(), bb.grad, atol=2e-3, rtol=3e-2)
    torch.testing.assert_close(bda.grad.float(), bdb.grad, atol=2e-3, rtol=3e-2)


def test_wide_deep_loss_fused_equals_unfused_path():
    """WideDeepDense.loss (fused head, tower told its top gradient is masked) == bce(forward(...)): value and grads."""
    from paddle_operator_b200.models.wide_deep import WideDeepDense, bce_with_logits

    torch.manual_seed(1)
    a = WideDeepDense(16, fused=True).cuda().to(torch.bfloat16)
    b = WideDeepDense(16, fused=True).cuda().to(torch.bfloat16)
    b.load_state_dict(a.state_dict())
    bsz = 512
    deep = torch.randn(bsz, 416, device="cuda").to(torch.bfloat16)
    wide = torch.randn(bsz, device="cuda")
    dense = torch.randn(bsz, 13, device="cuda")
    labels = (torch.rand(bsz, device="cuda") > 0.7).float()
    da, db = deep.clone().requires_grad_(True), deep.clone().requires_grad_(True)
    la = a.loss(da, wide, dense, labels)
    lb = bce_with_logits(b(db, wide, dense), labels)
    la.backward()
    lb.backward()
    torch.testing.assert_close(la.float(), lb.float(), atol=5e-3, rtol=5e-3)
    torch.testing.assert_close(da.grad.float(), db.grad.float(), atol=2e-5, rtol=0.1)
    for (n, pa), (_, pb) in zip(a.named_parameters(), b.named_parameters()):
        torch.testing.assert_close(pa.grad.float(), pb.grad.float(), atol=3e-3, rtol=0.1, msg=lambda m, n=n: f"{n}: {m}")
