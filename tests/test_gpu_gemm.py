"""tcgen05 / TMA fused linear kernel vs a plain PyTorch fp32 reference."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("m,n,k", [(128, 128, 64), (256, 400, 400), (1000, 400, 432), (4096, 1, 400), (77, 24, 40),
                                   (32768, 400, 400), (513, 1000, 2048)])
@pytest.mark.parametrize("relu,use_bias", [(False, False), (True, True)])
def test_linear_bf16_matches_fp32(m, n, k, relu, use_bias):
    from paddle_operator_b200.ops.gemm import linear_bf16

    torch.manual_seed(m + n + k)
    k8 = (k + 7) // 8 * 8
    a = (torch.randn(m, k8, device="cuda") * 0.5).to(torch.bfloat16)[:, :k]
    b = (torch.randn(n, k8, device="cuda") * 0.5).to(torch.bfloat16)[:, :k]
    bias = torch.randn(n, device="cuda") if use_bias else None
    out = linear_bf16(a, b, bias, relu=relu)
    ref = a.float() @ b.float().t()
    if use_bias:
        ref = ref + bias
    if relu:
        ref = torch.relu(ref)
    assert out.shape == (m, n) and out.dtype == torch.bfloat16
    torch.testing.assert_close(out.float(), ref, atol=3e-2 * (k ** 0.5) * 0.25 + 2e-2, rtol=2e-2)


def test_linear_mask_epilogue_and_autograd():
    from paddle_operator_b200.ops.gemm import FusedLinear, linear_bf16

    torch.manual_seed(0)
    a = torch.randn(300, 400, device="cuda").to(torch.bfloat16)
    b = torch.randn(208, 400, device="cuda").to(torch.bfloat16) * 0.1
    y = torch.randn(300, 208, device="cuda").to(torch.bfloat16)
    out = linear_bf16(a, b, mask=y)
    ref = (a.float() @ b.float().t()) * (y.float() > 0)
    torch.testing.assert_close(out.float(), ref, atol=0.1, rtol=2e-2)

    lin = FusedLinear(432, 400, relu=True).cuda().to(torch.bfloat16)
    ref_lin = torch.nn.Linear(432, 400).cuda().float()
    with torch.no_grad():
        ref_lin.weight.copy_(lin.weight.float())
        ref_lin.bias.copy_(lin.bias.float())
    x = torch.randn(512, 432, device="cuda").to(torch.bfloat16).requires_grad_(True)
    xr = x.detach().float().requires_grad_(True)
    g = torch.randn(512, 400, device="cuda").to(torch.bfloat16)
    lin(x).backward(g)
    torch.relu(ref_lin(xr)).backward(g.float())
    torch.testing.assert_close(x.grad.float(), xr.grad, atol=0.15, rtol=3e-2)
    torch.testing.assert_close(lin.weight.grad.float(), ref_lin.weight.grad, atol=0.5, rtol=3e-2)
    torch.testing.assert_close(lin.bias.grad.float(), ref_lin.bias.grad, atol=0.5, rtol=3e-2)


def test_fused_mlp_matches_torch():
    from paddle_operator_b200.ops.mlp import FusedMLP, colsum

    torch.manual_seed(3)
    x = torch.randn(1000, 400, device="cuda").to(torch.bfloat16)
    torch.testing.assert_close(colsum(x), x.float().sum(0), atol=0.5, rtol=1e-2)
    mlp = FusedMLP([432, 400, 400, 400]).cuda().to(torch.bfloat16)
    ref = [torch.nn.Linear(a, b).cuda().float() for a, b in ((432, 400), (400, 400), (400, 400))]
    with torch.no_grad():
        for r, l in zip(ref, mlp.layers):
            r.weight.copy_(l.weight.float())
            r.bias.copy_(l.bias.float())
    xi = (torch.randn(2048, 432, device="cuda") * 0.5).to(torch.bfloat16).requires_grad_(True)
    xr = xi.detach().float().requires_grad_(True)
    g = (torch.randn(2048, 400, device="cuda") * 0.1).to(torch.bfloat16)
    y = mlp(xi)
    y.backward(g)
    h = xr
    for r in ref:
        h = torch.relu(r(h))
    h.backward(g.float())
    torch.testing.assert_close(y.float(), h, atol=0.08, rtol=3e-2)
    torch.testing.assert_close(xi.grad.float(), xr.grad, atol=0.05, rtol=5e-2)
    for r, l in zip(ref, mlp.layers):
        # ReLU gates that sit within bf16 rounding of zero flip between the two paths; each flip moves a
        # sum over 2048 rows by one gradient element (~0.1), hence the absolute tolerances
        torch.testing.assert_close(l.weight.grad.float(), r.weight.grad, atol=1.0, rtol=5e-2)
        torch.testing.assert_close(l.bias.grad.float(), r.bias.grad, atol=1.0, rtol=5e-2)
