"""tcgen05 3x3-convolution kernels vs plain PyTorch fp32 references (every ResNet-50 bottleneck shape)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

# (H = W, Cin = Cout) of the stride-1 3x3 convolutions of ResNet-50, plus odd sizes that exercise ragged row
# chunks, a reduction length that is not a multiple of 16 and non-square channel counts
SHAPES = [(56, 64, 64), (28, 128, 128), (14, 256, 256), (7, 512, 512), (20, 32, 192), (9, 96, 64), (33, 64, 128)]


@pytest.mark.parametrize("hw,cin,cout", SHAPES)
@pytest.mark.parametrize("batch", [3, 16])
def test_conv3x3_wgrad_matches_fp32(hw, cin, cout, batch):
    from paddle_operator_b200.ops.conv import conv3x3_wgrad, conv3x3_wgrad_eligible

    torch.manual_seed(hw * 131 + cin + batch)
    cl = torch.channels_last
    x = torch.randn(batch, cin, hw, hw + (1 if hw == 20 else 0), device="cuda").to(torch.bfloat16).contiguous(memory_format=cl)
    dy = (torch.randn(batch, cout, x.shape[2], x.shape[3], device="cuda") * 0.1).to(torch.bfloat16).contiguous(memory_format=cl)
    w = torch.zeros(cout, cin, 3, 3, device="cuda", dtype=torch.bfloat16).contiguous(memory_format=cl)
    assert conv3x3_wgrad_eligible(x, w)
    dw = conv3x3_wgrad(x, dy, w.shape)
    ref = torch.nn.grad.conv2d_weight(x.float(), w.shape, dy.float(), padding=1)
    assert dw.shape == ref.shape and dw.is_contiguous(memory_format=cl)
    scale = float(ref.abs().max())
    torch.testing.assert_close(dw.float(), ref, atol=1e-2 * scale + 1e-3, rtol=2e-2)


@pytest.mark.parametrize("hw,cin,cout,stride", [(56, 64, 64, 1), (28, 128, 128, 1), (14, 256, 256, 1), (7, 512, 512, 1),
                                                (56, 128, 128, 2), (28, 256, 256, 2), (14, 512, 512, 2), (17, 64, 192, 1),
                                                (15, 128, 64, 2)])
def test_conv3x3_fprop_and_stats_match_fp32(hw, cin, cout, stride):
    from paddle_operator_b200.ops.conv import conv_im2col

    torch.manual_seed(hw + cin)
    cl = torch.channels_last
    x = torch.randn(5, cin, hw, hw, device="cuda").to(torch.bfloat16).contiguous(memory_format=cl)
    w = (torch.randn(cout, cin, 3, 3, device="cuda") * 0.05).to(torch.bfloat16).contiguous(memory_format=cl)
    y, partial = conv_im2col(x, w, stride, 1, stats=True)
    ref = torch.nn.functional.conv2d(x.float(), w.float(), stride=stride, padding=1)
    assert y.shape == ref.shape and y.is_contiguous(memory_format=cl)
    torch.testing.assert_close(y.float(), ref, atol=3e-2, rtol=2e-2)
    s = partial.sum(0)  # [2, C]: sum and sum of squares of the bf16-rounded output
    yf = y.float()
    torch.testing.assert_close(s[0], yf.sum((0, 2, 3)), atol=1e-2 * yf.abs().sum((0, 2, 3)).max().item() / 100 + 1e-2, rtol=1e-3)
    torch.testing.assert_close(s[1], (yf * yf).sum((0, 2, 3)), atol=1e-2, rtol=1e-3)


def test_strided_1x1_through_im2col():
    """The projection shortcut of a down-sampling block: 1x1, stride 2."""
    from paddle_operator_b200.ops.conv import conv_im2col

    torch.manual_seed(9)
    cl = torch.channels_last
    x = torch.randn(4, 256, 28, 28, device="cuda").to(torch.bfloat16).contiguous(memory_format=cl)
    w = (torch.randn(512, 256, 1, 1, device="cuda") * 0.05).to(torch.bfloat16).contiguous(memory_format=cl)
    y, _ = conv_im2col(x, w, 2, 0)
    torch.testing.assert_close(y.float(), torch.nn.functional.conv2d(x.float(), w.float(), stride=2), atol=3e-2, rtol=2e-2)


@pytest.mark.parametrize("hw,cin,cout", [(56, 64, 64), (28, 128, 128), (14, 256, 256), (7, 512, 512), (11, 64, 128)])
def test_conv3x3_dgrad_matches_fp32(hw, cin, cout):
    from paddle_operator_b200.ops.conv import conv3x3_dgrad

    torch.manual_seed(hw)
    cl = torch.channels_last
    dy = torch.randn(3, cout, hw, hw, device="cuda").to(torch.bfloat16).contiguous(memory_format=cl)
    w = (torch.randn(cout, cin, 3, 3, device="cuda") * 0.05).to(torch.bfloat16).contiguous(memory_format=cl)
    dx = conv3x3_dgrad(dy, w)
    ref = torch.nn.grad.conv2d_input((3, cin, hw, hw), w.float(), dy.float(), padding=1)
    torch.testing.assert_close(dx.float(), ref, atol=5e-2, rtol=2e-2)


def test_conv3x3_autograd_matches_library():
    from paddle_operator_b200.ops.conv import conv3x3 as _conv3x3

    def conv3x3(x, w):
        return _conv3x3(x, w)[0]

    torch.manual_seed(5)
    cl = torch.channels_last
    x = torch.randn(4, 64, 28, 28, device="cuda").to(torch.bfloat16).contiguous(memory_format=cl).requires_grad_(True)
    w = (torch.randn(128, 64, 3, 3, device="cuda") * 0.05).to(torch.bfloat16).contiguous(memory_format=cl).requires_grad_(True)
    g = torch.randn(4, 128, 28, 28, device="cuda").to(torch.bfloat16).contiguous(memory_format=cl)
    conv3x3(x, w).backward(g)
    xr, wr = x.detach().float().requires_grad_(True), w.detach().float().requires_grad_(True)
    torch.nn.functional.conv2d(xr, wr, padding=1).backward(g.float())
    torch.testing.assert_close(x.grad.float(), xr.grad, atol=0.1, rtol=3e-2)
    torch.testing.assert_close(w.grad.float(), wr.grad, atol=0.02 * float(wr.grad.abs().max()) + 1e-3, rtol=3e-2)
