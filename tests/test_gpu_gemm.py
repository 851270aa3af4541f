"""tcgen05 / TMA fused linear kernel vs a plain PyTorch fp32 reference."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _conv_gemm_for_every_size(monkeypatch):
    """The product only routes outputs larger than L2 through the 1x1-convolution GEMM; the tests use small shapes."""
    from paddle_operator_b200.ops import conv as conv_ops

    monkeypatch.setattr(conv_ops, "MIN_OUTPUT_BYTES", 0)


@pytest.mark.parametrize("m,no,nc", [(32768, 400, 432), (1000, 400, 400), (4096, 128, 64), (777, 1000, 2048), (300, 64, 512)])
def test_wgrad_mn_also_returns_the_bias_gradient(m, no, nc):
    """db = column sums of dy from the weight-gradient kernel (extra N = 16 MMA against a tile of ones)."""
    from paddle_operator_b200.ops.gemm import wgrad_mn

    torch.manual_seed(m + no + nc)
    dy = (torch.randn(m, no, device="cuda") * 0.5).to(torch.bfloat16)
    x = (torch.randn(m, nc, device="cuda") * 0.5).to(torch.bfloat16)
    db = torch.full((no,), 7.0, dtype=torch.bfloat16, device="cuda")
    dw, filled = wgrad_mn(dy, x, db_out=db)
    assert filled
    ref_db = dy.float().sum(0)
    torch.testing.assert_close(db.float(), ref_db, atol=0.02 * m ** 0.5 + 0.05, rtol=1e-2)
    torch.testing.assert_close(dw.float(), dy.float().t() @ x.float(), atol=0.05 * m ** 0.5, rtol=2e-2)
    assert torch.equal(dw, wgrad_mn(dy, x))  # the weight gradient itself is unchanged, bit for bit


@pytest.mark.parametrize("m,no,nc", [(32768, 400, 432), (1000, 256, 64), (777, 1000, 2048), (4096, 512, 1000), (300, 136, 16)])
def test_wgrad_mn_cta_pair_kernel(m, no, nc, monkeypatch):
    """cta_group::2 version (off by default, profiles/r2/wgrad_pair.md): 256 output channels per pair, the X columns
    split between the two CTAs, two instructions per K step for tiles wider than 256 columns, ragged edges."""
    from paddle_operator_b200.ops.gemm import wgrad_mn

    monkeypatch.setenv("PDJ_WGRAD_PAIR", "1")  # read by the library on every call
    torch.manual_seed(m + 3 * no + nc)
    dy = (torch.randn(m, no, device="cuda") * 0.5).to(torch.bfloat16)
    x = (torch.randn(m, nc, device="cuda") * 0.5).to(torch.bfloat16)
    db = torch.empty(no, dtype=torch.bfloat16, device="cuda")
    dw, filled = wgrad_mn(dy, x, db_out=db)
    assert filled
    torch.testing.assert_close(dw.float(), dy.float().t() @ x.float(), atol=0.05 * m ** 0.5, rtol=2e-2)
    torch.testing.assert_close(db.float(), dy.float().sum(0), atol=0.02 * m ** 0.5 + 0.05, rtol=1e-2)
    torch.testing.assert_close(wgrad_mn(dy, x).float(), dw.float(), atol=0.02 * m ** 0.5, rtol=1e-2)
    monkeypatch.setenv("PDJ_WGRAD_PAIR", "0")
    torch.testing.assert_close(wgrad_mn(dy, x).float(), dw.float(), atol=0.02 * m ** 0.5, rtol=1e-2)


@pytest.mark.parametrize("m,n,k", [(128, 64, 64), (1000, 432, 400), (4096, 400, 400), (513, 216, 72), (300, 1000, 136)])
def test_linear_with_weight_as_stored(m, n, k):
    """b_kn: out = a @ b with b given as [K, N] (dX = dZ W) - the MN-major B operand, incl. the mask epilogue."""
    from paddle_operator_b200.ops import gemm

    assert gemm.KN
    torch.manual_seed(m + n * 7 + k)
    a = (torch.randn(m, k, device="cuda") * 0.5).to(torch.bfloat16)
    b = (torch.randn(k, n, device="cuda") * 0.5).to(torch.bfloat16)
    gate = (torch.randn(m, n, device="cuda") > 0).to(torch.bfloat16)
    ref = a.float() @ b.float()
    tol = dict(atol=3e-2 * (k ** 0.5) * 0.25 + 2e-2, rtol=2e-2)
    torch.testing.assert_close(gemm.linear_bf16(a, b, b_kn=True).float(), ref, **tol)
    torch.testing.assert_close(gemm.linear_bf16(a, b, mask=gate, b_kn=True).float(), ref * gate.float(), **tol)
    # same bits as the K-major path on the transposed copy: the products and the accumulation order are identical
    assert torch.equal(gemm.linear_bf16(a, b, b_kn=True), gemm.linear_bf16(a, b.t().contiguous()))


@pytest.mark.parametrize("m,n,k", [(784, 64, 256), (1568, 256, 64), (3136, 512, 128), (200, 2048, 512)])
def test_conv1x1_input_gradient_reads_the_weight_as_stored(m, n, k, monkeypatch):
    from paddle_operator_b200.ops import conv as conv_ops

    monkeypatch.setattr(conv_ops, "CONV_KN", True)  # off by default in the product (slower on the ResNet step)
    torch.manual_seed(m + n + k)
    dy = (torch.randn(m, k, device="cuda") * 0.5).to(torch.bfloat16)
    w = (torch.randn(k, n, device="cuda") * 0.1).to(torch.bfloat16)  # [Cout, Cin]
    skip = torch.randn(m, n, device="cuda").to(torch.bfloat16)
    ref = dy.float() @ w.float()
    torch.testing.assert_close(conv_ops.conv1x1_gemm_add(dy, w, None, b_kn=True).float(), ref, atol=6e-2, rtol=3e-2)
    got = conv_ops.conv1x1_gemm_add(dy, w, skip, b_kn=True)
    torch.testing.assert_close(got.float(), ref + skip.float(), atol=8e-2, rtol=3e-2)
    assert torch.equal(got, conv_ops.conv1x1_gemm_add(dy, w.t().contiguous(), skip))


@pytest.mark.parametrize("m,n,k", [(256, 128, 64), (300, 400, 432), (1000, 1000, 2048), (4096, 512, 1024), (2048, 48, 4096),
                                   (257, 16, 40)])
def test_linear_cta_pair_mode_matches_fp32(m, n, k, monkeypatch):
    """tcgen05 cta_group::2: a CTA pair owns a 256-row tile (ragged last pair, odd pair counts, masks, narrow N)."""
    from paddle_operator_b200.ops.gemm import linear_bf16

    monkeypatch.setenv("PDJ_LINEAR_PAIR", "1")  # read by the library on every call
    torch.manual_seed(m * 3 + n + k)
    k8 = (k + 7) // 8 * 8
    a = (torch.randn(m, k8, device="cuda") * 0.5).to(torch.bfloat16)[:, :k]
    b = (torch.randn(n, k8, device="cuda") * 0.5).to(torch.bfloat16)[:, :k]
    bias = torch.randn(n, device="cuda")
    n8 = (n + 7) // 8 * 8
    gate = (torch.randn(m, n8, device="cuda") > 0).to(torch.bfloat16)[:, :n]
    ref = a.float() @ b.float().t() + bias
    tol = dict(atol=3e-2 * (k ** 0.5) * 0.25 + 2e-2, rtol=2e-2)
    torch.testing.assert_close(linear_bf16(a, b, bias, relu=True).float(), torch.relu(ref), **tol)
    if n % 8 == 0:
        torch.testing.assert_close(linear_bf16(a, b, bias, mask=gate).float(), ref * gate.float(), **tol)
    monkeypatch.setenv("PDJ_LINEAR_PAIR", "0")
    torch.testing.assert_close(linear_bf16(a, b, bias, relu=True).float(), torch.relu(ref), **tol)


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


# ---- 1x1 convolution GEMM with BatchNorm statistics in the epilogue (ops/conv.py) ----------------
@pytest.mark.gpu
@pytest.mark.parametrize("n,hw,cin,cout", [(4, 14, 64, 64), (2, 28, 256, 128), (3, 7, 128, 512), (2, 9, 72, 192),
                                           (8, 56, 64, 256)])
def test_conv1x1_stats_matches_conv2d(n, hw, cin, cout):
    from paddle_operator_b200.ops import conv as conv_ops

    torch.manual_seed(cin + cout)
    x = torch.randn(n, cin, hw, hw, device="cuda", dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
    w = (torch.randn(cout, cin, 1, 1, device="cuda", dtype=torch.bfloat16) * 0.1).contiguous(memory_format=torch.channels_last)
    assert conv_ops.conv1x1_eligible(x, w)
    y, partial = conv_ops.conv1x1_stats(x, w)
    ref = torch.nn.functional.conv2d(x.float(), w.float())
    assert y.shape == ref.shape and y.is_contiguous(memory_format=torch.channels_last)
    torch.testing.assert_close(y.float(), ref, atol=2e-2, rtol=2e-2)
    # the partial rows add up to the sums of the values as stored (bf16-rounded)
    yf = y.float()
    s_ref, q_ref = yf.sum(dim=(0, 2, 3)), (yf * yf).sum(dim=(0, 2, 3))
    torch.testing.assert_close(partial[:, 0].sum(0), s_ref, atol=1e-4 * float(yf.abs().sum(dim=(0, 2, 3)).max()), rtol=1e-3)
    torch.testing.assert_close(partial[:, 1].sum(0), q_ref, atol=1e-3, rtol=1e-3)
    # bit-reproducible
    y2, partial2 = conv_ops.conv1x1_stats(x, w)
    assert torch.equal(y, y2) and torch.equal(partial, partial2)


@pytest.mark.gpu
def test_conv1x1_bn_chain_matches_reference_forward_backward():
    from paddle_operator_b200.ops import bn as bn_ops
    from paddle_operator_b200.ops import conv as conv_ops

    torch.manual_seed(5)
    n, hw, cin, cout = 4, 14, 64, 128
    x = torch.randn(n, cin, hw, hw, device="cuda", dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
    res = torch.randn(n, cout, hw, hw, device="cuda", dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
    w = (torch.randn(cout, cin, 1, 1, device="cuda", dtype=torch.bfloat16) * 0.1).contiguous(memory_format=torch.channels_last)
    gamma = torch.rand(cout, device="cuda") + 0.5
    beta = torch.randn(cout, device="cuda") * 0.1
    xa, wa, ga, ba = (t.clone().requires_grad_(True) for t in (x, w, gamma, beta))
    rm, rv = torch.zeros(cout, device="cuda"), torch.ones(cout, device="cuda")
    y, partial = conv_ops.conv1x1_stats(xa, wa)
    z = bn_ops.bn_act(y, ga, ba, res, rm, rv, relu=True, partial=partial)
    # same chain with the statistics pass over y (the path without fusion): identical statistics source values
    xb, wb, gb, bb = (t.clone().requires_grad_(True) for t in (x, w, gamma, beta))
    rm2, rv2 = torch.zeros(cout, device="cuda"), torch.ones(cout, device="cuda")
    y_b = torch.nn.functional.conv2d(xb, wb)
    z_b = bn_ops.bn_act(y_b, gb, bb, res, rm2, rv2, relu=True)
    torch.testing.assert_close(z.float(), z_b.float(), atol=3e-2, rtol=3e-2)
    torch.testing.assert_close(rm, rm2, atol=1e-3, rtol=1e-3)
    torch.testing.assert_close(rv, rv2, atol=1e-3, rtol=1e-3)
    # fp32 oracle
    ref = bn_ops.bn_act_reference(torch.nn.functional.conv2d(x.float(), w.float()), gamma, beta, res, relu=True)
    torch.testing.assert_close(z.float(), ref, atol=6e-2, rtol=6e-2)
    g = torch.randn_like(z)
    z.backward(g)
    z_b.backward(g)
    torch.testing.assert_close(xa.grad.float(), xb.grad.float(), atol=3e-2, rtol=3e-2)
    torch.testing.assert_close(wa.grad.float(), wb.grad.float(), atol=0.3, rtol=5e-2)
    torch.testing.assert_close(ga.grad, gb.grad, atol=0.5, rtol=5e-2)


@pytest.mark.parametrize("n,hw,cin,cout", [(4, 14, 64, 16), (2, 28, 256, 64), (3, 7, 512, 128), (2, 9, 192, 72)])
def test_conv1x1_skip_gradient_is_folded_into_dgrad(n, hw, cin, cout):
    """y = conv(x) + skip path on x: the fused backward (one GEMM with addend) equals autograd's conv dgrad + add."""
    from paddle_operator_b200.ops import conv as conv_ops

    torch.manual_seed(cin * 3 + cout)
    x = torch.randn(n, cin, hw, hw, device="cuda", dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
    w = (torch.randn(cout, cin, 1, 1, device="cuda", dtype=torch.bfloat16) * 0.1).contiguous(memory_format=torch.channels_last)
    gy = torch.randn(n, cout, hw, hw, device="cuda", dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
    gs = torch.randn(n, cin, hw, hw, device="cuda", dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
    xa, wa = x.clone().requires_grad_(True), w.clone().requires_grad_(True)
    y, partial, skip = conv_ops.conv1x1_stats(xa, wa, carry_skip=True)
    assert skip.data_ptr() == xa.data_ptr()
    torch.autograd.backward([y, skip], [gy, gs])
    xb, wb = x.float().requires_grad_(True), w.float().requires_grad_(True)
    yb = torch.nn.functional.conv2d(xb, wb)
    torch.autograd.backward([yb, xb * 1.0], [gy.float(), gs.float()])
    torch.testing.assert_close(y.float(), yb, atol=3e-2, rtol=3e-2)
    torch.testing.assert_close(xa.grad.float(), xb.grad, atol=4e-2, rtol=3e-2)
    torch.testing.assert_close(wa.grad.float(), wb.grad, atol=0.25, rtol=5e-2)


@pytest.mark.parametrize("n,hw,cin,cout", [(4, 14, 64, 256), (2, 28, 128, 512), (3, 7, 512, 2048), (2, 9, 192, 128)])
def test_conv1x1_backward_without_skip_runs_on_own_kernels(n, hw, cin, cout):
    """conv3 of a bottleneck (no skip gradient to fold in): dgrad GEMM + MN-major wgrad vs an fp32 reference."""
    from paddle_operator_b200.ops import conv as conv_ops

    assert conv_ops.OWN_BWD
    torch.manual_seed(cin + cout)
    x = torch.randn(n, cin, hw, hw, device="cuda", dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
    w = (torch.randn(cout, cin, 1, 1, device="cuda", dtype=torch.bfloat16) * 0.1).contiguous(memory_format=torch.channels_last)
    gy = torch.randn(n, cout, hw, hw, device="cuda", dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
    xa, wa = x.clone().requires_grad_(True), w.clone().requires_grad_(True)
    before = conv_ops.launch_count()
    y, _ = conv_ops.conv1x1_stats(xa, wa)
    y.backward(gy)
    assert conv_ops.launch_count() - before >= 3  # wgrad (+reduce) and the dgrad GEMM ran here, not in the library
    xb, wb = x.float().requires_grad_(True), w.float().requires_grad_(True)
    torch.nn.functional.conv2d(xb, wb).backward(gy.float())
    torch.testing.assert_close(xa.grad.float(), xb.grad, atol=6e-2, rtol=3e-2)
    torch.testing.assert_close(wa.grad.float(), wb.grad, atol=0.25, rtol=5e-2)


def test_conv1x1_gemm_add_in_place():
    from paddle_operator_b200.ops import conv as conv_ops

    torch.manual_seed(11)
    m, n, k = 1000, 320, 136
    a = (torch.randn(m, k, device="cuda") * 0.3).to(torch.bfloat16)
    b = (torch.randn(n, k, device="cuda") * 0.3).to(torch.bfloat16)
    c = torch.randn(m, n, device="cuda").to(torch.bfloat16)
    ref = c.float() + a.float() @ b.float().t()
    out = conv_ops.conv1x1_gemm_add(a, b, c)
    torch.testing.assert_close(out.float(), ref, atol=6e-2, rtol=2e-2)
    c2 = c.clone()
    conv_ops.conv1x1_gemm_add(a, b, c2, out=c2)
    assert torch.equal(c2, out)


@pytest.mark.parametrize("stride,cin,c1,cd", [(2, 64, 32, 128), (1, 64, 64, 256), (2, 256, 128, 512)])
def test_conv1x1_dual_backward_sums_both_input_gradients(stride, cin, c1, cd):
    from paddle_operator_b200.ops import conv as conv_ops

    torch.manual_seed(cin + c1 + cd + stride)
    n, hw = 3, 12
    cl = torch.channels_last
    x = torch.randn(n, cin, hw, hw, device="cuda", dtype=torch.bfloat16).contiguous(memory_format=cl)
    w1 = (torch.randn(c1, cin, 1, 1, device="cuda", dtype=torch.bfloat16) * 0.1).contiguous(memory_format=cl)
    wd = (torch.randn(cd, cin, 1, 1, device="cuda", dtype=torch.bfloat16) * 0.1).contiguous(memory_format=cl)
    xa, w1a, wda = (t.clone().requires_grad_(True) for t in (x, w1, wd))
    y1, p1, yd, pd = conv_ops.conv1x1_dual(xa, w1a, wda, stride)
    xb, w1b, wdb = (t.float().requires_grad_(True) for t in (x, w1, wd))
    r1 = torch.nn.functional.conv2d(xb, w1b)
    rd = torch.nn.functional.conv2d(xb, wdb, stride=stride)
    torch.testing.assert_close(y1.float(), r1, atol=3e-2, rtol=3e-2)
    torch.testing.assert_close(yd.float(), rd, atol=3e-2, rtol=3e-2)
    g1 = torch.randn_like(y1)
    gd = torch.randn_like(yd)
    torch.autograd.backward([y1, yd], [g1, gd])
    torch.autograd.backward([r1, rd], [g1.float(), gd.float()])
    torch.testing.assert_close(xa.grad.float(), xb.grad, atol=6e-2, rtol=3e-2)
    torch.testing.assert_close(w1a.grad.float(), w1b.grad, atol=0.25, rtol=5e-2)
    torch.testing.assert_close(wda.grad.float(), wdb.grad, atol=0.25, rtol=5e-2)


@pytest.mark.parametrize("n,img", [(2, 64), (3, 224), (1, 40)])
def test_stem_s2d_conv_matches_conv2d(n, img):
    """4x4 stem over the space-to-depth input through TMA windows == F.conv2d; partial sums == sums of y."""
    from paddle_operator_b200.ops import conv as conv_ops

    torch.manual_seed(img)
    hi = img // 2 + 3
    x = torch.randn(n, 16, hi, hi, device="cuda", dtype=torch.bfloat16).contiguous(memory_format=torch.channels_last)
    w = (torch.randn(64, 16, 4, 4, device="cuda", dtype=torch.bfloat16) * 0.1).contiguous(memory_format=torch.channels_last)
    assert conv_ops.stem_s2d_eligible(x, w)
    xa, wa = x.clone().requires_grad_(True), w.clone().requires_grad_(True)
    y, partial = conv_ops.stem_s2d_conv(xa, wa)
    ref = torch.nn.functional.conv2d(x.float(), w.float())
    assert y.shape == ref.shape
    torch.testing.assert_close(y.float(), ref, atol=5e-2, rtol=3e-2)
    yf = y.float()
    torch.testing.assert_close(partial[:, 0].sum(0), yf.sum(dim=(0, 2, 3)), atol=1e-4 * float(yf.abs().sum(dim=(0, 2, 3)).max()),
                               rtol=1e-3)
    torch.testing.assert_close(partial[:, 1].sum(0), (yf * yf).sum(dim=(0, 2, 3)), atol=1e-3, rtol=1e-3)
    g = torch.randn_like(y)
    y.backward(g)
    wb = w.float().requires_grad_(True)
    torch.nn.functional.conv2d(x.float(), wb).backward(g.float())
    torch.testing.assert_close(wa.grad.float(), wb.grad, atol=0.5, rtol=5e-2)


@pytest.mark.parametrize("m,no,nc", [(32768, 400, 432), (32768, 400, 400), (1000, 64, 256), (50176, 256, 1024),
                                     (12544, 2048, 512), (300, 8, 8), (4096, 1000, 2048), (802816 // 4, 256, 64)])
def test_wgrad_mn_matches_fp32(m, no, nc):
    from paddle_operator_b200.ops.gemm import wgrad_mn

    torch.manual_seed(m + no)
    dy = (torch.randn(m, no, device="cuda") * 0.1).to(torch.bfloat16)
    x = torch.randn(m, nc + 8, device="cuda").to(torch.bfloat16)[:, :nc]  # a row pitch wider than the row
    dw = wgrad_mn(dy, x)
    ref = dy.float().t() @ x.float()
    assert dw.shape == (no, nc) and dw.dtype == torch.bfloat16
    torch.testing.assert_close(dw.float(), ref, atol=1e-2 * float(ref.abs().max()) + 1e-3, rtol=2e-2)
