"""CPU checks of the collective engine's host-side logic (flat buffers, buckets,
hooks, optimiser reference path) on a tiny ResNet."""
import torch

from paddle_operator_b200.models.resnet import resnet_tiny, to_training_dtype
from paddle_operator_b200.ops.optim import softmax_cross_entropy
from paddle_operator_b200.parallel.collective import CollectiveTrainer


def _loss(m, a, b):
    return softmax_cross_entropy(m(a), b)


def test_flat_layout_and_training_cpu():
    torch.manual_seed(0)
    model = to_training_dtype(resnet_tiny(num_classes=8, fused=True), torch.bfloat16).train()
    x = torch.randn(8, 3, 32, 32).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    y = torch.randint(0, 8, (8,))
    tr = CollectiveTrainer(model, _loss, (x, y), lr=0.05, bucket_bytes=4096, use_graph=False)
    flat = tr.flat
    assert len(flat.buckets) > 1
    assert flat.buckets[0].start == 0 and flat.buckets[-1].end == flat.numel
    for a, b in zip(flat.buckets, flat.buckets[1:]):
        assert a.end == b.start and a.start % 8 == 0
    # parameters are views into the flat buffer and keep channels-last strides
    for p in model.parameters():
        assert p.data_ptr() >= flat.param.data_ptr()
        assert p.data_ptr() < flat.param.data_ptr() + flat.numel * 2
        if p.dim() == 4:
            assert p.is_contiguous(memory_format=torch.channels_last)
    losses = [float(tr.step()) for _ in range(8)]
    assert losses[-1] < losses[0], losses
    assert all(b.pending == 0 for b in flat.buckets)
    # masters and bf16 params stay in sync
    torch.testing.assert_close(flat.param.float(), flat.master.to(torch.bfloat16).float(), atol=0, rtol=0)


def test_state_dict_roundtrip_cpu():
    torch.manual_seed(0)
    model = to_training_dtype(resnet_tiny(num_classes=8, fused=True), torch.bfloat16).train()
    x = torch.randn(4, 3, 32, 32).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    y = torch.randint(0, 8, (4,))
    tr = CollectiveTrainer(model, _loss, (x, y), use_graph=False)
    tr.step()
    sd = {k: (v.clone() if torch.is_tensor(v) else {n: b.clone() for n, b in v.items()}) for k, v in tr.state_dict().items()}
    l1 = float(tr.step())
    tr.load_state_dict(sd)
    l2 = float(tr.step())
    assert abs(l1 - l2) < 1e-6


def test_space_to_depth_stem_is_the_same_convolution():
    """7x7/2/pad3 over the image == 4x4/1/pad0 over the space-to-depth input with re-arranged weights."""
    from paddle_operator_b200.models.resnet import stem_s2d_mask, stem_weight_to_s2d
    from paddle_operator_b200.ops.optim import preprocess_u8_s2d, space_to_depth_reference

    torch.manual_seed(0)
    x = torch.randn(2, 3, 32, 48, dtype=torch.float64)
    w7 = torch.randn(5, 3, 7, 7, dtype=torch.float64)
    ref = torch.nn.functional.conv2d(x, w7, stride=2, padding=3)
    got = torch.nn.functional.conv2d(space_to_depth_reference(x), stem_weight_to_s2d(w7))
    assert got.shape == ref.shape
    torch.testing.assert_close(got, ref, atol=1e-10, rtol=1e-10)
    assert int(stem_s2d_mask(5).sum()) == 5 * 3 * 49
    # the uint8 pre-processing path (CPU branch) produces that layout
    img = torch.randint(0, 256, (2, 8, 12, 3), dtype=torch.uint8)
    y = preprocess_u8_s2d(img)
    assert y.shape == (2, 16, 7, 9) and float(y[:, 12:].abs().max()) == 0.0 and float(y[:, :, :2].abs().max()) == 0.0


def test_s2d_model_trains_and_keeps_structural_zeros():
    torch.manual_seed(0)
    model = to_training_dtype(resnet_tiny(num_classes=8, fused=True, s2d_stem=True), torch.bfloat16).train()
    x = torch.randn(4, 3, 32, 32).to(torch.bfloat16)
    y = torch.randint(0, 8, (4,))
    tr = CollectiveTrainer(model, _loss, (x, y), lr=0.05, use_graph=False)
    l0 = float(tr.step())
    for _ in range(5):
        l1 = float(tr.step())
    assert l1 < l0
    w = model.conv1.weight.float()
    assert float((w * (1 - model.stem_mask)).abs().max()) == 0.0


def test_resnet_gemm_1x1_falls_back_to_library_convolutions_on_cpu():
    """The tcgen05 convolution paths are CUDA/bf16 only: on CPU the same module graph must compute exactly
    what the plain model computes (same parameters, same outputs, same gradients)."""
    import torch

    from paddle_operator_b200.models.resnet import ResNet

    torch.manual_seed(0)
    a = ResNet((2, 1), 8, fused=True, width=16, gemm_1x1=False).train()
    b = ResNet((2, 1), 8, fused=True, width=16, gemm_1x1=True).train()
    b.load_state_dict(a.state_dict())
    x = torch.randn(2, 3, 32, 32)
    ya, yb = a(x), b(x)
    assert torch.allclose(ya, yb, atol=1e-5)
    ya.sum().backward()
    yb.sum().backward()
    for (n, pa), (_, pb) in zip(a.named_parameters(), b.named_parameters()):
        assert torch.allclose(pa.grad, pb.grad, atol=1e-4), n


def test_conv_ops_eligibility_rules():
    import torch

    from paddle_operator_b200.ops import conv as conv_ops

    x = torch.randn(2, 64, 8, 8).contiguous(memory_format=torch.channels_last)
    w = torch.randn(128, 64, 1, 1)
    assert not conv_ops.conv1x1_eligible(x, w)            # CPU / fp32
    y, partial = conv_ops.conv1x1_stats(x, w)
    assert partial is None and y.shape == (2, 128, 8, 8)
    y, partial, skip = conv_ops.conv1x1_stats(x, w, carry_skip=True)
    assert partial is None and skip is x
    y1, p1, yd, pd = conv_ops.conv1x1_dual(x, w, torch.randn(256, 64, 1, 1), 2)
    assert p1 is None and pd is None and yd.shape == (2, 256, 4, 4)
    ys, ps = conv_ops.stem_s2d_conv(torch.randn(1, 16, 19, 19), torch.randn(64, 16, 4, 4))
    assert ps is None and ys.shape == (1, 64, 16, 16)


def test_sink_scope_bookkeeping_without_a_gpu():
    """ops/_sink.py: the overlap scope is only open inside a trainer step and sinks are ignored for CPU parameters."""
    import torch

    from paddle_operator_b200.ops import _sink

    p = torch.nn.Parameter(torch.zeros(4, 4))
    p._grad_sink = torch.zeros(4, 4)
    assert _sink.sink_of(p) is None and _sink.sink_of(None) is None  # CPU tensors / wrong dtype never sink
    assert not _sink.overlap_enabled()
    _sink.begin_step(True)
    assert _sink.overlap_enabled() == _sink.WGRAD_OVERLAP
    _sink.begin_step(False)
    assert not _sink.overlap_enabled()
    _sink.end_step()
    assert not _sink.overlap_enabled() and not _sink._scope["events"] and not _sink._scope["keep"]
    g = torch.ones(4, 4)
    assert _sink.is_sunk(g, g) and not _sink.is_sunk(g.clone(), g) and not _sink.is_sunk(None, g)
