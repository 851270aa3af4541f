"""Gradient sinks (ops/_sink.py) and the deterministic weighted column-sum kernel behind the bias / head gradients."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("m,n,ne", [(1000, 400, 13), (32768, 400, 0), (257, 1000, 0), (64, 8, 31), (5000, 432, 5)])
def test_colsum_w_matches_fp32_reference(m, n, ne):
    from paddle_operator_b200.ops.mlp import colsum_w

    torch.manual_seed(m + n)
    x = torch.randn(m, n, device="cuda").to(torch.bfloat16)
    w = torch.randn(m, device="cuda")
    g = torch.full((1,), 0.5, device="cuda")
    xe = torch.randn(m, ne, device="cuda") if ne else None
    a = torch.empty(n, dtype=torch.float32, device="cuda")
    b = torch.empty(n // 2, dtype=torch.bfloat16, device="cuda")  # a second window onto the same columns, as bf16
    e = torch.empty(max(ne, 1), dtype=torch.float32, device="cuda")
    o = torch.empty(1, dtype=torch.bfloat16, device="cuda")
    dests = [(a, 0), (b, n // 2), (o, n + ne)] + ([(e, n)] if ne else [])
    colsum_w(x, dests, weights=w, gscale=g, extra=xe, ones=True)
    ref = (0.5 * w)[:, None].double() * x.double()
    torch.testing.assert_close(a.double(), ref.sum(0), atol=2e-3 * m ** 0.5, rtol=1e-4)
    torch.testing.assert_close(b.double(), ref.sum(0)[n // 2:], atol=2e-3 * m ** 0.5, rtol=1e-2)
    torch.testing.assert_close(o.double(), (0.5 * w).double().sum().view(1), atol=0.05 * m ** 0.5, rtol=1e-2)
    if ne:
        torch.testing.assert_close(e.double(), ((0.5 * w)[:, None].double() * xe.double()).sum(0), atol=1e-3 * m ** 0.5, rtol=1e-4)
    # same bits on a second run (fixed reduction order), and the ticket is back at zero
    a2 = torch.empty_like(a)
    colsum_w(x, [(a2, 0)], weights=w, gscale=g, extra=xe, ones=True)
    assert torch.equal(a, a2)
    # unweighted form = bias gradient
    colsum_w(x, [(a2, 0)])
    torch.testing.assert_close(a2.double(), x.double().sum(0), atol=2e-3 * m ** 0.5, rtol=1e-4)


def _train(sink: bool, monkeypatch, steps: int = 3):
    from paddle_operator_b200.ops import _sink
    from paddle_operator_b200.ops.mlp import FusedMLP
    from paddle_operator_b200.parallel.collective import CollectiveTrainer

    monkeypatch.setattr(_sink, "ENABLED", sink)
    copies = []
    real = torch._foreach_copy_
    monkeypatch.setattr(torch, "_foreach_copy_", lambda d, s, *a, **k: (copies.append(len(d)), real(d, s, *a, **k))[1])
    torch.manual_seed(0)
    model = FusedMLP([64, 128, 64]).cuda().to(torch.bfloat16).train()
    x = torch.randn(256, 64, device="cuda").to(torch.bfloat16)
    y = torch.randn(256, 64, device="cuda")

    def loss_fn(m, xx, yy):
        return ((m(xx).float() - yy) ** 2).mean()

    tr = CollectiveTrainer(model, loss_fn, (x, y), comm=None, lr=0.05, momentum=0.9, weight_decay=0.0, use_graph=False)
    losses = [float(tr.step()) for _ in range(steps)]
    return losses, tr.flat.param.clone(), list(copies)


def test_trainer_gradients_land_in_the_flat_buffer_without_a_copy(monkeypatch):
    l1, p1, c1 = _train(True, monkeypatch)
    l0, p0, c0 = _train(False, monkeypatch)
    assert not c1, f"_foreach_copy_ still ran with sinks: {c1}"
    assert c0 and all(n > 0 for n in c0)
    assert l1 == l0 and torch.equal(p1, p0)  # same kernels, same bits: only the destination differs
    assert l1[-1] < l1[0]
