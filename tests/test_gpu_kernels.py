"""GPU numerics: every hand-written kernel against a plain PyTorch fp32
reference of the same op.  Multi-rank collectives are exercised on ONE GPU by
the emulation mode (all ranks in one cooperative launch over per-rank arenas on
the same device), as the profiling guide prescribes for boxes with fewer GPUs
than ranks."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda", 0)


def _nhwc(n, c, h, w, dev, scale=1.0, shift=0.0):
    x = (torch.randn(n, c, h, w, device=dev) * scale + shift).to(torch.bfloat16)
    return x.contiguous(memory_format=torch.channels_last)


@pytest.mark.parametrize("shape", [(8, 64, 56, 56), (4, 256, 14, 14), (2, 2048, 7, 7), (3, 8, 5, 7), (16, 512, 28, 28)])
@pytest.mark.parametrize("relu,res", [(True, False), (True, True), (False, False), (False, True)])
def test_bn_act_fwd_bwd(dev, shape, relu, res):
    from paddle_operator_b200.ops.bn import bn_act, bn_act_reference

    torch.manual_seed(0)
    n, c, h, w = shape
    x = _nhwc(n, c, h, w, dev, 1.5, 0.3).requires_grad_(True)
    r = _nhwc(n, c, h, w, dev).requires_grad_(True) if res else None
    g = (torch.rand(c, device=dev) + 0.5).requires_grad_(True)
    b = (torch.randn(c, device=dev) * 0.1).requires_grad_(True)
    rm, rv = torch.zeros(c, device=dev), torch.ones(c, device=dev)
    y = bn_act(x, g, b, r, rm, rv, relu=relu)
    dy = _nhwc(n, c, h, w, dev)
    y.backward(dy)

    xr = x.detach().float().requires_grad_(True)
    rr = r.detach().float().requires_grad_(True) if res else None
    gr, br = g.detach().clone().requires_grad_(True), b.detach().clone().requires_grad_(True)
    yr = bn_act_reference(xr, gr, br, rr, relu)
    yr.backward(dy.float())

    assert y.dtype == torch.bfloat16 and y.is_contiguous(memory_format=torch.channels_last)
    torch.testing.assert_close(y.float(), yr, atol=2e-2, rtol=2e-2)
    torch.testing.assert_close(x.grad.float(), xr.grad, atol=3e-2, rtol=3e-2)
    m = n * h * w
    tol = 2e-2 * (m ** 0.5)
    torch.testing.assert_close(g.grad, gr.grad, atol=tol, rtol=2e-2)
    torch.testing.assert_close(b.grad, br.grad, atol=tol, rtol=2e-2)
    if res:
        torch.testing.assert_close(r.grad.float(), rr.grad, atol=2e-2, rtol=2e-2)
    # running statistics follow torch's convention (momentum 0.1, unbiased variance)
    xm = x.detach().float()
    torch.testing.assert_close(rm, 0.1 * xm.mean(dim=(0, 2, 3)), atol=1e-3, rtol=1e-3)
    torch.testing.assert_close(rv, 0.9 + 0.1 * xm.var(dim=(0, 2, 3), unbiased=True), atol=2e-3, rtol=2e-3)


@pytest.mark.parametrize("shape,k,s,p", [((4, 64, 112, 112), 3, 2, 1), ((2, 16, 9, 11), 3, 2, 1), ((2, 8, 8, 8), 2, 2, 0),
                                          ((1, 24, 13, 7), 3, 1, 1)])
def test_maxpool_nhwc(dev, shape, k, s, p):
    from paddle_operator_b200.ops.pool import max_pool2d_nhwc

    torch.manual_seed(3)
    x = _nhwc(*shape, dev).requires_grad_(True)
    y = max_pool2d_nhwc(x, k, s, p)
    xr = x.detach().float().requires_grad_(True)
    yr = torch.nn.functional.max_pool2d(xr, k, s, p)
    assert y.shape == yr.shape and y.is_contiguous(memory_format=torch.channels_last)
    torch.testing.assert_close(y.float(), yr, atol=0, rtol=0)
    dy = torch.randn_like(yr).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
    y.backward(dy)
    yr.backward(dy.float())
    # ties are rare with random bf16 data but possible: compare sums exactly and values where unique
    torch.testing.assert_close(x.grad.float(), xr.grad, atol=2e-2, rtol=2e-2)


def test_sgd_momentum(dev):
    from paddle_operator_b200.ops.optim import sgd_momentum_

    torch.manual_seed(1)
    n = 8 * 1000 + 8 * 37
    w = torch.randn(n, device=dev)
    v = torch.randn(n, device=dev) * 0.1
    g = (torch.randn(n, device=dev) * 0.01).to(torch.bfloat16)
    p = torch.zeros(n, device=dev, dtype=torch.bfloat16)
    hyper = torch.tensor([0.1, 0.9, 1e-4, 1.0], device=dev)
    w0, v0 = w.clone(), v.clone()
    sgd_momentum_(w, v, g, p, hyper, n)
    gg = g.float() + 1e-4 * w0
    v1 = 0.9 * v0 + gg
    w1 = w0 - 0.1 * v1
    torch.testing.assert_close(v, v1, atol=1e-6, rtol=1e-5)
    torch.testing.assert_close(w, w1, atol=1e-6, rtol=1e-5)
    torch.testing.assert_close(p.float(), w1.to(torch.bfloat16).float(), atol=0, rtol=0)


def test_preprocess_u8(dev):
    from paddle_operator_b200.ops.optim import IMAGENET_MEAN, IMAGENET_STD, preprocess_u8_nhwc

    x = torch.randint(0, 256, (3, 17, 19, 3), dtype=torch.uint8, device=dev)
    y = preprocess_u8_nhwc(x)
    ref = (x.float() / 255.0 - torch.tensor(IMAGENET_MEAN, device=dev)) / torch.tensor(IMAGENET_STD, device=dev)
    assert y.shape == (3, 3, 17, 19)
    torch.testing.assert_close(y.permute(0, 2, 3, 1).float(), ref, atol=2e-2, rtol=1e-2)


def test_preprocess_u8_s2d(dev):
    from paddle_operator_b200.ops.optim import IMAGENET_MEAN, IMAGENET_STD, preprocess_u8_s2d, space_to_depth_reference

    x = torch.randint(0, 256, (3, 16, 20, 3), dtype=torch.uint8, device=dev)
    y = preprocess_u8_s2d(x)
    norm = (x.float() / 255.0 - torch.tensor(IMAGENET_MEAN, device=dev)) / torch.tensor(IMAGENET_STD, device=dev)
    ref = space_to_depth_reference(norm.permute(0, 3, 1, 2))
    assert y.shape == ref.shape == (3, 16, 11, 13)
    torch.testing.assert_close(y.float(), ref, atol=2e-2, rtol=1e-2)


def test_softmax_xent(dev):
    from paddle_operator_b200.ops.optim import softmax_cross_entropy

    torch.manual_seed(2)
    logits = (torch.randn(64, 1000, device=dev) * 3).to(torch.bfloat16).requires_grad_(True)
    labels = torch.randint(0, 1000, (64,), device=dev)
    loss = softmax_cross_entropy(logits, labels)
    loss.backward()
    lr = logits.detach().float().requires_grad_(True)
    ref = torch.nn.functional.cross_entropy(lr, labels)
    ref.backward()
    torch.testing.assert_close(loss, ref, atol=1e-3, rtol=1e-3)
    torch.testing.assert_close(logits.grad.float(), lr.grad, atol=1e-4, rtol=2e-2)


# ---------------------------------------------------------------------------
@pytest.fixture(scope="module")
def comms(dev):
    from paddle_operator_b200.parallel.symm import Communicator

    made = {}

    def get(n):
        if n not in made:
            made[n] = Communicator.emulated(n, arena_bytes=48 << 20)
        return made[n]

    yield get
    for c in made.values():
        c.close()


def _fill(comm, off, n, dtype=torch.bfloat16, seed=0):
    torch.manual_seed(seed)
    vals = []
    for r in range(comm.nranks):
        t = comm.tensor(off, (n,), dtype, rank=r)
        v = torch.randn(n, device=t.device).to(dtype)
        t.copy_(v)
        vals.append(v)
    return vals


@pytest.mark.parametrize("nranks", [2, 4, 8])
@pytest.mark.parametrize("n", [8, 8 * 1031, 1 << 20])
def test_allreduce_twoshot_inplace(comms, nranks, n):
    from paddle_operator_b200.parallel.symm import AR_TWOSHOT

    comm = comms(nranks)
    comm.reset_allocations()
    off = comm.alloc(n * 2)
    for it in range(2):  # twice: epochs must keep working across launches
        vals = _fill(comm, off, n, seed=it)
        ref = torch.stack([v.float() for v in vals]).sum(0) / nranks
        comm.all_reduce_bf16(off, off, n, scale=1.0 / nranks, algo=AR_TWOSHOT, nblocks=8)
        torch.cuda.synchronize()
        for r in range(nranks):
            got = comm.tensor(off, (n,), torch.bfloat16, rank=r).float()
            torch.testing.assert_close(got, ref, atol=2e-2, rtol=2e-2)
        # all ranks hold bit-identical results
        g0 = comm.tensor(off, (n,), torch.bfloat16, rank=0)
        for r in range(1, nranks):
            assert torch.equal(g0, comm.tensor(off, (n,), torch.bfloat16, rank=r))


@pytest.mark.parametrize("nranks", [2, 4, 8])
def test_allreduce_sgd_fused_matches_reduce_then_sgd(comms, nranks):
    """K1b: reduce-scatter + SGD-momentum + master / parameter broadcast in one kernel, on a sub-range (a bucket)
    of the flat buffers, against all-reduce followed by the plain update; two steps in a row."""
    comm = comms(nranks)
    comm.reset_allocations()
    n, first, cnt = 8 * 2000, 8 * 300, 8 * 1203  # bucket = [first, first + cnt)
    g_off, p_off = comm.alloc(n * 2), comm.alloc(n * 2)
    m_off, v_off = comm.alloc(n * 4), comm.alloc(n * 4)
    dev = comm.tensor(g_off, (1,), torch.bfloat16).device
    hyper = torch.tensor([0.1, 0.9, 1e-2, 1.0], device=dev)
    torch.manual_seed(1)
    w0 = torch.randn(n, device=dev)
    for r in range(nranks):
        comm.tensor(m_off, (n,), torch.float32, rank=r).copy_(w0)
        comm.tensor(v_off, (n,), torch.float32, rank=r).zero_()
        comm.tensor(p_off, (n,), torch.bfloat16, rank=r).copy_(w0.to(torch.bfloat16))
    w_ref, v_ref = w0.clone(), torch.zeros(n, device=dev)
    wd_split = first + 8 * 600  # decay on the first part of the bucket only
    for it in range(2):
        grads = _fill(comm, g_off, n, seed=10 + it)
        gsum = torch.stack([g.float() for g in grads]).sum(0) / nranks
        sl = slice(first, first + cnt)
        dec = torch.zeros(n, device=dev)
        dec[:wd_split] = 1e-2
        gj = dec[sl] * w_ref[sl] + gsum[sl]
        v_ref[sl] = 0.9 * v_ref[sl] + gj
        w_ref[sl] = w_ref[sl] - 0.1 * v_ref[sl]
        comm.allreduce_sgd(g_off, p_off, m_off, v_off, first, cnt, wd_split, hyper.data_ptr(), 1.0 / nranks, nblocks=6)
        torch.cuda.synchronize()
        for r in range(nranks):
            m = comm.tensor(m_off, (n,), torch.float32, rank=r)
            torch.testing.assert_close(m, w_ref, atol=1e-5, rtol=1e-5)
            assert torch.equal(comm.tensor(p_off, (n,), torch.bfloat16, rank=r), m.to(torch.bfloat16))
        # masters and parameters are bit-identical on every rank; untouched elements stayed untouched
        m0 = comm.tensor(m_off, (n,), torch.float32, rank=0)
        for r in range(1, nranks):
            assert torch.equal(m0, comm.tensor(m_off, (n,), torch.float32, rank=r))
        assert torch.equal(m0[:first], w0[:first]) and torch.equal(m0[first + cnt:], w0[first + cnt:])
        # momentum: every element's owner holds the reference value
        per = (cnt // 8 + nranks - 1) // nranks * 8
        for r in range(nranks):
            a, b = first + r * per, min(first + (r + 1) * per, first + cnt)
            if b > a:
                torch.testing.assert_close(comm.tensor(v_off, (n,), torch.float32, rank=r)[a:b], v_ref[a:b], atol=1e-5, rtol=1e-5)
    assert comm.error_flags() == [0] * comm.nranks


@pytest.mark.parametrize("nranks", [2, 8])
def test_allreduce_oneshot_and_f32(comms, nranks):
    from paddle_operator_b200.parallel.symm import AR_ONESHOT, AR_TWOSHOT

    comm = comms(nranks)
    comm.reset_allocations()
    n = 8 * 513
    a = comm.alloc(n * 2)
    b = comm.alloc(n * 2)
    c = comm.alloc(n * 4)
    vals = _fill(comm, a, n, seed=5)
    ref = torch.stack([v.float() for v in vals]).sum(0)
    comm.all_reduce_bf16(a, b, n, scale=1.0, algo=AR_ONESHOT, nblocks=4)
    comm.all_reduce_bf16(a, c, n, scale=0.5, algo=AR_TWOSHOT, out_f32=True, nblocks=4)
    torch.cuda.synchronize()
    for r in range(nranks):
        torch.testing.assert_close(comm.tensor(b, (n,), torch.bfloat16, rank=r).float(), ref, atol=5e-2, rtol=2e-2)
        torch.testing.assert_close(comm.tensor(c, (n,), torch.float32, rank=r), ref * 0.5, atol=1e-5, rtol=1e-5)
        assert torch.equal(comm.tensor(a, (n,), torch.bfloat16, rank=r), vals[r])  # input untouched


@pytest.mark.parametrize("nranks", [2, 4, 8])
def test_alltoall_broadcast_barrier(comms, nranks):
    comm = comms(nranks)
    comm.reset_allocations()
    chunk = 16 * 129  # bytes
    a = comm.alloc(chunk * nranks)
    b = comm.alloc(chunk * nranks)
    src = []
    for r in range(nranks):
        t = comm.tensor(a, (nranks, chunk), torch.uint8, rank=r)
        v = torch.randint(0, 256, (nranks, chunk), dtype=torch.uint8, device=t.device)
        t.copy_(v)
        src.append(v)
    comm.all_to_all(a, b, chunk, nblocks=4)
    comm.barrier()
    torch.cuda.synchronize()
    for r in range(nranks):
        got = comm.tensor(b, (nranks, chunk), torch.uint8, rank=r)
        for p in range(nranks):
            assert torch.equal(got[p], src[p][r])
    root = nranks - 1
    comm.broadcast(root, a, chunk * nranks, nblocks=4)
    torch.cuda.synchronize()
    for r in range(nranks):
        assert torch.equal(comm.tensor(a, (nranks, chunk), torch.uint8, rank=r), src[root])
    assert all(e == 0 for e in comm.error_flags())


def test_device_barrier_and_allreduce_stress(comms):
    """a stress test of the device barrier.  500 back-to-back barriers and 200 in-place
    all-reduces of changing sizes on 8 emulated ranks, no host synchronisation in between: a lost or
    reordered flag would either trap (spin timeout -> CUDA error) or leave a wrong sum / an error flag."""
    from paddle_operator_b200.parallel.symm import AR_ONESHOT, AR_TWOSHOT

    comm = comms(8)
    comm.reset_allocations()
    n_max = 1 << 16
    a = comm.alloc(n_max * 2)
    b = comm.alloc(n_max * 2)
    for _ in range(500):
        comm.barrier()
    torch.manual_seed(3)
    for it in range(200):
        n = 8 * (1 + (it * 37) % 4096)
        for r in range(comm.nranks):
            comm.tensor(a, (n,), torch.bfloat16, rank=r).fill_(float(r + 1 + it % 3))
        algo = AR_TWOSHOT if it % 2 == 0 else AR_ONESHOT
        out = a if algo == AR_TWOSHOT else b
        comm.all_reduce_bf16(a, out, n, scale=1.0 / comm.nranks, algo=algo, nblocks=8)  # 64 co-resident blocks
        if it % 20 == 19:  # sum over r of (r + 1 + k) / 8 = 4.5 + k
            exp = 4.5 + it % 3
            for r in (0, comm.nranks - 1):
                got = comm.tensor(out, (n,), torch.bfloat16, rank=r).float()
                assert torch.all(got == exp), (it, r, float(got[0]), exp)
    torch.cuda.synchronize()
    assert all(e == 0 for e in comm.error_flags())


# ---------------------------------------------------------------------------
@pytest.mark.parametrize("gemm_1x1", [False, True])
def test_resnet_fused_matches_torch(dev, gemm_1x1, monkeypatch):
    """Few-step loss parity: fused model + engine vs stock torch modules + torch SGD.  With gemm_1x1 the
    64-channel 1x1 convolutions of the tiny model run on the tcgen05 kernel with fused BN statistics."""
    from paddle_operator_b200.models.resnet import resnet_tiny, to_training_dtype
    from paddle_operator_b200.ops.optim import softmax_cross_entropy
    from paddle_operator_b200.parallel.collective import CollectiveTrainer

    torch.manual_seed(0)
    from paddle_operator_b200.models.resnet import ResNet
    from paddle_operator_b200.ops import conv as conv_ops

    monkeypatch.setattr(conv_ops, "MIN_OUTPUT_BYTES", 0)  # small activations: take the GEMM path anyway
    # two blocks per stage, so that the second one has a plain skip connection (fused dgrad + add path)
    layers = (2, 2) if gemm_1x1 else (1, 1)
    ref = ResNet(layers, 16, fused=False, width=16).to(dev).float().train()
    fused = ResNet(layers, 16, fused=True, width=16, gemm_1x1=gemm_1x1).to(dev)
    with torch.no_grad():
        for pf, pr in zip(fused.parameters(), ref.parameters()):
            pf.copy_(pr)
    fused = to_training_dtype(fused, torch.bfloat16).train()
    x = torch.randn(16, 3, 64, 64, device=dev)
    y = torch.randint(0, 16, (16,), device=dev)
    xb = x.to(torch.bfloat16).contiguous(memory_format=torch.channels_last)

    tr = CollectiveTrainer(fused, lambda m, a, b: softmax_cross_entropy(m(a), b), (xb, y), lr=0.05, momentum=0.9,
                           weight_decay=1e-4, use_graph=False)
    opt = torch.optim.SGD(ref.parameters(), lr=0.05, momentum=0.9, weight_decay=1e-4)
    lf, lr_ = [], []
    for _ in range(6):
        lf.append(float(tr.step().item()))
        opt.zero_grad()
        loss = torch.nn.functional.cross_entropy(ref(xb.float()), y)
        loss.backward()
        opt.step()
        lr_.append(float(loss.item()))
    assert tr.kernels_per_step > 0
    assert lf[-1] < lf[0], lf  # it learns
    for a, b in zip(lf, lr_):
        assert abs(a - b) < 0.08 * max(1.0, abs(b)), (lf, lr_)


def test_graft_smoke():
    import __graft_entry__ as g

    g.smoke()
