"""K2/K3 numerics on one GPU through the multi-rank emulation: fused lookup and
fused sparse SGD/Adagrad update vs a plain PyTorch fp32 restatement."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _mk(nranks, rows=4096, dim=16, batch=96, slots=26, **kw):
    from paddle_operator_b200.parallel.ps import ShardedEmbedding
    from paddle_operator_b200.parallel.symm import Communicator

    comm = Communicator.emulated(nranks, arena_bytes=32 << 20)
    emb = ShardedEmbedding(comm, rows, dim, batch, slots, nblocks=4, **kw)
    return comm, emb


@pytest.mark.parametrize("nranks", [1, 2, 8])
@pytest.mark.parametrize("dim", [4, 8, 16, 64])
def test_lookup_matches_reference(nranks, dim):
    from paddle_operator_b200.parallel.ps import embedding_reference

    comm, emb = _mk(nranks, dim=dim, slots=5 if dim == 64 else 26)
    try:
        deep_t, wide_t = emb.full_tables()
        torch.manual_seed(1)
        ids = [torch.randint(0, emb.num_rows, (96, emb.slots), device="cuda") for _ in range(nranks)]
        for r in range(nranks):
            emb.ids_buffer(96, r).copy_(ids[r])
        emb.lookup_raw(96)
        torch.cuda.synchronize()
        for r in range(nranks):
            d_ref, w_ref = embedding_reference(deep_t, wide_t, ids[r])
            got = emb.out_deep(96, r).float().view(96, emb.slots, dim)
            torch.testing.assert_close(got, d_ref.to(torch.bfloat16).float(), atol=0, rtol=0)
            torch.testing.assert_close(emb.out_wide(96, r), w_ref, atol=1e-5, rtol=1e-5)
    finally:
        comm.close()


@pytest.mark.parametrize("nranks", [1, 2, 4])
@pytest.mark.parametrize("opt", ["sgd", "adagrad", "async"])
def test_push_update_matches_reference(nranks, opt):
    from paddle_operator_b200.parallel.ps import ADAGRAD, SGD

    kw = {"sgd": dict(optimizer=SGD), "adagrad": dict(optimizer=ADAGRAD), "async": dict(optimizer=SGD, async_sgd=True)}[opt]
    comm, emb = _mk(nranks, rows=512, lr=0.1, eps=1e-6, **kw)  # few rows -> many repeated ids, within and across ranks
    try:
        B, S, D = 96, emb.slots, emb.dim
        deep_ref, wide_ref = emb.full_tables()
        acc_d, acc_w = torch.zeros_like(deep_ref), torch.zeros_like(wide_ref)
        for step in range(3):
            torch.manual_seed(10 + step)
            ids = [torch.randint(0, emb.num_rows, (B, S), device="cuda") for _ in range(nranks)]
            gd = [(torch.randn(B, S * D, device="cuda") * 0.1).to(torch.bfloat16) for _ in range(nranks)]
            gw = [torch.randn(B, device="cuda") * 0.1 for _ in range(nranks)]
            for r in range(nranks):
                emb.ids_buffer(B, r).copy_(ids[r])
                emb.grad_deep_in(B, r).copy_(gd[r])
                emb.grad_wide_in(B, r).copy_(gw[r])
            emb.push_raw(B)
            torch.cuda.synchronize()
            # reference: sum the gradients of repeated ids over all trainers, then one optimizer step per row
            G = torch.zeros_like(deep_ref)
            Gw = torch.zeros_like(wide_ref)
            for r in range(nranks):
                G.index_add_(0, ids[r].reshape(-1), gd[r].float().view(B * S, D))
                Gw.index_add_(0, ids[r].reshape(-1), gw[r].repeat_interleave(S))
            if opt == "adagrad":
                acc_d += G * G
                acc_w += Gw * Gw
                touched = torch.zeros(emb.num_rows, dtype=torch.bool, device="cuda")
                for r in range(nranks):
                    touched[ids[r].reshape(-1)] = True
                deep_ref = torch.where(touched[:, None], deep_ref - 0.1 * G * torch.rsqrt(acc_d + 1e-6), deep_ref)
                wide_ref = torch.where(touched, wide_ref - 0.1 * Gw * torch.rsqrt(acc_w + 1e-6), wide_ref)
            else:
                deep_ref = deep_ref - 0.1 * G
                wide_ref = wide_ref - 0.1 * Gw
            deep_t, wide_t = emb.full_tables()
            torch.testing.assert_close(deep_t, deep_ref, atol=2e-5, rtol=1e-4)
            torch.testing.assert_close(wide_t, wide_ref, atol=2e-5, rtol=1e-4)
            if opt != "async":
                uniq = sum(int(torch.unique(torch.cat([i.reshape(-1) for i in ids]))[lambda u: u % nranks == r].numel())
                           for r in range(nranks)) if False else int(torch.unique(torch.cat([i.reshape(-1) for i in ids])).numel())
                assert sum(emb.touched_last_step(r) for r in range(nranks)) == uniq
                # staging rows and flags are clean again
                for r in range(nranks):
                    assert all(float(g.abs().max()) == 0.0 for g in emb.staging(r))
                    assert int(comm.tensor(emb.t.flag_off, (emb._flag_bytes,), torch.uint8, r).sum()) == 0
    finally:
        comm.close()


def test_autograd_lookup_and_fused_update_single_rank():
    """End-to-end through autograd on one real rank: the table moves against the gradient."""
    from paddle_operator_b200.models.wide_deep import WideDeepDense, bce_with_logits
    from paddle_operator_b200.parallel.ps import SGD, ShardedEmbedding
    from paddle_operator_b200.parallel.symm import Communicator

    comm = Communicator.emulated(1, arena_bytes=32 << 20)
    try:
        emb = ShardedEmbedding(comm, 2048, 16, 64, 26, optimizer=SGD, lr=0.5, nblocks=4)
        model = WideDeepDense(16).cuda().to(torch.bfloat16)
        opt = torch.optim.SGD(model.parameters(), lr=0.05)
        torch.manual_seed(0)
        ids = torch.randint(0, 2048, (64, 26), device="cuda")
        dense = torch.randn(64, 13, device="cuda")
        labels = (torch.rand(64, device="cuda") > 0.5).float()
        losses = []
        for _ in range(30):
            deep, wide = emb(ids)
            loss = bce_with_logits(model(deep, wide, dense), labels)
            opt.zero_grad()
            loss.backward()
            opt.step()
            losses.append(float(loss))
        assert losses[-1] < 0.5 * losses[0], losses
    finally:
        comm.close()


@pytest.mark.parametrize("nranks", [1, 2])
@pytest.mark.parametrize("how", ["batched", "tma", "loop"])
def test_lookup_writes_padded_tower_input_in_place(nranks, how, monkeypatch):
    """set_output_layout: rows of 432 = 26 x 16 embedding values | 13 dense features (fp32 -> bf16) | 3 zeros."""
    from paddle_operator_b200.parallel.ps import embedding_reference

    monkeypatch.setenv("PDJ_EMB_LOOKUP", how)  # the three lookup kernels (csrc/embedding.cu); batched is the default
    monkeypatch.setenv("PDJ_EMB_PUSH", "loop" if how == "loop" else "batched")
    comm, emb = _mk(nranks, dim=16, slots=26)
    try:
        emb.set_output_layout(432, 13)
        deep_t, wide_t = emb.full_tables()
        torch.manual_seed(3)
        ids = [torch.randint(0, emb.num_rows, (96, 26), device="cuda") for _ in range(nranks)]
        dense = [torch.randn(96, 13, device="cuda") for _ in range(nranks)]
        for r in range(nranks):
            emb.ids_buffer(96, r).copy_(ids[r])
            emb.dense_buffer(96, r).copy_(dense[r])
            emb.out_deep(96, r).fill_(7.0)  # stale contents must be overwritten, pad included
        emb.lookup_raw(96)
        torch.cuda.synchronize()
        for r in range(nranks):
            d_ref, w_ref = embedding_reference(deep_t, wide_t, ids[r])
            got = emb.out_deep(96, r).float()
            assert got.shape == (96, 432)
            torch.testing.assert_close(got[:, :416].view(96, 26, 16), d_ref.to(torch.bfloat16).float(), atol=0, rtol=0)
            torch.testing.assert_close(got[:, 416:429], dense[r].to(torch.bfloat16).float(), atol=0, rtol=0)
            assert float(got[:, 429:].abs().max()) == 0.0
            torch.testing.assert_close(emb.out_wide(96, r), w_ref, atol=1e-5, rtol=1e-5)
        # push with a pitched gradient (the gradient of that padded input) equals push of its first 416 columns
        g = [(torch.randn(96, 432, device="cuda") * 0.1).to(torch.bfloat16) for _ in range(nranks)]
        stage = emb.comm.alloc(96 * 432 * 2)
        for r in range(nranks):
            emb.comm.tensor(stage, (96, 432), torch.bfloat16, r).copy_(g[r])
            emb.grad_wide_in(96, r).zero_()
        before = emb.full_tables()[0].clone()
        emb.push_raw(96, 1.0, gdeep_off=stage, g_pitch=432)
        torch.cuda.synchronize()
        after = emb.full_tables()[0]
        ref = before.clone()
        G = torch.zeros_like(ref)
        for r in range(nranks):
            G.index_add_(0, ids[r].reshape(-1), g[r][:, :416].float().reshape(-1, 16))
        touched = G.abs().sum(1) > 0
        assert bool((after[~touched] == before[~touched]).all()) and bool((after[touched] != before[touched]).any())
    finally:
        comm.close()


@pytest.mark.parametrize("nranks,shards", [(2, 3), (4, 2), (2, 8), (1, 2)])
def test_shard_count_follows_the_pserver_count(nranks, shards):
    """P = spec.ps.replicas shards over W GPUs, P != W: shard j = id % P on GPU j % W.  Lookup and a sync SGD push
    agree with the unsharded reference, and the per-shard checkpoints are named by shard, not by trainer."""
    from paddle_operator_b200.parallel.ps import SGD, embedding_reference

    comm, emb = _mk(nranks, rows=1000, num_shards=shards, optimizer=SGD, lr=0.5)
    try:
        assert emb.num_shards == shards and emb.num_rows % shards == 0
        held = [emb.shards_of(r) for r in range(nranks)]
        assert sorted(j for h in held for j in h) == list(range(shards))
        deep_t, wide_t = emb.full_tables()
        torch.manual_seed(2)
        B, S, D = 96, emb.slots, emb.dim
        ids = [torch.randint(0, emb.num_rows, (B, S), device="cuda") for _ in range(nranks)]
        for r in range(nranks):
            emb.ids_buffer(B, r).copy_(ids[r])
        emb.lookup_raw(B)
        torch.cuda.synchronize()
        for r in range(nranks):
            d_ref, w_ref = embedding_reference(deep_t, wide_t, ids[r])
            torch.testing.assert_close(emb.out_deep(B, r).float().view(B, S, D), d_ref.to(torch.bfloat16).float(), atol=0, rtol=0)
            torch.testing.assert_close(emb.out_wide(B, r), w_ref, atol=1e-5, rtol=1e-5)
        gd = [(torch.randn(B, S * D, device="cuda") * 0.1).to(torch.bfloat16) for _ in range(nranks)]
        for r in range(nranks):
            emb.grad_deep_in(B, r).copy_(gd[r])
            emb.grad_wide_in(B, r).zero_()
        emb.push_raw(B, 1.0)
        torch.cuda.synchronize()
        G = torch.zeros_like(deep_t)
        for r in range(nranks):
            G.index_add_(0, ids[r].reshape(-1), gd[r].float().reshape(-1, D))
        torch.testing.assert_close(emb.full_tables()[0], deep_t - 0.5 * G, atol=1e-4, rtol=1e-4)
        sds = emb.shard_state_dicts()  # rank 0's shards
        assert sorted(sds) == held[0] and all(sd["deep"].shape == (emb.shard_rows, D) for sd in sds.values())
    finally:
        comm.close()


@pytest.mark.parametrize("nranks", [1, 2])
@pytest.mark.parametrize("dim", [8, 16, 32, 64])
def test_lookup_reads_the_bf16_view_kept_by_apply(nranks, dim):
    """Record layout with the bf16 read view: after push + apply a lookup returns bf16(updated fp32 weights), and
    refresh_shadow() re-derives the view after a direct write to shard()."""
    from paddle_operator_b200.parallel.ps import ADAGRAD, embedding_reference

    comm, emb = _mk(nranks, rows=777, dim=dim, slots=7, optimizer=ADAGRAD, lr=0.3, eps=1e-6)
    try:
        assert emb.has_shadow and emb.t.shadow_off != 0
        B, S, D = 96, emb.slots, dim
        torch.manual_seed(dim)
        for step in range(2):
            ids = [torch.randint(0, emb.num_rows, (B, S), device="cuda") for _ in range(nranks)]
            for r in range(nranks):
                emb.ids_buffer(B, r).copy_(ids[r])
                emb.grad_deep_in(B, r).copy_((torch.randn(B, S * D, device="cuda") * 0.5).to(torch.bfloat16))
                emb.grad_wide_in(B, r).copy_(torch.randn(B, device="cuda"))
            emb.push_raw(B)
            emb.lookup_raw(B)
            torch.cuda.synchronize()
            deep_t, wide_t = emb.full_tables()
            for r in range(nranks):
                d_ref, w_ref = embedding_reference(deep_t, wide_t, ids[r])
                got = emb.out_deep(B, r).float().view(B, S, D)
                torch.testing.assert_close(got, d_ref.to(torch.bfloat16).float(), atol=0, rtol=0)
                torch.testing.assert_close(emb.out_wide(B, r), w_ref, atol=1e-5, rtol=1e-5)
        for r in range(nranks):
            emb.shard(r).mul_(2.0)
            emb.refresh_shadow(r)
        emb.lookup_raw(B)
        torch.cuda.synchronize()
        deep_t, wide_t = emb.full_tables()
        d_ref, _ = embedding_reference(deep_t, wide_t, ids[0])
        torch.testing.assert_close(emb.out_deep(B, 0).float().view(B, S, D), d_ref.to(torch.bfloat16).float(), atol=0, rtol=0)
    finally:
        comm.close()
