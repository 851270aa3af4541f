"""Wide&Deep samples/s measurement used by `bench.py --workload wide_deep`
(BASELINE.json config 3): PS mode with GPU-sharded tables, product vs a stock
torch + NCCL all-to-all implementation of the same sharded embedding."""
from __future__ import annotations

import json
import time

import torch
import torch.distributed as dist

from ..models.wide_deep import DENSE_FEATURES, SPARSE_SLOTS, WideDeepDense, bce_with_logits
from .bench_resnet import _max_over_ranks, _sync_all, _time_region  # noqa: F401


class _NcclShardedJob:
    """Baseline: row-sharded table, ids and rows exchanged with NCCL all_to_all_single,
    local F.embedding, torch Adagrad on the touched rows, DDP for the dense part."""

    def __init__(self, batch, rows, dim, rank, world, dev):
        self.batch, self.rank, self.world, self.dev, self.dim = batch, rank, world, dev, dim
        self.rps = (rows + world - 1) // world
        self.num_rows = self.rps * world
        g = torch.Generator(device=dev).manual_seed(rank)
        self.deep = torch.randn(self.rps, dim, device=dev, generator=g) * 0.01
        self.wide = torch.randn(self.rps, device=dev, generator=g) * 0.01
        self.acc_d = torch.zeros_like(self.deep)
        self.acc_w = torch.zeros_like(self.wide)
        torch.manual_seed(0)
        dense = WideDeepDense(dim).to(dev).to(torch.bfloat16)
        self.model = dense
        if world > 1:
            self.model = torch.nn.parallel.DistributedDataParallel(dense, device_ids=[dev.index])
        self.opt = torch.optim.SGD(dense.parameters(), lr=0.05, momentum=0.9, fused=True)
        gen = torch.Generator().manual_seed(rank)
        self.ids = torch.randint(0, self.num_rows, (batch, SPARSE_SLOTS), generator=gen).to(dev)
        self.dense_x = torch.randn(batch, DENSE_FEATURES, generator=gen).to(dev)
        self.labels = (torch.rand(batch, generator=gen) > 0.75).float().to(dev)
        self.loss = torch.zeros((), device=dev)
        self._loss_host = torch.zeros(1).pin_memory()
        self.lr, self.eps = 0.05, 1e-6

    def _exchange(self, flat_ids):
        w = self.world
        owner = flat_ids % w
        order = torch.argsort(owner, stable=True)
        sorted_ids = flat_ids[order]
        counts = torch.bincount(owner, minlength=w)
        if w == 1:
            return order, sorted_ids, counts.tolist(), counts.tolist()
        rcounts = torch.empty_like(counts)
        dist.all_to_all_single(rcounts, counts)
        sc, rc = counts.tolist(), rcounts.tolist()
        recv = torch.empty(sum(rc), dtype=flat_ids.dtype, device=self.dev)
        dist.all_to_all_single(recv, sorted_ids, rc, sc)
        return order, recv, sc, rc

    def step_device(self):
        w, D = self.world, self.dim
        flat = self.ids.reshape(-1)
        order, req, sc, rc = self._exchange(flat)
        local_rows = req // w
        rows_d = self.deep[local_rows]
        rows_w = self.wide[local_rows]
        if w > 1:
            back_d = torch.empty(sum(sc), D, device=self.dev)
            back_w = torch.empty(sum(sc), device=self.dev)
            dist.all_to_all_single(back_d, rows_d, sc, rc)
            dist.all_to_all_single(back_w, rows_w, sc, rc)
        else:
            back_d, back_w = rows_d, rows_w
        inv = torch.empty_like(order)
        inv[order] = torch.arange(order.numel(), device=self.dev)
        deep = back_d[inv].view(self.batch, SPARSE_SLOTS * D).to(torch.bfloat16).requires_grad_(True)
        wide_rows = back_w[inv].view(self.batch, SPARSE_SLOTS)
        wide = wide_rows.sum(1).requires_grad_(True)
        self.opt.zero_grad(set_to_none=True)
        loss = bce_with_logits(self.model(deep, wide, self.dense_x), self.labels)
        loss.backward()
        self.opt.step()
        # push: gradients travel back to the owners
        gd = (deep.grad.float().view(-1, D) / w)[order]
        gw = (wide.grad.repeat_interleave(SPARSE_SLOTS) / w)[order]
        if w > 1:
            rgd = torch.empty(sum(rc), D, device=self.dev)
            rgw = torch.empty(sum(rc), device=self.dev)
            dist.all_to_all_single(rgd, gd, rc, sc)
            dist.all_to_all_single(rgw, gw, rc, sc)
        else:
            rgd, rgw = gd, gw
        uniq, inverse = torch.unique(local_rows, return_inverse=True)
        G = torch.zeros(uniq.numel(), D, device=self.dev).index_add_(0, inverse, rgd)
        Gw = torch.zeros(uniq.numel(), device=self.dev).index_add_(0, inverse, rgw)
        a = self.acc_d[uniq] + G * G
        self.acc_d[uniq] = a
        self.deep[uniq] -= self.lr * G * torch.rsqrt(a + self.eps)
        aw = self.acc_w[uniq] + Gw * Gw
        self.acc_w[uniq] = aw
        self.wide[uniq] -= self.lr * Gw * torch.rsqrt(aw + self.eps)
        self.loss = loss.detach()
        return self.loss

    def train_step(self, ids, dense, labels):
        self.ids = ids.to(self.dev, non_blocking=True)
        self.dense_x = dense.to(self.dev, non_blocking=True)
        self.labels = labels.to(self.dev, non_blocking=True)
        loss = self.step_device()
        self._loss_host.copy_(loss.reshape(1), non_blocking=True)
        return self._loss_host

    h2d_bytes_per_step = property(lambda s: s.batch * (SPARSE_SLOTS * 8 + DENSE_FEATURES * 4 + 4))
    d2h_bytes_per_step = 4
    launches_per_step = 0


def run(args, rank: int, world: int, dev, ClockSampler) -> int:
    batch = args.batch or 32768
    rows, dim = 1 << 24, 16
    if args.impl == "product":
        from .wide_deep import WideDeepJob

        job = WideDeepJob.from_env(batch=batch, rows=rows, emb_dim=dim, use_graph=not args.no_graph)
        job.prepare()
        nvls = bool(job.comm.has_multicast)
    else:
        job = _NcclShardedJob(batch, rows, dim, rank, world, dev)
        nvls = False
    steps, warmup = args.steps, max(args.warmup, 3)
    for _ in range(warmup):
        job.step_device()
    sampler = ClockSampler() if rank == 0 else None
    if sampler:
        sampler.start()
        time.sleep(0.05)
    ms = _time_region(job.step_device, steps, world, dev)
    clocks = sampler.stop() if sampler else None
    loss_dev = float(job.trainer.loss.item()) if args.impl == "product" else float(job.loss.item())

    gen_seed = [1]
    if args.impl == "product":
        hb = [job.synthetic_batch(s) for s in (1, 2, 3)]
    else:
        g = torch.Generator().manual_seed(100 + rank)
        hb = [(torch.randint(0, job.num_rows, (batch, SPARSE_SLOTS), generator=g).pin_memory(),
               torch.randn(batch, DENSE_FEATURES, generator=g).pin_memory(),
               (torch.rand(batch, generator=g) > 0.75).float().pin_memory()) for _ in range(3)]
    k, last = [0], [None]

    def e2e_step():
        b = hb[k[0] % len(hb)]
        k[0] += 1
        last[0] = job.train_step(*b)

    for _ in range(3):
        e2e_step()
    ms_e2e = _time_region(e2e_step, steps, world, dev)
    loss_e2e = float(last[0].item())
    samples = batch * world * steps
    out = {
        "metric": "wide_deep_train_samples_per_sec",
        "value": round(samples / (ms / 1e3), 1),
        "unit": "samples/s",
        "n_gpus": world, "steps": steps, "warmup": warmup, "ms_per_step": round(ms / steps, 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16 (MLP) / fp32 (tables)",
        "data": "synthetic (uniform random ids over the table, random dense features and labels; random-init weights)",
        "impl": args.impl,
        "config": {"model": "Wide&Deep: 26 sparse slots x dim 16 + 13 dense -> 400-400-400-1, Adagrad (sparse) + SGD momentum (dense)",
                   "table_rows": rows, "global_batch": batch * world, "per_gpu_batch": batch, "seq_len": None,
                   "parallelism": f"ps{world}+dp{world} (one table shard per GPU)",
                   "l2": "table (1 GiB+) and gathered rows exceed the 126 MB L2; ids differ per rank",
                   "cuda_graph": bool(args.impl == "product" and not args.no_graph), "nvls": nvls},
        "clocks": clocks,
        "e2e": {"value": round(samples / (ms_e2e / 1e3), 1), "unit": "samples/s", "ms_per_step": round(ms_e2e / steps, 3),
                "h2d_bytes_per_step": job.h2d_bytes_per_step, "d2h_bytes_per_step": job.d2h_bytes_per_step},
        "gpu_launches": job.launches_per_step * steps,
        "loss": {"device": loss_dev, "e2e": loss_e2e},
    }
    if rank == 0:
        print(json.dumps(out), flush=True)
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier(device_ids=[dev.index])
        dist.destroy_process_group()
    return 0
