"""Wide&Deep samples/s measurement used by bench.py (BASELINE.json config 3): PS mode with
GPU-sharded tables.

Arms (`measure(..., impl=...)`):
  product     fused lookup / push kernels over NVSwitch peer memory, tcgen05 tower, CUDA graph;
  nccl        stock torch + NCCL: row-sharded table, ids / rows / gradients exchanged with
              fixed-capacity `all_to_all_single`, sort-based duplicate merge and Adagrad on the
              touched rows, DDP for the dense part - no host synchronisation inside the step;
  nccl_graph  the same step captured in one CUDA graph (the dense gradients all-reduced inside it).
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist

from ..models.wide_deep import DENSE_FEATURES, SPARSE_SLOTS, WideDeepDense, bce_with_logits
from .bench_common import job_key, measure_job, release


class NcclShardedJob:
    """Baseline: the same sharding as the product (row `id` on shard id % world at local row
    id // world) on stock torch ops.  Every buffer has a fixed shape: each rank may send up to `cap`
    ids to each peer (5 % above the mean of a uniform draw; `overflowed` reports a violation), unused
    slots carry id -1 and resolve to a spare row, so nothing in the step waits for the host."""

    def __init__(self, batch, rows, dim, rank, world, dev, graph: bool = False, lr: float = 0.05, eps: float = 1e-6):
        self.batch, self.rank, self.world, self.dev, self.dim = batch, rank, world, dev, dim
        self.rps = (rows + world - 1) // world
        self.num_rows = self.rps * world
        n = batch * SPARSE_SLOTS
        self.n = n
        self.cap = n if world == 1 else int(n / world * 1.05) + 64
        self.graph_mode = graph
        g = torch.Generator(device=dev).manual_seed(rank)
        # row `rps` is the spare row that padding slots read and "update"
        self.deep = torch.randn(self.rps + 1, dim, device=dev, generator=g) * 0.01
        self.wide = torch.randn(self.rps + 1, device=dev, generator=g) * 0.01
        self.acc_d = torch.zeros_like(self.deep)
        self.acc_w = torch.zeros_like(self.wide)
        torch.manual_seed(0)
        dense = WideDeepDense(dim).to(dev)
        if dev.type == "cuda":
            dense = dense.to(torch.bfloat16)
        self.net = dense
        self.model = dense
        if world > 1 and not graph:
            self.model = torch.nn.parallel.DistributedDataParallel(
                dense, device_ids=[dev.index] if dev.type == "cuda" else None)
        self.opt = torch.optim.SGD(dense.parameters(), lr=0.05, momentum=0.9, fused=(dev.type == "cuda"))
        gen = torch.Generator().manual_seed(rank)
        self.ids = torch.randint(0, self.num_rows, (batch, SPARSE_SLOTS), generator=gen).to(dev)
        self.dense_x = torch.randn(batch, DENSE_FEATURES, generator=gen).to(dev)
        self.labels = (torch.rand(batch, generator=gen) > 0.75).float().to(dev)
        self.loss = torch.zeros((), device=dev)
        self.overflowed = torch.zeros((), dtype=torch.bool, device=dev)
        self._loss_host = torch.zeros(1).pin_memory() if dev.type == "cuda" else torch.zeros(1)
        self.lr, self.eps = lr, eps
        self._graph = None
        if graph:
            self._capture()

    # -- pull ---------------------------------------------------------------------------------
    def _route(self, flat_ids):
        """-> (order, slot, local_rows): position i of the owner-sorted ids sits in send slot slot[i];
        local_rows are the rows this rank serves (w * cap entries, spare row for unused slots)."""
        w, cap, n = self.world, self.cap, self.n
        owner = flat_ids % w
        order = torch.argsort(owner, stable=True)
        sorted_ids, sorted_owner = flat_ids[order], owner[order]
        counts = torch.zeros(w, dtype=torch.int64, device=self.dev).scatter_add_(0, owner, torch.ones_like(owner))
        starts = torch.cumsum(counts, 0) - counts
        pos = torch.arange(n, device=self.dev) - starts[sorted_owner]
        fits = pos < cap
        self.overflowed |= ~fits.all()
        slot = torch.where(fits, sorted_owner * cap + pos, torch.full_like(pos, w * cap))  # last slot = trash
        send = torch.full((w * cap + 1,), -1, dtype=torch.int64, device=self.dev).scatter_(0, slot, sorted_ids)
        recv = torch.empty(w * cap, dtype=torch.int64, device=self.dev)
        dist.all_to_all_single(recv, send[:w * cap])
        local_rows = torch.where(recv >= 0, recv // w, torch.full_like(recv, self.rps))
        return order, slot.clamp_(max=w * cap - 1), local_rows

    def _a2a(self, t):
        if self.world == 1:
            return t
        out = torch.empty_like(t)
        dist.all_to_all_single(out, t)
        return out

    # -- push: merge duplicate rows with a sort, then Adagrad on one entry per touched row -------
    def _sparse_adagrad(self, local_rows, g_deep, g_wide):
        m, D = local_rows.numel(), self.dim
        r_sorted, perm = torch.sort(local_rows)
        first = torch.ones(m, dtype=torch.bool, device=self.dev)
        first[1:] = r_sorted[1:] != r_sorted[:-1]
        seg = torch.cumsum(first, 0) - 1
        G = torch.zeros(m, D, device=self.dev).index_add_(0, seg, g_deep[perm])
        Gw = torch.zeros(m, device=self.dev).index_add_(0, seg, g_wide[perm])
        seg_row = torch.full((m,), self.rps, dtype=torch.int64, device=self.dev).scatter_(0, seg, r_sorted)
        a = self.acc_d[seg_row] + G * G
        self.acc_d[seg_row] = a
        self.deep.index_add_(0, seg_row, -self.lr * G * torch.rsqrt(a + self.eps))
        aw = self.acc_w[seg_row] + Gw * Gw
        self.acc_w[seg_row] = aw
        self.wide.index_add_(0, seg_row, -self.lr * Gw * torch.rsqrt(aw + self.eps))

    def _allreduce_dense_grads(self) -> None:
        grads = [p.grad for p in self.net.parameters()]
        flat = torch.cat([g.reshape(-1) for g in grads])
        dist.all_reduce(flat)
        flat.mul_(1.0 / self.world)
        views, off = [], 0
        for g in grads:
            views.append(flat[off:off + g.numel()].view(g.shape))
            off += g.numel()
        torch._foreach_copy_(grads, views)

    def _body(self) -> None:
        w, D, n = self.world, self.dim, self.n
        act_dtype = next(self.net.parameters()).dtype
        if w == 1:  # every row is local: a plain gather
            local_rows = self.ids.reshape(-1)
            deep, wide_rows = self.deep[local_rows], self.wide[local_rows]
        else:
            order, slot, local_rows = self._route(self.ids.reshape(-1))
            back_d = self._a2a(self.deep[local_rows])
            back_w = self._a2a(self.wide[local_rows])
            deep = torch.empty(n, D, device=self.dev).index_copy_(0, order, back_d[slot])
            wide_rows = torch.empty(n, device=self.dev).index_copy_(0, order, back_w[slot])
        deep = deep.view(self.batch, SPARSE_SLOTS * D).to(act_dtype).requires_grad_(True)
        wide = wide_rows.view(self.batch, SPARSE_SLOTS).sum(1).requires_grad_(True)
        loss = bce_with_logits(self.model(deep, wide, self.dense_x), self.labels)
        loss.backward()
        if self.graph_mode and w > 1:
            self._allreduce_dense_grads()
        self.opt.step()
        # gradients travel back to the owners in the slots the ids went out in
        gd = deep.grad.float().view(n, D) / w
        gw = wide.grad.repeat_interleave(SPARSE_SLOTS) / w
        if w > 1:
            send_d = torch.zeros(w * self.cap, D, device=self.dev).index_copy_(0, slot, gd[order])
            send_w = torch.zeros(w * self.cap, device=self.dev).index_copy_(0, slot, gw[order])
            gd, gw = self._a2a(send_d), self._a2a(send_w)
        self._sparse_adagrad(local_rows, gd, gw)
        self.loss.copy_(loss.detach().float())

    def _capture(self) -> None:
        s = torch.cuda.Stream(self.dev)
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(3):
                self.opt.zero_grad(set_to_none=True)
                self._body()
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize(self.dev)
        if self.world > 1:
            dist.barrier(device_ids=[self.dev.index])
            torch.cuda.synchronize(self.dev)
        self.opt.zero_grad(set_to_none=True)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, capture_error_mode="thread_local"):
            self._body()
        self._graph = g
        torch.cuda.synchronize(self.dev)

    def step_device(self):
        if self._graph is not None:
            self._graph.replay()
        else:
            self.opt.zero_grad(set_to_none=True)
            self._body()
        return self.loss

    def train_step(self, ids, dense, labels):
        self._h2d_bytes = sum(t.numel() * t.element_size() for t in (ids, dense, labels))
        self.ids.copy_(ids, non_blocking=True)
        self.dense_x.copy_(dense, non_blocking=True)
        self.labels.copy_(labels, non_blocking=True)
        loss = self.step_device()
        self._loss_host.copy_(loss.reshape(1), non_blocking=True)
        return self._loss_host

    def full_tables(self):
        """This rank's shard without the spare row (tests)."""
        return self.deep[:self.rps], self.wide[:self.rps]

    h2d_bytes_per_step = property(lambda s: getattr(s, "_h2d_bytes", s.batch * (SPARSE_SLOTS * 8 + DENSE_FEATURES * 4 + 4)))
    d2h_bytes_per_step = 4
    launches_per_step = 0


IMPL_NAMES = {
    "product": "this repository (all-to-all fused into TMA-fed lookup / sparse update, tcgen05 tower, CUDA graph)",
    "nccl": "torch + NCCL fixed-capacity all_to_all_single, sort-merge Adagrad, DDP dense part; eager, no host sync",
    "nccl_graph": "the same torch + NCCL step captured in one CUDA graph",
}


def measure(args, impl: str, rank: int, world: int, dev, ClockSampler) -> dict:
    batch = args.wd_batch or 32768
    rows, dim = 1 << 24, 16
    nvls = False
    if impl == "product":
        from ..parallel.symm import Communicator
        from .wide_deep import WideDeepJob

        comm = None
        if world > 1:
            comm = Communicator.from_env(WideDeepJob.arena_bytes_needed(rows, dim, batch, world), key=job_key("wd"))
        job = WideDeepJob(batch=batch, rows=rows, emb_dim=dim, comm=comm, device=dev, use_graph=not args.no_graph)
        job.prepare()
        nvls = bool(job.comm.has_multicast)
        hb = [job.synthetic_batch(s) for s in (1, 2, 3)]
        num_rows = job.emb.num_rows
    else:
        job = NcclShardedJob(batch, rows, dim, rank, world, dev, graph=(impl == "nccl_graph"))
        num_rows = job.num_rows
        g = torch.Generator().manual_seed(100 + rank)
        ids_dtype = torch.int32 if os.environ.get("PDJ_WD_IDS", "int32") == "int32" else torch.int64  # 16.7 M rows fit 32 bits
        hb = [(torch.randint(0, num_rows, (batch, SPARSE_SLOTS), generator=g).to(ids_dtype).pin_memory(),
               torch.randn(batch, DENSE_FEATURES, generator=g).pin_memory(),
               (torch.rand(batch, generator=g) > 0.75).float().pin_memory()) for _ in range(3)]
    steps, warmup = args.steps, max(args.warmup, 3)
    rec = measure_job(job, hb, steps, warmup, world, rank, dev, ClockSampler, batch * world)
    if impl != "product" and bool(job.overflowed.item()):
        rec["invalid"] = "a destination received more ids than the fixed all-to-all capacity"
    rec.update({
        "metric": "wide_deep_train_samples_per_sec",
        "unit": "samples/s",
        "impl": impl,
        "impl_name": IMPL_NAMES[impl],
        "dtype": "bf16 (MLP) / fp32 (tables)",
        "data": "synthetic (uniform random ids over the table - int32 on the host, widened on the device; random dense features "
                "and labels; random-init weights)",
        "config": {"model": "Wide&Deep: 26 sparse slots x dim 16 + 13 dense -> 400-400-400-1, Adagrad (sparse) + SGD momentum (dense)",
                   "table_rows": num_rows, "global_batch": batch * world, "per_gpu_batch": batch, "seq_len": None,
                   "parallelism": f"ps{world}+dp{world} (one table shard per GPU)",
                   "l2": "table (1 GiB+) and gathered rows exceed the 126 MB L2; ids differ per rank",
                   "cuda_graph": bool((impl == "product" and not args.no_graph) or impl == "nccl_graph"),
                   "nvls": nvls},
    })
    torch.cuda.synchronize(dev)
    if impl == "product":
        comm = job.comm
        del job
        comm.close()
    else:
        del job
    release(dev)
    return rec
