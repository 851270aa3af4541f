"""Wide&Deep PS-mode workload: the public training API a trainer rank uses.

    job = WideDeepJob.from_env(batch=32768)
    job.prepare()
    for ids, dense, labels in loader:        # pinned host tensors
        loss = job.train_step(ids, dense, labels)

Sparse parameters live in a `ShardedEmbedding` (one shard per GPU, pulled and
pushed by the fused kernels of csrc/embedding.cu); the dense MLP is
data-parallel through `CollectiveTrainer` (flat buffers, fused all-reduce,
fused SGD).  The whole step is one CUDA graph.

Reference parity: the TRAINER program of deploy/examples/wide_and_deep.yaml
(there: Paddle's demo image); rank / world come from PADDLE_TRAINER_ID /
PADDLE_TRAINERS_NUM (controllers/paddlejob_helper.go:264,319-322).
"""
from __future__ import annotations

import os
from typing import Optional

import torch

from ..models.wide_deep import DENSE_FEATURES, SPARSE_SLOTS, WideDeepDense
from ..parallel.collective import CollectiveTrainer
from ..parallel.ps import ADAGRAD, ShardedEmbedding
from ..parallel.symm import Communicator


class _Net(torch.nn.Module):
    """Dense module + a handle on the sharded embedding (which is not a Parameter)."""

    def __init__(self, dense: WideDeepDense, emb: ShardedEmbedding):
        super().__init__()
        self.dense = dense
        self._emb = [emb]

    def forward(self, ids, dense_x):
        deep, wide = self._emb[0](ids)
        return self.dense(deep, wide, dense_x)

    def loss(self, ids, dense_x, labels):
        deep, wide = self._emb[0](ids)
        return self.dense.loss(deep, wide, dense_x, labels)


def _loss_fn(model, ids, dense_x, labels):
    return model.loss(ids, dense_x, labels)


class WideDeepJob:
    def __init__(self, batch: int = 32768, rows: int = 1 << 24, emb_dim: int = 16, comm: Optional[Communicator] = None,
                 device: Optional[torch.device] = None, lr: float = 0.05, sparse_lr: float = 0.05,
                 optimizer: int = ADAGRAD, async_sgd: bool = False, use_graph: bool = True, seed: int = 0,
                 arena_bytes: Optional[int] = None, emb_blocks: int = 0, num_shards: Optional[int] = None):
        self.device = device or torch.device("cuda", torch.cuda.current_device())
        self.batch, self.rows, self.emb_dim = batch, rows, emb_dim
        world = comm.nranks if comm is not None else 1
        if comm is None:
            # single rank: a private arena, same kernels (every row is local)
            need = arena_bytes or self.arena_bytes_needed(rows, emb_dim, batch, 1)
            comm = Communicator.create(0, 1, self.device.index or 0, need, key=f"wd-{os.getpid()}")
            self._own_comm = True
        else:
            self._own_comm = False
        self.comm = comm
        self.rank, self.world = comm.rank, world
        self.emb = ShardedEmbedding(comm, rows, emb_dim, batch, SPARSE_SLOTS, has_wide=True, optimizer=optimizer,
                                    lr=sparse_lr, async_sgd=async_sgd, seed=seed, nblocks=emb_blocks or 148,
                                    num_shards=num_shards)
        self.emb.grad_scale = 1.0 / world
        torch.manual_seed(seed)
        dense = WideDeepDense(emb_dim, fused=True).to(self.device).to(torch.bfloat16)
        self.model = _Net(dense, self.emb).train()
        # the step's ids live where the lookup kernel reads them (no per-step copy into the arena)
        self.ids = None
        # the lookup kernel writes the tower's padded input [26 x 16 embeddings | 13 dense | 3 zeros] in place: the
        # dense features live in the arena where it reads them (no concatenation / conversion / fill kernels)
        self.emb.set_output_layout(dense.in_dim, DENSE_FEATURES)
        self.ids = self.emb.ids_buffer(batch)
        self.ids.zero_()
        self.dense_x = self.emb.dense_buffer(batch)
        self.dense_x.zero_()
        self.labels = torch.zeros((batch,), dtype=torch.float32, device=self.device)
        self.trainer = CollectiveTrainer(self.model, _loss_fn, (self.ids, self.dense_x, self.labels),
                                         comm=comm if world > 1 else None, lr=lr, momentum=0.9, weight_decay=0.0,
                                         use_graph=use_graph, bucket_bytes=4 << 20, comm_blocks=8,
                                         # embedding kernels + 2 tcgen05 GEMMs (forward, dX) per hidden layer
                                         # embedding kernels + per hidden layer: forward GEMM, dX GEMM, bias colsum
                                         extra_kernels_per_step=(3 if async_sgd else 4) + 3 * len(dense.layers) + 2)  # + fused head / loss, its dz
        self._attach_bias_masters()
        self.trainer.on_rehome.append(self._attach_bias_masters)
        self._copy_stream = torch.cuda.Stream(self.device)
        self._stage = [(torch.empty_like(self.ids), torch.empty_like(self.dense_x), torch.empty_like(self.labels))
                       for _ in range(2)]
        self._copied = [torch.cuda.Event() for _ in range(2)]
        self._consumed = [torch.cuda.Event() for _ in range(2)]
        for e in self._consumed:
            e.record()
        self._slot = 0
        self._loss_host = torch.zeros(64, dtype=torch.float32).pin_memory()
        self._loss_i = 0

    def _attach_bias_masters(self) -> None:
        """The tcgen05 epilogue adds fp32 biases: point it at the trainer's master copies."""
        from ..ops.mlp import FusedMLP

        for m in self.model.modules():
            if isinstance(m, FusedMLP):
                m.bias_f32 = [self.trainer.flat.master_view(l.bias) for l in m.layers]

    @staticmethod
    def arena_bytes_needed(rows: int, dim: int, batch: int, world: int) -> int:
        r = (rows + world - 1) // world
        head = (2 * dim + 12 + 63) // 64 * 64
        table = r * ((head + 12 * dim + 127) // 128 * 128) + r  # one record per row + flag bytes
        io = batch * SPARSE_SLOTS * (8 + dim * 2 * 2 + 4 * world) + batch * 8
        return int(table + io + (96 << 20))

    @classmethod
    def from_env(cls, batch: int = 32768, rows: int = 1 << 24, emb_dim: int = 16, **kw) -> "WideDeepJob":
        env = os.environ
        world = int(env.get("WORLD_SIZE", env.get("PADDLE_TRAINERS_NUM", "1")))
        comm = None
        # P table shards = the job's pserver count (spec.ps.replicas reaches the ranks as PADDLE_PSERVERS_IP_PORT_LIST,
        # controllers/paddlejob_helper.go:258-260); shard j lives on GPU j mod world
        ps = [e for e in env.get("PADDLE_PSERVERS_IP_PORT_LIST", "").split(",") if e.strip()]
        if ps and "num_shards" not in kw:
            kw["num_shards"] = len(ps)
        if world > 1:
            slots = -(-(kw.get("num_shards") or world) // world)  # shard slots per GPU
            comm = Communicator.from_env(cls.arena_bytes_needed(rows, emb_dim, batch, world) * max(1, slots * world // (kw.get("num_shards") or world)))
        return cls(batch=batch, rows=rows, emb_dim=emb_dim, comm=comm, **kw)

    def synthetic_batch(self, seed: int = 0, device=None):
        g = torch.Generator(device="cpu").manual_seed(seed * 7919 + self.rank)
        ids = torch.randint(0, self.emb.num_rows, (self.batch, SPARSE_SLOTS), generator=g, dtype=torch.int64)
        dense = torch.randn((self.batch, DENSE_FEATURES), generator=g)
        labels = (torch.rand((self.batch,), generator=g) > 0.75).float()
        if device is not None:
            return ids.to(device), dense.to(device), labels.to(device)
        if self.emb.num_rows < (1 << 31) and os.environ.get("PDJ_WD_IDS", "int32") == "int32":
            ids = ids.to(torch.int32)  # what a data loader hands over for a table this size: half the PCIe bytes
        return ids.pin_memory(), dense.pin_memory(), labels.pin_memory()

    def prepare(self) -> None:
        ids, dense, labels = self.synthetic_batch(0, self.device)
        self.ids.copy_(ids)
        self.dense_x.copy_(dense)
        self.labels.copy_(labels)
        self.trainer.capture()

    def step_device(self) -> torch.Tensor:
        return self.trainer.step()

    def train_step(self, ids: torch.Tensor, dense: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        """One step on a pinned host batch: ids [B, 26] int64 or int32 (a table of fewer than 2^31 rows needs no
        more; half the bytes over PCIe, widened on the device), dense [B, 13] fp32, labels [B] fp32.  The copy of
        batch k + 1 runs on a side stream under step k; returns a pinned one-element tensor holding the loss."""
        s = self._slot
        self._slot ^= 1
        main = torch.cuda.current_stream()
        if self._stage[s][0].dtype != ids.dtype:  # first int32 batch: staging buffers of the host dtype
            torch.cuda.synchronize(self.device)
            self._stage = [(torch.empty(self.ids.shape, dtype=ids.dtype, device=self.device), st[1], st[2])
                           for st in self._stage]
        st = self._stage[s]
        with torch.cuda.stream(self._copy_stream):
            self._copy_stream.wait_event(self._consumed[s])
            st[0].copy_(ids, non_blocking=True)
            st[1].copy_(dense, non_blocking=True)
            st[2].copy_(labels, non_blocking=True)
            self._copied[s].record(self._copy_stream)
        main.wait_event(self._copied[s])
        self.ids.copy_(st[0])
        self.dense_x.copy_(st[1])
        self.labels.copy_(st[2])
        self._consumed[s].record(main)
        self._h2d_bytes = sum(t.numel() * t.element_size() for t in (ids, dense, labels))
        loss = self.step_device()
        i = self._loss_i
        self._loss_i = (i + 1) % self._loss_host.numel()
        out = self._loss_host[i:i + 1]
        out.copy_(loss.reshape(1), non_blocking=True)
        return out

    @property
    def h2d_bytes_per_step(self) -> int:
        # bytes of the last batch handed to train_step (int64 ids until one has been seen)
        return getattr(self, "_h2d_bytes", self.batch * (SPARSE_SLOTS * 8 + DENSE_FEATURES * 4 + 4))

    d2h_bytes_per_step = 4

    @property
    def launches_per_step(self) -> int:
        # lookup + push + apply + count reset, plus the trainer's all-reduce / fused SGD launches
        return self.trainer.kernels_per_step

    # -- checkpoints in the reference's layout: model.pt (dense part) + one ps-<j>.pt per table shard + step.json ---
    def save_checkpoint(self, root: str, step: int) -> None:
        """Every rank writes the shards it holds (named by pserver index j, not by trainer rank), rank 0 the dense
        part; step.json marks the checkpoint complete once every rank has passed the barrier."""
        from ..utils import checkpoint as ckpt

        torch.cuda.synchronize(self.device)
        if self.world > 1:
            self.trainer.sync_state()
            torch.cuda.synchronize(self.device)
        for j, sd in self.emb.shard_state_dicts().items():
            ckpt.save_ps_shard(root, j, sd)
        if self.rank == 0:
            ckpt.save_dp(root, step, self.world, {"trainer": self.trainer.state_dict(), "step": step,
                                                  "nshards": self.emb.num_shards})
        if self.world > 1:
            self.comm.barrier()
            torch.cuda.synchronize(self.device)
        if self.rank == 0:
            ckpt.commit(root, step, self.world)

    def load_checkpoint(self, root: str) -> Optional[int]:
        """Resumes from `root` when it holds a complete checkpoint with the same shard count; returns its step."""
        from ..utils import checkpoint as ckpt

        meta = ckpt.latest(root)
        dp = ckpt.load_dp(root, map_location=self.device) if meta else None
        if dp is None or int(dp.get("nshards", -1)) != self.emb.num_shards:
            return None
        for j in self.emb.shards_of(self.rank):
            sd = ckpt.load_ps_shard(root, j, map_location=self.device)
            if sd is None:
                return None
            self.emb.load_shard_state_dict(sd)
        self.trainer.load_state_dict(dp["trainer"])
        return int(dp["step"])

    def close(self) -> None:
        torch.cuda.synchronize(self.device)
        if self._own_comm:
            self.comm.close()
