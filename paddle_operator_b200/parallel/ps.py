"""PS mode: the parameter-server role as GPU-sharded embedding tables.

`ShardedEmbedding` owns one shard of a row-sharded table (plus staging rows,
Adagrad accumulators and the touched-row list) inside the symmetric arena of a
`Communicator`, and exposes

    deep, wide = emb(ids)          # pull: fused peer-memory gather (K2)
    loss.backward()                # push + owner-side optimizer (K3), inside backward

The dense part of the model is ordinary data parallelism (parallel/collective).

Reference parity: PS mode = `spec.ps` present (controllers/paddlejob_helper.go:
191-193); roles PSERVER/TRAINER (api/v1/paddlejob_types.go:44-48); what they do:
docs/design-arch.md:7-9,49-74 ("trainers ... pull parameters ... push gradients;
pservers update parameters", sync vs async SGD).  Mapping used here: shard j of
the table lives on GPU j mod ngpus, with P = `ps.replicas` shards (`num_shards`;
PADDLE_PSERVERS_IP_PORT_LIST names them); the PSERVER ranks still exist as processes for
ordering/status, but the "server" is memory plus the owner-side half of the kernels, and
checkpoints are one `ps-<j>.pt` per shard.
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional

import torch

from ..ops import _lib
from ..ops._lib import PdjEmbTable, check
from .symm import Communicator

SGD, ADAGRAD = 0, 1


def embedding_reference(table_deep, table_wide, ids):
    """fp32 oracle: table_* are the FULL (unsharded) tables."""
    deep = table_deep[ids]  # [B, S, D]
    wide = table_wide[ids].sum(dim=1) if table_wide is not None else None
    return deep, wide


class ShardedEmbedding:
    def __init__(self, comm: Communicator, num_rows: int, dim: int, max_batch: int, slots: int,
                 has_wide: bool = True, optimizer: int = ADAGRAD, lr: float = 0.05, eps: float = 1e-6,
                 async_sgd: bool = False, init_scale: float = 0.01, seed: int = 0, nblocks: int = 0,
                 num_shards: Optional[int] = None):
        """`num_shards` = P, the PSERVER count of the job (spec.ps.replicas; the reference sizes its server side
        from it, controllers/paddlejob_helper.go:225-235,258-260).  Shard j = id % P lives on GPU j % nranks; a GPU
        holds ceil(P / nranks) shard slots.  Default: one shard per rank."""
        self.comm, self.dim, self.slots, self.max_batch = comm, dim, slots, max_batch
        self.nranks = comm.nranks
        self.num_shards = int(num_shards) if num_shards else self.nranks
        self.slots_per_rank = (self.num_shards + self.nranks - 1) // self.nranks
        self.shard_rows = (num_rows + self.num_shards - 1) // self.num_shards   # rows of one shard
        self.rows_per_shard = self.shard_rows * self.slots_per_rank             # rows one GPU holds
        self.num_rows = self.shard_rows * self.num_shards
        self.has_wide, self.optimizer, self.lr, self.eps, self.async_sgd = has_wide, optimizer, lr, eps, async_sgd
        self.nblocks = nblocks or 128
        self.cstruct = comm.channel_struct(1)  # never shares barrier slots with the dense all-reduce
        self.grad_scale = 1.0  # trainers set 1/world so that summed pushes equal the global-batch mean
        self.lib = _lib.load()
        R, D = self.rows_per_shard, dim
        a = comm.alloc
        t = PdjEmbTable()
        # One record per row (default), 128-byte aligned:
        #   head (64 B for dim 16): | bf16 copy of the deep weights | w wide | acc wide | g wide | pad |   <- a lookup reads
        #   body:                   | g deep | w deep fp32 | acc deep |                                     only the head
        # so a lookup is ONE 64-byte request per row (it used to be a 64-byte and a 32-byte one, and twice the bytes
        # over NVLink), and the optimizer pass reads and writes whole aligned records instead of six arrays.  The bf16
        # view is maintained by the apply kernel; asynchronous SGD updates weights with remote reductions and cannot
        # keep it, so it gets none.  PDJ_EMB_LAYOUT=arrays keeps six separate arrays, =record_fp32 the record
        # without the bf16 view (the A/B arms).
        self.record_floats = 0
        self.has_shadow = False
        layout = os.environ.get("PDJ_EMB_LAYOUT", "record")
        t.shadow_off = 0
        if layout in ("record", "record_fp32"):
            self.has_shadow = layout == "record" and not async_sgd and 8 <= D <= 64
            head = (2 * D + 12 + 63) // 64 * 64 if self.has_shadow else 0
            if self.has_shadow:
                self.record_floats = (head + 12 * D + 127) // 128 * 32
                base = a(R * self.record_floats * 4)
                t.shadow_off = base
                t.wide_off, t.acc_wide_off, t.gwide_off = base + 2 * D, base + 2 * D + 4, base + 2 * D + 8
                # g deep directly after the head: a push touches g wide (head) and g deep in the same 128-byte line
                t.gdeep_off, t.deep_off, t.acc_deep_off = base + head, base + head + 4 * D, base + head + 8 * D
            else:
                self.record_floats = ((3 * D + 12) * 4 + 127) // 128 * 32
                base = a(R * self.record_floats * 4)
                t.deep_off, t.wide_off = base, base + D * 4
                t.acc_deep_off, t.acc_wide_off = base + (D + 4) * 4, base + (2 * D + 4) * 4
                t.gdeep_off, t.gwide_off = base + (2 * D + 8) * 4, base + (3 * D + 8) * 4
            self._rec_base = base
        else:
            t.deep_off = a(R * D * 4)
            t.wide_off = a(R * 4)
            t.gdeep_off = a(R * D * 4)
            t.gwide_off = a(R * 4)
            t.acc_deep_off = a(R * D * 4)
            t.acc_wide_off = a(R * 4)
        self._flag_bytes = (R + 511) // 512 * 512  # the apply kernel scans 512-row windows, 16 flags per lane
        t.flag_off = a(self._flag_bytes)  # one byte per row: "touched this step"
        t.list_off = self.record_floats * 4  # record stride in bytes; 0 = separate arrays
        t.count_off = a(256)
        t.rows_per_shard, t.list_cap, t.dim, t.has_wide = self.shard_rows, self.num_shards, D, int(has_wide)
        self.t = t
        # per-step I/O buffers (also in the arena so the emulation can address every rank's)
        self.ids_off = a(max_batch * slots * 8)
        # rows of the lookup output may be wider than slots * dim (set_output_layout): room for 32 extra columns
        self.out_pitch, self.dense_n, self.dense_off = slots * D, 0, 0
        self.out_deep_off = a(max_batch * (slots * D + 32) * 2)
        self.out_wide_off = a(max_batch * 4)
        self.gdeep_in_off = a(max_batch * slots * D * 2)
        self.gwide_in_off = a(max_batch * 4)
        self._init(init_scale, seed)

    # -- views -----------------------------------------------------------------
    def _ranks(self):
        return range(self.nranks) if self.comm.emulate else [self.comm.rank]

    def _part(self, off: int, width: int, rank: Optional[int] = None) -> torch.Tensor:
        """[rows_per_shard, width] fp32 view of one part of the table (width 0: a vector).  Strided in the record
        layout, so index / copy_ / clone it - do not .view() it."""
        R, F = self.rows_per_shard, self.record_floats
        if not F:
            return self.comm.tensor(off, (R, width) if width else (R,), torch.float32, rank)
        rec = self.comm.tensor(self._rec_base, (R, F), torch.float32, rank)
        c = (off - self._rec_base) // 4
        return rec[:, c:c + width] if width else rec[:, c]

    def _shadow(self, rank: Optional[int] = None) -> torch.Tensor:
        """[rows_per_shard, dim] bf16 view of the read view at the head of the records."""
        rec = self.comm.tensor(self._rec_base, (self.rows_per_shard, self.record_floats * 2), torch.bfloat16, rank)
        return rec[:, :self.dim]

    def refresh_shadow(self, rank: Optional[int] = None) -> None:
        """Re-derives the bf16 read view from the fp32 weights.  Call after writing `shard()` directly (the apply
        kernel, `_init` and the checkpoint loaders do it themselves)."""
        if self.has_shadow:
            self._shadow(rank).copy_(self.shard(rank))

    def shard(self, rank: Optional[int] = None) -> torch.Tensor:
        return self._part(self.t.deep_off, self.dim, rank)

    def shard_wide(self, rank: Optional[int] = None) -> torch.Tensor:
        return self._part(self.t.wide_off, 0, rank)

    def staging(self, rank: Optional[int] = None):
        """(deep, wide) gradient staging rows: the sums pushed this step, zero between steps."""
        return self._part(self.t.gdeep_off, self.dim, rank), self._part(self.t.gwide_off, 0, rank)

    def _init(self, scale: float, seed: int) -> None:
        dev = f"cuda:{self.comm.device}"
        for r in self._ranks():
            g = torch.Generator(device=dev).manual_seed(seed * 1000 + r)
            deep = torch.randn(self.rows_per_shard, self.dim, generator=g, device=dev) * scale
            wide = torch.randn(self.rows_per_shard, generator=g, device=dev) * scale
            if self.record_floats:
                self.comm.tensor(self._rec_base, (self.rows_per_shard * self.record_floats,), torch.float32, r).zero_()
            else:
                for part in (*self.staging(r), *self._acc(r)):
                    part.zero_()
            self.shard(r).copy_(deep)
            self.shard_wide(r).copy_(wide)
            self.refresh_shadow(r)
            for off, n in ((self.t.flag_off, self._flag_bytes // 4), (self.t.count_off, 64)):
                self.comm.tensor(off, (n,), torch.float32, r).zero_()

    def shards_of(self, rank: int) -> list:
        """The shard indices GPU `rank` holds, in slot order."""
        return [j for j in range(rank, self.num_shards, self.nranks)]

    def full_tables(self):
        """Gathers the whole table on this device (tests / checkpoints); needs the emulation
        or peer mappings, both of which let any rank read any shard."""
        deep = torch.empty(self.num_rows, self.dim, device=f"cuda:{self.comm.device}")
        wide = torch.empty(self.num_rows, device=f"cuda:{self.comm.device}")
        P, R = self.num_shards, self.shard_rows
        for r in range(self.nranks):
            for k, j in enumerate(self.shards_of(r)):
                deep[j::P] = self.shard(r)[k * R:(k + 1) * R]
                wide[j::P] = self.shard_wide(r)[k * R:(k + 1) * R]
        return deep, wide

    # -- raw kernel launches (rank-local unless emulated) -----------------------------
    def ids_buffer(self, batch: int, rank: Optional[int] = None) -> torch.Tensor:
        return self.comm.tensor(self.ids_off, (batch, self.slots), torch.int64, rank)

    def set_output_layout(self, pitch: int, dense_n: int = 0) -> None:
        """Makes the lookup write rows of `pitch` bf16 values: the slots * dim embedding values, then `dense_n` dense
        features taken (fp32 -> bf16) from `dense_buffer()`, then zeros - the padded input of the MLP tower in
        place (csrc/embedding.cu::emb_write_tail).  Backward then reads gradients with the same pitch."""
        used = self.slots * self.dim
        if pitch % 8 != 0 or not used + dense_n <= pitch <= used + 32:
            raise ValueError(f"pitch {pitch} must be a multiple of 8 in [{used + dense_n}, {used + 32}]")
        self.out_pitch, self.dense_n = pitch, dense_n
        if dense_n and not self.dense_off:
            self.dense_off = self.comm.alloc(self.max_batch * dense_n * 4)

    def dense_buffer(self, batch: int, rank: Optional[int] = None) -> torch.Tensor:
        return self.comm.tensor(self.dense_off, (batch, self.dense_n), torch.float32, rank)

    def lookup_raw(self, batch: int, barrier: bool = True, stream=None) -> None:
        c = self.comm
        check(self.lib.pdj_emb_lookup2(ctypes.byref(self.cstruct), ctypes.byref(self.t), self.ids_off,
                                       self.out_deep_off, self.out_wide_off, batch, self.slots, int(barrier),
                                       self.out_pitch, self.dense_off, self.dense_n,
                                       self.nblocks, int(c.emulate), c._stream(stream)), "emb_lookup")

    def push_raw(self, batch: int, grad_scale: float = 1.0, stream=None, gdeep_off: Optional[int] = None,
                 g_pitch: Optional[int] = None, gwide_off: Optional[int] = None) -> None:
        """`gdeep_off` / `g_pitch`: read the deep gradient from there instead of the staging buffer.  On a real rank
        any device address works (pass address - arena base: the kernel adds the base back), which lets backward
        hand the MLP's input gradient over without a copy."""
        c = self.comm
        check(self.lib.pdj_emb_push2(ctypes.byref(self.cstruct), ctypes.byref(self.t), self.ids_off,
                                     self.gdeep_in_off if gdeep_off is None else gdeep_off,
                                     self.gwide_in_off if gwide_off is None else gwide_off, batch, self.slots,
                                     int(self.async_sgd), self.lr, grad_scale,
                                     self.slots * self.dim if g_pitch is None else g_pitch,
                                     self.nblocks, int(c.emulate), c._stream(stream)), "emb_push")
        if not self.async_sgd:
            check(self.lib.pdj_emb_apply(ctypes.byref(self.cstruct), ctypes.byref(self.t), self.optimizer, self.lr,
                                         self.eps, 4 * 148, int(c.emulate), c._stream(stream)), "emb_apply")

    def out_deep(self, batch: int, rank: Optional[int] = None) -> torch.Tensor:
        return self.comm.tensor(self.out_deep_off, (batch, self.out_pitch), torch.bfloat16, rank)

    def out_wide(self, batch: int, rank: Optional[int] = None) -> torch.Tensor:
        return self.comm.tensor(self.out_wide_off, (batch,), torch.float32, rank)

    def grad_deep_in(self, batch: int, rank: Optional[int] = None) -> torch.Tensor:
        return self.comm.tensor(self.gdeep_in_off, (batch, self.slots * self.dim), torch.bfloat16, rank)

    def grad_wide_in(self, batch: int, rank: Optional[int] = None) -> torch.Tensor:
        return self.comm.tensor(self.gwide_in_off, (batch,), torch.float32, rank)

    def touched_last_step(self, rank: Optional[int] = None) -> int:
        return int(self.comm.tensor(self.t.count_off, (2,), torch.int32, rank)[1].item())

    # -- autograd face (one real rank) --------------------------------------------------
    def __call__(self, ids: torch.Tensor):
        """ids [B, S] int64 on this rank's device -> (deep bf16 [B, S*D], wide fp32 [B]).
        The sparse optimizer step runs inside backward."""
        return _LookupFn.apply(self, ids, torch.zeros((), device=ids.device, requires_grad=True))

    def _acc(self, rank: Optional[int] = None):
        return self._part(self.t.acc_deep_off, self.dim, rank), self._part(self.t.acc_wide_off, 0, rank)

    def shard_state_dicts(self) -> dict:
        """{shard index j: state} for the shards THIS rank holds - one `ps-<j>.pt` each (utils/checkpoint.py), named
        by pserver index as the reference's server side would be, whatever the number of trainers."""
        r, R = self.comm.rank, self.shard_rows
        acc_d, acc_w = self._acc()
        out = {}
        for k, j in enumerate(self.shards_of(r)):
            sl = slice(k * R, (k + 1) * R)
            out[j] = {"deep": self.shard(r)[sl].clone(), "wide": self.shard_wide(r)[sl].clone(),
                      "acc_deep": acc_d[sl].clone(), "acc_wide": acc_w[sl].clone(), "shard": j,
                      "nshards": self.num_shards}
        return out

    def load_shard_state_dict(self, sd: dict) -> None:
        j = int(sd["shard"])
        if int(sd.get("nshards", self.num_shards)) != self.num_shards or j % self.nranks != self.comm.rank:
            raise ValueError(f"shard {j} of {sd.get('nshards')} does not belong on rank {self.comm.rank} "
                             f"of a {self.num_shards}-shard table")
        k, R = j // self.nranks, self.shard_rows
        sl = slice(k * R, (k + 1) * R)
        acc_d, acc_w = self._acc()
        self.shard()[sl].copy_(sd["deep"])
        self.shard_wide()[sl].copy_(sd["wide"])
        acc_d[sl].copy_(sd["acc_deep"])
        acc_w[sl].copy_(sd["acc_wide"])
        self.refresh_shadow()

    def state_dict(self) -> dict:
        """Everything this rank holds in one dict (one shard per rank: the round-1 format)."""
        r = self.comm.rank
        acc_d, acc_w = self._acc()
        return {"deep": self.shard(r).clone(), "wide": self.shard_wide(r).clone(), "acc_deep": acc_d.clone(),
                "acc_wide": acc_w.clone(), "shard": r, "nshards": self.num_shards}

    def load_state_dict(self, sd: dict) -> None:
        acc_d, acc_w = self._acc()
        self.shard().copy_(sd["deep"])
        self.shard_wide().copy_(sd["wide"])
        acc_d.copy_(sd["acc_deep"])
        acc_w.copy_(sd["acc_wide"])
        self.refresh_shadow()


class _LookupFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, emb: ShardedEmbedding, ids: torch.Tensor, _anchor: torch.Tensor):
        b = ids.shape[0]
        buf = emb.ids_buffer(b)
        if ids.data_ptr() != buf.data_ptr() or ids.dtype != buf.dtype:  # a caller may keep its ids in the arena buffer itself
            buf.copy_(ids)
        emb.lookup_raw(b)
        ctx.emb, ctx.batch = emb, b
        return emb.out_deep(b), emb.out_wide(b)

    @staticmethod
    def backward(ctx, g_deep, g_wide):
        emb, b = ctx.emb, ctx.batch
        base = emb.comm._bases[emb.comm.rank]
        real = not emb.comm.emulate  # on a real rank any device address works: the kernel adds the arena base back
        gwide_off = None
        if real and g_wide is not None and g_wide.dtype == torch.float32 and g_wide.is_contiguous() and g_wide.data_ptr() % 4 == 0:
            gwide_off = (g_wide.data_ptr() - base) % (1 << 64)  # no staging copy of the wide gradient either
        else:
            emb.grad_wide_in(b).copy_(g_wide if g_wide is not None else torch.zeros(b, device=g_deep.device))
        pitch = g_deep.stride(0)
        if (real and g_deep.dim() == 2 and g_deep.stride(1) == 1 and pitch % 4 == 0
                and g_deep.data_ptr() % 8 == 0 and pitch >= emb.slots * emb.dim):
            # no staging copy: the push kernel reads the gradient tensor where autograd left it
            off = (g_deep.data_ptr() - base) % (1 << 64)
            emb.push_raw(b, emb.grad_scale, gdeep_off=off, g_pitch=pitch, gwide_off=gwide_off)
        else:
            emb.grad_deep_in(b).copy_(g_deep[:, :emb.slots * emb.dim])
            emb.push_raw(b, emb.grad_scale, gwide_off=gwide_off)
        return None, None, None
