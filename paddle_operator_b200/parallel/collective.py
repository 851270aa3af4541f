"""Collective-mode (data-parallel) training engine.

One process per GPU.  Parameters, fp32 masters, momentum and gradients live in
flat buffers; the gradient buffer sits in the symmetric arena so that the fused
all-reduce kernels (csrc/collectives.cu) can reduce it in place across ranks,
bucket by bucket, on a side stream while backward is still running.  The whole
step (forward, loss, backward, bucket all-reduces, fused SGD) is captured once
in a CUDA graph and replayed.

Reference parity: "Collective" mode of the reference is worker.replicas > 1
with no ps (controllers/paddlejob_helper.go:191-199); it hands each worker
PADDLE_TRAINER_ID / PADDLE_TRAINERS_NUM / PADDLE_TRAINER_ENDPOINTS
(controllers/paddlejob_helper.go:261-265,319-322) and Paddle's fleet does the
rest with NCCL.  This module is "the rest", B200-first.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable, Optional

import torch

from ..ops import optim as pdj_optim
from ..ops import _sink
from ..ops._sink import is_sunk
from ..utils import trace
from .symm import Communicator


def _round_up(n: int, m: int) -> int:
    return (n + m - 1) // m * m


@dataclass
class _Bucket:
    start: int  # element offset in the flat buffers
    end: int
    params: list = field(default_factory=list)
    gviews: list = field(default_factory=list)
    pending: int = 0
    grads: list = field(default_factory=list)
    event: Optional[torch.cuda.Event] = None


class FlatParams:
    """Flat bf16 parameter / gradient buffers with fp32 master + momentum.

    Layout follows reversed(model.parameters()) so that buckets complete in
    address order during backward.  Every parameter starts on a 16-byte
    boundary.  Convolution weights keep channels-last strides inside the flat
    buffer.
    """

    def __init__(self, model: torch.nn.Module, device, comm: Optional[Communicator], bucket_bytes: int):
        params = [p for p in model.parameters() if p.requires_grad]
        order = list(reversed(params))
        offs, cur = [], 0
        for p in order:
            offs.append(cur)
            cur += _round_up(p.numel(), 8)
        self.numel = _round_up(cur, 8 * 8)  # whole 16-byte vectors per rank for up to 8 ranks
        self.comm = comm
        self.device = device
        self._order, self._offs = order, offs
        self._allocate(comm)
        self.buckets: list[_Bucket] = []
        self.views: dict[int, tuple[torch.Tensor, torch.Tensor, torch.Tensor]] = {}
        cap = max(bucket_bytes // 2, 8)
        b = _Bucket(0, 0)
        for p, off in zip(order, offs):
            n = p.numel()
            pv = self._view(self.param, off, p)
            gv = self._view(self.grad, off, p)
            mv = self.master[off:off + n]
            with torch.no_grad():
                pv.copy_(p.data)
                mv.copy_(p.data.reshape(-1) if p.dim() != 4 else p.data.permute(0, 2, 3, 1).reshape(-1))
            p.data = pv
            p._grad_sink = gv  # kernels write this parameter's gradient here directly (ops/_sink.py)
            self.views[id(p)] = (pv, gv, mv)
            b.params.append(p)
            b.gviews.append(gv)
            b.end = off + _round_up(n, 8)
            if b.end - b.start >= cap:
                self.buckets.append(b)
                b = _Bucket(b.end, b.end)
        if b.params:
            self.buckets.append(b)
        self.buckets[-1].end = self.numel
        # masters were filled from the (bf16 or fp32) init values; make param == bf16(master)
        self.param.copy_(self.master.to(torch.bfloat16))
        self.bucket_of = {id(p): b for b in self.buckets for p in b.params}

    def _allocate(self, comm: Optional[Communicator]) -> None:
        """The four flat buffers.  With a communicator they ALL live in the symmetric arena, at offsets that are the
        same on every rank: the fused bucket kernel (csrc/collectives.cu, K1b) reads peers' gradients and writes
        peers' parameters and masters directly."""
        n = self.numel
        if comm is not None:
            self.param_off, self.grad_off = comm.alloc(n * 2), comm.alloc(n * 2)
            self.master_off, self.mom_off = comm.alloc(n * 4), comm.alloc(n * 4)
            self.param = comm.tensor(self.param_off, (n,), torch.bfloat16)
            self.grad = comm.tensor(self.grad_off, (n,), torch.bfloat16)
            self.master = comm.tensor(self.master_off, (n,), torch.float32)
            self.momentum = comm.tensor(self.mom_off, (n,), torch.float32)
            for t in (self.param, self.grad, self.master, self.momentum):
                t.zero_()
        else:
            self.param_off = self.grad_off = self.master_off = self.mom_off = 0
            self.param = torch.zeros(n, dtype=torch.bfloat16, device=self.device)
            self.grad = torch.zeros(n, dtype=torch.bfloat16, device=self.device)
            self.master = torch.zeros(n, dtype=torch.float32, device=self.device)
            self.momentum = torch.zeros(n, dtype=torch.float32, device=self.device)

    def rehome(self, comm: Optional[Communicator]) -> None:
        """Moves the buffers into `comm`'s arena (or into private memory for comm = None), keeping their contents,
        and re-points the parameters and the per-parameter views.  The old communicator must still be open."""
        old = (self.param, self.master, self.momentum)
        self.comm = comm
        self._allocate(comm)
        with torch.no_grad():
            self.param.copy_(old[0])
            self.master.copy_(old[1])
            self.momentum.copy_(old[2])
        for b in self.buckets:
            b.gviews = []
        for p, off in zip(self._order, self._offs):
            pv, gv = self._view(self.param, off, p), self._view(self.grad, off, p)
            p.data = pv
            p._grad_sink = gv
            self.views[id(p)] = (pv, gv, self.master[off:off + p.numel()])
            self.bucket_of[id(p)].gviews.append(gv)

    @staticmethod
    def _view(flat: torch.Tensor, off: int, p: torch.Tensor) -> torch.Tensor:
        n = p.numel()
        s = flat[off:off + n]
        if p.dim() == 4:
            o, i, kh, kw = p.shape
            return s.view(o, kh, kw, i).permute(0, 3, 1, 2)  # channels-last strides
        return s.view(p.shape)

    def master_view(self, p: torch.Tensor) -> torch.Tensor:
        return self.views[id(p)][2]


class CollectiveTrainer:
    """Data-parallel SGD trainer with fused all-reduce and CUDA-graph replay.

    loss_fn(model, *inputs) -> scalar loss tensor.  `example_inputs` are static
    device tensors; `step()` replays the captured graph on whatever they hold.
    """

    def __init__(self, model: torch.nn.Module, loss_fn: Callable, example_inputs: tuple,
                 comm: Optional[Communicator] = None, lr: float = 0.1, momentum: float = 0.9,
                 weight_decay: float = 1e-4, bucket_bytes: int = 16 << 20, use_graph: bool = True,
                 comm_blocks: int = 24, algo: Optional[int] = None, extra_kernels_per_step: int = 0):
        self.model = model
        self.loss_fn = loss_fn
        self.inputs = example_inputs
        self.comm = comm if (comm is not None and comm.nranks > 1) else None
        self.world = comm.nranks if comm is not None else 1
        self.device = example_inputs[0].device
        self.flat = FlatParams(model, self.device, self.comm, bucket_bytes)
        self.hyper = torch.tensor([lr, momentum, weight_decay, 1.0], dtype=torch.float32, device=self.device)
        self._hyper_host = torch.empty(4, dtype=torch.float32).pin_memory() if self.device.type == "cuda" else None
        self.loss = torch.zeros((), dtype=torch.float32, device=self.device)
        self.use_graph = use_graph and self.device.type == "cuda"
        self.comm_blocks = comm_blocks
        self.algo = algo
        self.graph: Optional[torch.cuda.CUDAGraph] = None
        self.kernels_per_step = 0  # launches of OUR kernels per step (counted at capture/eager time)
        self.extra_kernels_per_step = extra_kernels_per_step  # the job's own kernels outside modules (loss, embedding)
        self._count = 0
        self.comm_stream = torch.cuda.Stream(self.device) if self.comm is not None else None
        self.on_rehome: list = []  # callbacks of the job that re-attach views of the flat buffers after a re-form
        self._attach_masters()
        self._install_hooks()

    # -- setup ---------------------------------------------------------------
    def _attach_masters(self) -> None:
        from ..ops.bn import FusedBnAct

        for m in self.model.modules():
            if isinstance(m, FusedBnAct):
                m.weight_f32 = self.flat.master_view(m.weight)
                m.bias_f32 = self.flat.master_view(m.bias)

    def _install_hooks(self) -> None:
        for b in self.flat.buckets:
            for p in b.params:
                p.register_post_accumulate_grad_hook(self._make_hook(b))

    def _make_hook(self, bucket: _Bucket):
        def hook(p: torch.Tensor):
            gv = self.flat.views[id(p)][1]
            if not is_sunk(p.grad, gv):  # else the producing kernel already wrote it in place
                bucket.grads.append((gv, p.grad))
            p.grad = None
            bucket.pending -= 1
            if bucket.pending == 0:
                self._bucket_ready(bucket)

        return hook

    def _bucket_ready(self, b: _Bucket) -> None:
        if b.grads:
            _sink.wait_side_work(torch.cuda.current_stream()) if self.device.type == "cuda" else None
            dst = [d for d, _ in b.grads]
            src = [s for _, s in b.grads]
            torch._foreach_copy_(dst, src)
            b.grads.clear()
        if self.comm is None:
            return
        main = torch.cuda.current_stream()
        ev = torch.cuda.Event()
        ev.record(main)
        self.comm_stream.wait_event(ev)
        _sink.wait_side_work(self.comm_stream)  # parameter gradients computed on a side stream (ops/_sink.py)
        n = b.end - b.start
        f = self.flat
        # one kernel per bucket: reduce-scatter + SGD-momentum on the owner's slice + broadcast of the new master and
        # parameter (K1b).  Reduced gradients never come back and nobody runs an optimizer pass over the whole model.
        with trace.span("bucket_allreduce_sgd", bytes=n * 2):
            self.comm.allreduce_sgd(f.grad_off, f.param_off, f.master_off, f.mom_off, b.start, n, f.numel,
                                    self.hyper.data_ptr(), 1.0 / self.world, algo=self.algo, nblocks=self.comm_blocks,
                                    stream=self.comm_stream)
        self._count += 1
        done = torch.cuda.Event()
        done.record(self.comm_stream)
        b.event = done

    # -- one step ---------------------------------------------------------------
    def _step_body(self) -> None:
        from ..ops import conv as _conv

        conv0 = _conv.launch_count()
        for b in self.flat.buckets:
            b.pending = len(b.params)
            b.event = None
        cuda = self.device.type == "cuda"
        if cuda:
            # side-stream parameter gradients pay on one rank (0.544 -> 0.518 ms per Wide&Deep step); with a
            # communicator the bucket collective already overlaps the embedding push and the extra stream only adds
            # contention (0.699 -> 0.710 ms at N = 2), so it stays off there (PDJ_WGRAD_OVERLAP=2 forces it)
            _sink.begin_step(self.comm is None or _sink.WGRAD_OVERLAP_FORCE)
        try:
            loss = self.loss_fn(self.model, *self.inputs)
            loss.backward()
        except BaseException:
            if cuda:
                _sink.end_step()  # never leave the side-stream scope open behind a failed step
            raise
        main = torch.cuda.current_stream() if cuda else None
        if cuda:
            _sink.wait_side_work(main)  # the optimizer below / the next step read what the side stream wrote
        for b in self.flat.buckets:
            if b.pending != 0:
                raise RuntimeError("a parameter received no gradient; every parameter must take part in the step")
            if b.event is not None:
                main.wait_event(b.event)
        if self.comm is None:
            pdj_optim.sgd_momentum_(self.flat.master, self.flat.momentum, self.flat.grad, self.flat.param, self.hyper,
                                    self.flat.numel)
            self._count += 1
        # (with a communicator the update already happened, bucket by bucket, inside the fused collective)
        self._count += _conv.launch_count() - conv0  # tcgen05 convolution kernels launched by the model this step
        self.loss.copy_(loss.detach().float())
        if cuda:
            _sink.end_step()

    def capture(self, warmup: int = 2) -> None:
        """Runs `warmup` eager steps on a side stream, then captures the step."""
        if not self.use_graph:
            return
        s = torch.cuda.Stream(self.device)
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for _ in range(warmup):
                self._step_body()
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize(self.device)
        if self.comm is not None:
            self.comm.barrier()
            torch.cuda.synchronize(self.device)
        self.graph = torch.cuda.CUDAGraph()
        self._count = 0
        with torch.cuda.graph(self.graph):
            self._step_body()
        self.kernels_per_step = self._count + self._model_kernel_count()
        torch.cuda.synchronize(self.device)

    def _model_kernel_count(self) -> int:
        """Hand-written kernel launches per step inside the model (fused BN:
        2 forward + 2 backward per layer, max-pool 1 + 1) plus what the job declares."""
        from ..ops.bn import FusedBnAct

        from ..ops.pool import FusedMaxPool2d

        n = sum(4 for m in self.model.modules() if isinstance(m, FusedBnAct))
        n += sum(2 for m in self.model.modules() if isinstance(m, FusedMaxPool2d))
        return n + self.extra_kernels_per_step

    def step(self) -> torch.Tensor:
        """Runs one optimisation step on the current contents of the static
        inputs; returns the (device) loss tensor of that step."""
        if self.graph is not None:
            self.graph.replay()
        else:
            self._count = 0
            self._step_body()
            self.kernels_per_step = self._count + self._model_kernel_count()
        return self.loss

    def set_lr(self, lr: float) -> None:
        if self._hyper_host is not None:
            self._hyper_host.copy_(torch.tensor([lr, 0, 0, 0]))
            self.hyper[0:1].copy_(self._hyper_host[0:1], non_blocking=True)
        else:
            self.hyper[0] = lr

    # -- elastic re-form -----------------------------------------------------------
    def reform(self, comm: Optional[Communicator]) -> None:
        """Switches to a new communicator (new world size).  The flat buffers live in the arena, so they are copied
        device-to-device into the new one (the OLD communicator must still be open: create the new one, call this,
        then close the old one); BN statistics stay where they are.  The step graph is captured again for the new
        pointers.  Call sync_state() on every rank of the old group first: momentum is only current on its owner."""
        torch.cuda.synchronize(self.device)
        self.graph = None
        self.comm = comm if (comm is not None and comm.nranks > 1) else None
        self.world = comm.nranks if comm is not None else 1
        self.comm_stream = torch.cuda.Stream(self.device) if self.comm is not None else None
        self.flat.rehome(self.comm)
        self._attach_masters()
        for fn in self.on_rehome:
            fn()
        if self.use_graph:
            self.capture(warmup=0)

    def broadcast_state(self, root: int = 0, chunk_bytes: int = 0) -> None:
        """K6: every rank ends up with the root's masters, momentum, hyper-parameters and
        model buffers, moved GPU-to-GPU through the arena (no disk, no host).

        The two 100 MB flat buffers already live in the arena and are pulled from the root in place (one launch
        each).  The small tensors (hyper-parameters and a hundred BatchNorm statistics for a ResNet-50) are packed
        back to back, in 16-byte aligned segments, into one staging window of the arena and travel in one launch
        instead of one launch and one barrier each.  Call sync_state() on the old group first: the root must hold
        every rank's momentum slice."""
        if self.comm is None:
            return
        comm = self.comm
        # masters and momentum sit in the arena at the same offset on every rank: pulled from the root in place
        nbytes = self.flat.numel * 4
        comm.broadcast(root, self.flat.master_off, nbytes)
        comm.broadcast(root, self.flat.mom_off, nbytes)
        # the rest (hyper-parameters, BatchNorm statistics: ~100 small tensors) is packed into one staging window
        tensors = [self.hyper] + [b for _, b in self.model.named_buffers()]
        flats = []
        for t in tensors:
            f = t.detach().reshape(-1) if t.is_contiguous() else t.detach().contiguous().reshape(-1)
            flats.append((t, f.view(torch.uint8)))
        total = sum((f.numel() + 15) // 16 * 16 for _, f in flats)
        room = (comm.arena_bytes - comm._cursor - 4096) // 16 * 16
        cap = min(total, room) if chunk_bytes <= 0 else min(total, room, chunk_bytes // 16 * 16)
        if cap < 16:
            raise MemoryError("symmetric arena has no room left for the state broadcast")
        off = comm.alloc(cap)
        stage = comm.tensor(off, (cap,), torch.uint8)
        # segments: (flat tensor, start, length) in order; a window holds as many whole or partial segments as fit
        i, pos = 0, 0
        while i < len(flats):
            window, used = [], 0
            while i < len(flats) and used < cap:
                f = flats[i][1]
                n = min(f.numel() - pos, cap - used)
                n_al = (n + 15) // 16 * 16
                if used + n_al > cap:
                    n = (cap - used) // 16 * 16
                    n_al = n
                    if n == 0:
                        break
                window.append((i, pos, n, used))
                used += n_al
                pos += n
                if pos >= f.numel():
                    i, pos = i + 1, 0
            if comm.rank == root:
                for (k, p0, n, at) in window:
                    stage[at:at + n].copy_(flats[k][1][p0:p0 + n])
            comm.broadcast(root, off, (used + 15) // 16 * 16)
            if comm.rank != root:
                for (k, p0, n, at) in window:
                    flats[k][1][p0:p0 + n].copy_(stage[at:at + n])
            if i < len(flats):
                comm.barrier()  # the stage is reused by the next window
        if comm.rank != root:
            for t, f in flats:
                if not t.is_contiguous():
                    t.copy_(f.view(t.dtype).view(t.shape))
        with torch.no_grad():
            self.flat.param.copy_(self.flat.master.to(torch.bfloat16))
        torch.cuda.synchronize(self.device)

    def sync_state(self) -> None:
        """Collective (every rank, same point of the step sequence): makes the momentum buffer complete on every
        rank.  Inside the fused bucket kernel each element's momentum is kept by the rank that owns its slice
        only; before a checkpoint, a re-form or a state broadcast the owners' slices are pulled together
        (K6 per slice).  Masters and parameters are always complete everywhere."""
        if self.comm is None:
            return
        comm, f, nr = self.comm, self.flat, self.world
        for b in f.buckets:
            n_vec = (b.end - b.start) // 8
            per = (n_vec + nr - 1) // nr
            for r in range(nr):
                beg, end = r * per, min((r + 1) * per, n_vec)
                if end > beg:
                    comm.broadcast(r, f.mom_off + (b.start + beg * 8) * 4, (end - beg) * 32, nblocks=16)

    # -- state (checkpoint / elastic re-shard) -------------------------------------
    def state_dict(self) -> dict:
        return {"master": self.flat.master, "momentum": self.flat.momentum, "hyper": self.hyper,
                "buffers": {k: v for k, v in self.model.named_buffers()}}

    def load_state_dict(self, sd: dict) -> None:
        with torch.no_grad():
            self.flat.master.copy_(sd["master"])
            self.flat.momentum.copy_(sd["momentum"])
            self.hyper.copy_(sd["hyper"])
            self.flat.param.copy_(self.flat.master.to(torch.bfloat16))
            bufs = dict(self.model.named_buffers())
            for k, v in sd.get("buffers", {}).items():
                if k in bufs:
                    bufs[k].copy_(v)
