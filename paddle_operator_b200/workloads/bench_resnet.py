"""ResNet-50 img/s measurement used by bench.py.

Arms (`measure(..., impl=...)`):
  product    this repository's engine (tcgen05 convolutions, fused BN kernels, fused NVSwitch
             all-reduce overlapped with backward, fused SGD, the whole step in a CUDA graph);
  ddp        the yardstick BASELINE.md names: the same network out of stock torch modules under
             torch DDP/NCCL (+cuDNN/cuBLAS), fused torch SGD, eager;
  ddp_graph  the same stock network with the whole step (forward, backward, NCCL all-reduce of the
             gradients, SGD) captured in a CUDA graph - separates "launch overhead removed by graph
             capture" from "faster kernels" in the product / baseline ratio.

Timing rules: see bench_common.py.  The working set (activations, several GB per step) is far larger
than the 126 MB L2, so no explicit flush is needed.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from .bench_common import job_key, measure_job, release


def _host_batches(batch: int, image: int, n: int = 3):
    g = torch.Generator().manual_seed(1234)
    out = []
    for _ in range(n):
        x = torch.randint(0, 256, (batch, image, image, 3), dtype=torch.uint8, generator=g).pin_memory()
        y = torch.randint(0, 1000, (batch,), dtype=torch.int64, generator=g).pin_memory()
        out.append((x, y))
    return out


def _build_product(args, world, dev):
    from ..parallel.symm import AR_NVLS, AR_TWOSHOT, Communicator
    from .resnet import ResNetJob

    comm = Communicator.from_env(arena_bytes=384 << 20, key=job_key("resnet")) if world > 1 else None
    algo = {"auto": None, "twoshot": AR_TWOSHOT, "nvls": AR_NVLS}[args.algo]
    job = ResNetJob(batch=args.batch or 256, comm=comm, device=dev, use_graph=not args.no_graph,
                    bucket_bytes=int(args.bucket_mb * (1 << 20)), comm_blocks=args.comm_blocks, algo=algo)
    job.prepare()
    return job, comm


class _DdpJob:
    """The yardstick: stock torch modules, NCCL gradient all-reduce, fused torch SGD.

    graph=False: torch DistributedDataParallel, eager (the arm BASELINE.md names).
    graph=True:  the whole step in one CUDA graph; gradients are all-reduced with one
                 dist.all_reduce per dtype on a flattened buffer inside the captured step."""

    def __init__(self, batch, world, dev, graph: bool = False):
        from ..models.resnet import resnet50, to_training_dtype

        torch.manual_seed(0)
        self.dev, self.batch, self.world, self.graph_mode = dev, batch, world, graph
        model = to_training_dtype(resnet50(fused=False).to(dev), torch.bfloat16, bn_fp32=True).train()
        self.net = model
        self.model = model
        if world > 1 and not graph:
            self.model = torch.nn.parallel.DistributedDataParallel(model, device_ids=[dev.index],
                                                                   gradient_as_bucket_view=True)
        self.opt = torch.optim.SGD(model.parameters(), lr=0.1, momentum=0.9, weight_decay=1e-4, fused=True)
        self.x = torch.randn((batch, 3, 224, 224), device=dev).to(torch.bfloat16).contiguous(
            memory_format=torch.channels_last)
        self.y = torch.randint(0, 1000, (batch,), device=dev)
        self.mean = torch.tensor([0.485, 0.456, 0.406], device=dev).view(1, 1, 1, 3)
        self.std = torch.tensor([0.229, 0.224, 0.225], device=dev).view(1, 1, 1, 3)
        self.loss = torch.zeros((), device=dev)
        self._loss_host = torch.zeros(1).pin_memory()
        self._graph = None
        if graph:
            self._capture()

    def _allreduce_grads(self) -> None:
        by_dtype: dict = {}
        for p in self.net.parameters():
            by_dtype.setdefault(p.grad.dtype, []).append(p.grad)
        for grads in by_dtype.values():
            flat = torch.cat([g.reshape(-1) for g in grads])
            dist.all_reduce(flat)
            flat.mul_(1.0 / self.world)
            views, off = [], 0
            for g in grads:
                views.append(flat[off:off + g.numel()].view(g.shape))
                off += g.numel()
            torch._foreach_copy_(grads, views)

    def _body(self) -> None:
        loss = torch.nn.functional.cross_entropy(self.model(self.x).float(), self.y)
        loss.backward()
        if self.graph_mode and self.world > 1:
            self._allreduce_grads()
        self.opt.step()
        self.loss.copy_(loss.detach())

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
        # grads are None at capture: backward allocates them from the graph's pool and every replay
        # overwrites them (the documented whole-network capture recipe)
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

    def train_step(self, images_u8, labels):
        xs = images_u8.to(self.dev, non_blocking=True)
        self.y.copy_(labels.to(self.dev, non_blocking=True))
        x = ((xs.float() / 255.0 - self.mean) / self.std).to(torch.bfloat16)
        self.x.copy_(x.permute(0, 3, 1, 2))
        loss = self.step_device()
        self._loss_host.copy_(loss.reshape(1), non_blocking=True)
        return self._loss_host

    h2d_bytes_per_step = property(lambda s: s.batch * 224 * 224 * 3 + s.batch * 8)
    d2h_bytes_per_step = 4
    launches_per_step = 0


IMPL_NAMES = {
    "product": "this repository (tcgen05 convolutions, fused BN, fused NVSwitch all-reduce, CUDA graph)",
    "ddp": "torch DDP/NCCL eager (stock modules, cuDNN NHWC, fused SGD)",
    "ddp_graph": "stock torch modules + NCCL gradient all-reduce + fused SGD, whole step in a CUDA graph",
}


def measure(args, impl: str, rank: int, world: int, dev, ClockSampler) -> dict:
    """Runs one arm and returns its record (no printing, the process group stays up)."""
    batch = args.batch or 256
    comm = None
    if impl == "product":
        job, comm = _build_product(args, world, dev)
    else:
        job = _DdpJob(batch, world, dev, graph=(impl == "ddp_graph"))
    steps, warmup = args.steps, max(args.warmup, 3)
    rec = measure_job(job, _host_batches(batch, 224), steps, warmup, world, rank, dev, ClockSampler, batch * world)
    rec.update({
        "metric": "resnet50_train_images_per_sec",
        "unit": "img/s",
        "impl": impl,
        "impl_name": IMPL_NAMES[impl],
        "config": {
            "model": "ResNet-50 v1.5, NHWC, SGD momentum 0.9 wd 1e-4"
                     + (", fp32 master weights" if impl == "product" else "")
                     + (" (7x7 stem evaluated as the equivalent 4x4 conv over space-to-depth input)"
                        if getattr(job, "s2d", False) else ""),
            "global_batch": batch * world,
            "per_gpu_batch": batch,
            "seq_len": None,
            "image": "3x224x224",
            "parallelism": f"dp{world}",
            "l2": "no explicit flush: each step streams several GB of activations, far larger than the 126 MB L2",
            "cuda_graph": bool((impl == "product" and not args.no_graph) or impl == "ddp_graph"),
            "allreduce": (args.algo if (world > 1 and impl == "product") else ("nccl" if world > 1 else "none")),
            "nvls": bool(comm is not None and comm.has_multicast),
        },
    })
    torch.cuda.synchronize(dev)
    del job
    if comm is not None:
        comm.close()
    release(dev)
    return rec
