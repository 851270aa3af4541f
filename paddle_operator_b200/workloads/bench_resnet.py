"""ResNet-50 img/s measurement used by bench.py (product and ddp arms).

Timing rules followed: >= 3 warm-up steps, CUDA events on the launching stream
bracketed by barrier + synchronize, max over ranks, clocks sampled during the
timed region, working set (activations, several GB per step) far larger than
the 126 MB L2 so no explicit flush is needed.
"""
from __future__ import annotations

import json
import time

import torch
import torch.distributed as dist


def _max_over_ranks(ms: float, world: int, dev) -> float:
    if world == 1:
        return ms
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def _sync_all(world: int, dev) -> None:
    if world > 1:
        dist.barrier(device_ids=[dev.index])
    torch.cuda.synchronize(dev)


def _host_batches(batch: int, image: int, n: int = 3):
    g = torch.Generator().manual_seed(1234)
    out = []
    for _ in range(n):
        x = torch.randint(0, 256, (batch, image, image, 3), dtype=torch.uint8, generator=g).pin_memory()
        y = torch.randint(0, 1000, (batch,), dtype=torch.int64, generator=g).pin_memory()
        out.append((x, y))
    return out


def _time_region(fn, steps: int, world: int, dev) -> float:
    _sync_all(world, dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1)
    _sync_all(world, dev)
    return _max_over_ranks(ms, world, dev)


def _build_product(args, world, dev):
    from ..parallel.symm import AR_NVLS, AR_TWOSHOT, Communicator
    from .resnet import ResNetJob

    comm = Communicator.from_env(arena_bytes=256 << 20) if world > 1 else None
    algo = {"auto": None, "twoshot": AR_TWOSHOT, "nvls": AR_NVLS}[args.algo]
    job = ResNetJob(batch=args.batch or 256, comm=comm, device=dev, use_graph=not args.no_graph,
                    bucket_bytes=int(args.bucket_mb * (1 << 20)), comm_blocks=args.comm_blocks, algo=algo)
    job.prepare()
    return job, comm


class _DdpJob:
    """The yardstick: stock torch modules, DDP/NCCL, fused torch SGD, eager."""

    def __init__(self, batch, world, dev):
        from ..models.resnet import resnet50, to_training_dtype

        torch.manual_seed(0)
        self.dev, self.batch = dev, batch
        model = to_training_dtype(resnet50(fused=False).to(dev), torch.bfloat16, bn_fp32=True).train()
        self.model = model
        if world > 1:
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

    def step_device(self):
        self.opt.zero_grad(set_to_none=True)
        loss = torch.nn.functional.cross_entropy(self.model(self.x).float(), self.y)
        loss.backward()
        self.opt.step()
        self.loss = loss.detach()
        return self.loss

    def train_step(self, images_u8, labels):
        xs = images_u8.to(self.dev, non_blocking=True)
        self.y = labels.to(self.dev, non_blocking=True)
        x = ((xs.float() / 255.0 - self.mean) / self.std).to(torch.bfloat16)
        self.x = x.permute(0, 3, 1, 2)
        loss = self.step_device()
        self._loss_host.copy_(loss.reshape(1), non_blocking=True)
        return self._loss_host

    h2d_bytes_per_step = property(lambda s: s.batch * 224 * 224 * 3 + s.batch * 8)
    d2h_bytes_per_step = 4
    launches_per_step = 0


def run(args, rank: int, world: int, dev, ClockSampler) -> int:
    batch = args.batch or 256
    comm = None
    if args.impl == "product":
        job, comm = _build_product(args, world, dev)
    else:
        job = _DdpJob(batch, world, dev)
    steps, warmup = args.steps, max(args.warmup, 3)

    # ---- device-timed: inputs resident on the device -------------------------
    for _ in range(warmup):
        job.step_device()
    sampler = ClockSampler() if rank == 0 else None
    if sampler:
        sampler.start()
        time.sleep(0.05)
    ms = _time_region(job.step_device, steps, world, dev)
    clocks = sampler.stop() if sampler else None
    loss_dev = float(job.loss.item()) if hasattr(job, "loss") else float(job.trainer.loss.item())

    # ---- end to end: pinned host batch in, loss out, every step --------------
    hb = _host_batches(batch, 224)
    k = [0]
    last = [None]

    def e2e_step():
        x, y = hb[k[0] % len(hb)]
        k[0] += 1
        last[0] = job.train_step(x, y)

    for _ in range(3):
        e2e_step()
    ms_e2e = _time_region(e2e_step, steps, world, dev)
    loss_e2e = float(last[0].item())

    imgs = batch * world * steps
    value = imgs / (ms / 1e3)
    e2e_value = imgs / (ms_e2e / 1e3)
    launches = job.launches_per_step * steps
    out = {
        "metric": "resnet50_train_images_per_sec",
        "value": round(value, 1),
        "unit": "img/s",
        "n_gpus": world,
        "steps": steps,
        "warmup": warmup,
        "ms_per_step": round(ms / steps, 3),
        "higher_is_better": True,
        "scaling": "weak",
        "vs_baseline": None,
        "dtype": "bf16",
        "data": "synthetic (random uint8 224x224 images, random labels; random-init weights)",
        "impl": args.impl,
        "config": {
            "model": "ResNet-50 v1.5, NHWC, SGD momentum 0.9 wd 1e-4, fp32 master weights"
                     + (" (7x7 stem evaluated as the equivalent 4x4 conv over space-to-depth input)"
                        if getattr(job, "s2d", False) else ""),
            "global_batch": batch * world,
            "per_gpu_batch": batch,
            "seq_len": None,
            "image": "3x224x224",
            "parallelism": f"dp{world}",
            "l2": "no explicit flush: each step streams several GB of activations, far larger than the 126 MB L2",
            "cuda_graph": bool(args.impl == "product" and not args.no_graph),
            "allreduce": (args.algo if world > 1 else "none"),
            "nvls": bool(comm is not None and comm.has_multicast),
        },
        "clocks": clocks,
        "e2e": {
            "value": round(e2e_value, 1),
            "unit": "img/s",
            "ms_per_step": round(ms_e2e / steps, 3),
            "h2d_bytes_per_step": job.h2d_bytes_per_step,
            "d2h_bytes_per_step": job.d2h_bytes_per_step,
        },
        "gpu_launches": launches,
        "loss": {"device": loss_dev, "e2e": loss_e2e},
    }
    if rank == 0:
        print(json.dumps(out), flush=True)
    if comm is not None:
        torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier(device_ids=[dev.index])
        dist.destroy_process_group()
    return 0
