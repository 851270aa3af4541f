"""Timing helpers shared by the benchmark arms (bench.py).

Rules followed by every arm: >= 3 warm-up steps, CUDA events on the launching
stream bracketed by barrier + synchronize on both sides, max over ranks, clocks
sampled with nvidia-smi while the timed region runs.
"""
from __future__ import annotations

import gc
import os
import time

import torch
import torch.distributed as dist


def max_over_ranks(ms: float, world: int, dev) -> float:
    if world == 1:
        return ms
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sync_all(world: int, dev) -> None:
    if world > 1:
        dist.barrier(device_ids=[dev.index])
    torch.cuda.synchronize(dev)


def time_region(fn, steps: int, world: int, dev) -> float:
    """Device milliseconds for `steps` calls of fn, max over ranks."""
    sync_all(world, dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize(dev)
    ms = e0.elapsed_time(e1)
    sync_all(world, dev)
    return max_over_ranks(ms, world, dev)


def measure_job(job, host_batches, steps: int, warmup: int, world: int, rank: int, dev, ClockSampler,
                units_per_step: int) -> dict:
    """Device-timed and end-to-end throughput of one job object.

    `job.step_device()` runs a step on device-resident inputs; `job.train_step(*batch)` is the
    public API: pinned host batch in, loss out, every step.  `units_per_step` = samples per step over
    the whole job (all ranks)."""
    for _ in range(warmup):
        job.step_device()
    sampler = ClockSampler() if rank == 0 else None
    if sampler:
        sampler.start()
        time.sleep(0.05)
    ms = time_region(job.step_device, steps, world, dev)
    clocks = sampler.stop() if sampler else None
    loss_obj = job.loss if hasattr(job, "loss") else job.trainer.loss
    loss_dev = float(loss_obj.item())

    k, last = [0], [None]

    def e2e_step():
        b = host_batches[k[0] % len(host_batches)]
        k[0] += 1
        last[0] = job.train_step(*b)

    for _ in range(3):
        e2e_step()
    ms_e2e = time_region(e2e_step, steps, world, dev)
    loss_e2e = float(last[0].item())
    return {
        "value": round(units_per_step * steps / (ms / 1e3), 1),
        "ms_per_step": round(ms / steps, 3),
        "clocks": clocks,
        "e2e": {
            "value": round(units_per_step * steps / (ms_e2e / 1e3), 1),
            "ms_per_step": round(ms_e2e / steps, 3),
            "h2d_bytes_per_step": int(job.h2d_bytes_per_step),
            "d2h_bytes_per_step": int(job.d2h_bytes_per_step),
        },
        "gpu_launches": int(job.launches_per_step) * steps,
        "loss": {"device": loss_dev, "e2e": loss_e2e},
    }


def release(dev) -> None:
    """Frees what the previous arm held (graphs, pools) before the next one allocates."""
    gc.collect()
    torch.cuda.synchronize(dev)
    torch.cuda.empty_cache()


def job_key(suffix: str) -> str:
    env = os.environ
    base = env.get("PDJ_JOB_KEY") or f"{env.get('MASTER_ADDR', 'local')}-{env.get('MASTER_PORT', '0')}"
    return f"{base}-{suffix}"


def ratio(a, b):
    return round(a / b, 3) if (a and b) else None
