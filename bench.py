#!/usr/bin/env python
"""Headline benchmark: ResNet-50 Collective-mode training throughput (img/s),
synthetic 224x224, bf16, fixed per-GPU batch (weak scaling), on N GPUs of one
box.  BASELINE.json config 2; see BASELINE.md section 3.

  python bench.py --gpus 1 --steps 20 --warmup 5
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
      --master-port 29500 bench.py --gpus 8 --steps 20 --warmup 5

--impl product    (default) this repository's engine: fused BN/add/ReLU kernels,
                  fused NVSwitch all-reduce overlapped with backward, fused SGD,
                  whole step in a CUDA graph.
--impl ddp        the yardstick BASELINE.md names: the same network out of stock
                  torch modules under torch DDP/NCCL (+cuDNN/cuBLAS), eager.
--impl reference  the unmodified reference.  It is a Go Kubernetes operator with
                  no Python package and no GPU path, so this prints
                  {"impl": "reference", "unavailable": ...} and exits 0.
--workload wide_deep switches to the PS-mode Wide&Deep benchmark (samples/s).

Rank 0 prints exactly one JSON line on stdout.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="product", choices=["product", "ddp", "reference"])
    ap.add_argument("--workload", default="resnet50", choices=["resnet50", "wide_deep"])
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch (default 256 images / 32768 samples)")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--algo", default="auto", choices=["auto", "twoshot", "nvls"])
    ap.add_argument("--bucket-mb", type=float, default=16.0)
    ap.add_argument("--comm-blocks", type=int, default=24)
    return ap.parse_args()


def reference_unavailable() -> None:
    why = ("reference is a Go Kubernetes operator (no setup.py/pyproject, no Python or GPU code): "
           "pip install of /root/reference fails and there is no Go toolchain, Kubernetes or Paddle on the box")
    print(json.dumps({"impl": "reference", "unavailable": why}))


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None

    @staticmethod
    def gpus_in_use() -> str:
        """Only the GPUs this job runs on: idle neighbours on a bigger box sit
        at idle clocks and would drag the median down."""
        world = int(os.environ.get("WORLD_SIZE", "1"))
        visible = [v.strip() for v in os.environ.get("CUDA_VISIBLE_DEVICES", "").split(",") if v.strip()]
        ids = visible[:world] if visible else [str(i) for i in range(world)]
        return ",".join(ids)

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", self.gpus_in_use(), f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "50"],
                                      stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self) -> dict:
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        try:
            self.p.wait(5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2])); pw.append(float(c[3]))
            except ValueError:
                continue
            for nm, v in zip(names, c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        try:
            os.unlink(self.f.name)
        except OSError:
            pass
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "power_w_max": max(pw),
                "samples": len(sm), "reasons": sorted(reasons)}


def main() -> int:
    args = parse_args()
    if args.impl == "reference":
        if int(os.environ.get("RANK", "0")) == 0:
            reference_unavailable()
        return 0

    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            # convenience: re-launch ourselves under torchrun
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
                   "--master-addr", "127.0.0.1", "--master-port", os.environ.get("MASTER_PORT", "29511"),
                   os.path.abspath(__file__)] + sys.argv[1:]
            return subprocess.call(cmd)
        raise SystemExit(f"WORLD_SIZE={world} but --gpus {args.gpus}")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (use gpurun)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)

    if args.workload == "wide_deep":
        from paddle_operator_b200.workloads import bench_wide_deep

        return bench_wide_deep.run(args, rank, world, dev, ClockSampler)
    from paddle_operator_b200.workloads import bench_resnet

    return bench_resnet.run(args, rank, world, dev, ClockSampler)


if __name__ == "__main__":
    sys.exit(main())
