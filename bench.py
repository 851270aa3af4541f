#!/usr/bin/env python
"""Headline benchmark: ResNet-50 Collective-mode training throughput (img/s),
synthetic 224x224, bf16, fixed per-GPU batch (weak scaling), on N GPUs of one
box.  BASELINE.json config 2; see BASELINE.md section 3.

  python bench.py --gpus 1 --steps 20 --warmup 5
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
      --master-port 29500 bench.py --gpus 8 --steps 20 --warmup 5

--impl product    (default) this repository's engine: tcgen05 convolutions, fused
                  BN/add/ReLU kernels, fused NVSwitch all-reduce overlapped with
                  backward, fused SGD, whole step in a CUDA graph.
                  With the default --arms all the same invocation then also runs,
                  with the same steps / warm-up, and attaches to the one JSON line:
                    baselines.ddp_eager   torch DDP/NCCL, stock modules, eager (the
                                          yardstick BASELINE.md names; vs_baseline =
                                          value / its value)
                    baselines.ddp_graph   the same stock step captured in a CUDA graph
                    workloads.wide_deep   the PS-mode Wide&Deep product, with its own
                                          torch + NCCL all-to-all baselines (eager
                                          without host syncs, and graph-captured)
                    allreduce_check       (N >= 2) every all-reduce algorithm incl. NVLS
                                          against dist.all_reduce on 64 MiB of random
                                          bf16; a mismatch makes the run exit 3.
                  --arms primary runs the product arm only.
--impl ddp | ddp_graph   one baseline arm alone.
--impl reference  the unmodified reference.  It is a Go Kubernetes operator with
                  no Python package and no GPU path, so this prints
                  {"impl": "reference", "unavailable": ...} and exits 0.
--workload wide_deep  the PS-mode Wide&Deep benchmark (samples/s) as the primary
                  line (--impl product | nccl | nccl_graph).

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
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="product", choices=["product", "ddp", "ddp_graph", "nccl", "nccl_graph", "reference"])
    ap.add_argument("--arms", default="all", choices=["all", "primary"],
                    help="all: baselines, Wide&Deep and the all-reduce check ride along with the product arm")
    ap.add_argument("--wd-batch", type=int, default=0, help="per-GPU Wide&Deep batch (default 32768)")
    ap.add_argument("--budget-s", type=float, default=900.0,
                    help="wall-clock guard: print what has been measured so far and exit if the run takes longer")
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

    from paddle_operator_b200.workloads import bench_resnet, bench_wide_deep
    from paddle_operator_b200.workloads.bench_common import ratio

    wd_primary = args.workload == "wide_deep"
    if wd_primary and args.batch and not args.wd_batch:
        args.wd_batch = args.batch
    mod = bench_wide_deep if wd_primary else bench_resnet
    impl = args.impl
    if wd_primary and impl in ("ddp", "ddp_graph"):
        impl = {"ddp": "nccl", "ddp_graph": "nccl_graph"}[impl]
    if not wd_primary and impl in ("nccl", "nccl_graph"):
        impl = {"nccl": "ddp", "nccl_graph": "ddp_graph"}[impl]

    out: dict = {}
    printed = threading.Event()

    def emit(note=None) -> None:
        if printed.is_set():
            return
        printed.set()
        if note:
            out["truncated"] = note
        if rank == 0:
            print(json.dumps(out), flush=True)

    def guard() -> None:  # a hung baseline arm must not cost the primary measurement
        emit(f"wall-clock budget of {args.budget_s:.0f} s exceeded; arms not listed were not measured")
        os._exit(0)

    timer = threading.Timer(args.budget_s, guard)
    timer.daemon = True
    timer.start()

    rec = mod.measure(args, impl, rank, world, dev, ClockSampler)
    out.update({
        "metric": rec["metric"], "value": rec["value"], "unit": rec["unit"], "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": rec["ms_per_step"], "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": rec.get("dtype", "bf16"),
        "data": rec.get("data", "synthetic (random uint8 224x224 images, random labels; random-init weights)"),
        "impl": rec["impl"], "config": rec["config"], "clocks": rec["clocks"],
        "e2e": dict(rec["e2e"], unit=rec["unit"]), "gpu_launches": rec["gpu_launches"], "loss": rec["loss"],
    })
    rc = 0
    if args.arms == "all" and impl == "product" and not wd_primary:
        def arm(module, name):
            """One ride-along arm; a failure is recorded, not fatal (it fails on every rank alike)."""
            try:
                r = module.measure(args, name, rank, world, dev, ClockSampler)
                return {k: r[k] for k in ("value", "ms_per_step", "e2e", "clocks", "impl_name", "loss", "gpu_launches")
                        if k in r} | ({"invalid": r["invalid"]} if "invalid" in r else {})
            except Exception as e:  # noqa: BLE001
                return {"unavailable": f"{type(e).__name__}: {e}"[:300]}

        def versus(prod, base):
            if "value" in base and "invalid" not in base:
                base["ratio"] = ratio(prod["value"], base["value"])
                base["e2e_ratio"] = ratio(prod["e2e"]["value"], base["e2e"]["value"])

        if world > 1:
            from paddle_operator_b200.workloads.bench_common import job_key
            from paddle_operator_b200.workloads.comm_check import allreduce_check

            chk = allreduce_check(dev, rank, world, key=job_key("archeck"))
            out["allreduce_check"] = chk
            if not chk["ok"]:
                rc = 3
        # the yardstick first, then the second workload, then the graph-captured baselines
        base = {"ddp_eager": arm(bench_resnet, "ddp")}
        versus(rec, base["ddp_eager"])
        out["baselines"] = base
        if "ratio" in base["ddp_eager"]:
            out["vs_baseline"] = base["ddp_eager"]["ratio"]
            out["vs_baseline_is"] = "value / baselines.ddp_eager.value, both measured in this invocation"
        wd = arm(bench_wide_deep, "product")
        wd.update({"metric": "wide_deep_train_samples_per_sec", "unit": "samples/s", "per_gpu_batch": args.wd_batch or 32768,
                   "vs_baseline": None})
        out["workloads"] = {"wide_deep": wd}
        if "value" in wd:
            wd["baselines"] = {"nccl_eager": arm(bench_wide_deep, "nccl")}
            versus(wd, wd["baselines"]["nccl_eager"])
            wd["vs_baseline"] = wd["baselines"]["nccl_eager"].get("ratio")
        base["ddp_graph"] = arm(bench_resnet, "ddp_graph")
        versus(rec, base["ddp_graph"])
        if "value" in wd:
            wd["baselines"]["nccl_graph"] = arm(bench_wide_deep, "nccl_graph")
            versus(wd, wd["baselines"]["nccl_graph"])
    timer.cancel()
    emit()
    if world > 1:
        torch.cuda.synchronize(dev)
        dist.barrier(device_ids=[dev.index])
        dist.destroy_process_group()
    return rc


if __name__ == "__main__":
    sys.exit(main())
