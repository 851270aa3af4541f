"""Interleaved A/B timing of ResNet-50 step variants in ONE process, so that clock / power-cap drift
between separate runs (about +-1 % on this box) does not hide small effects.

    python profiles/ab_resnet.py base gemm:0 gemm:150      # variant = name[:min MB for the 1x1 GEMM path]

Every round times each variant for --steps steps (CUDA events, device-resident inputs); the table
shows the median over rounds.
"""
import argparse
import json
import statistics
import sys

import torch

sys.path.insert(0, ".")


def build(spec: str, batch: int):
    from paddle_operator_b200.ops import conv as conv_ops
    from paddle_operator_b200.workloads.resnet import ResNetJob

    name, _, arg = spec.partition(":")
    if name == "base":
        job = ResNetJob(batch=batch, gemm_1x1=False)
    elif name == "gemm":
        conv_ops.MIN_OUTPUT_BYTES = int(float(arg or 0) * (1 << 20))  # read at graph-capture time (prepare)
        job = ResNetJob(batch=batch, gemm_1x1=True)
    else:
        raise SystemExit(f"unknown variant {spec}")
    job.prepare()
    return job


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("variants", nargs="+")
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--rounds", type=int, default=7)
    args = ap.parse_args()
    jobs = [(v, build(v, args.batch)) for v in args.variants]
    for _, j in jobs:
        for _ in range(5):
            j.step_device()
    torch.cuda.synchronize()
    times = {v: [] for v, _ in jobs}
    for r in range(args.rounds):
        order = jobs if r % 2 == 0 else jobs[::-1]
        for v, j in order:
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            j.step_device()
            a.record()
            for _ in range(args.steps):
                j.step_device()
            b.record()
            torch.cuda.synchronize()
            times[v].append(a.elapsed_time(b) / args.steps)
    out = {}
    for v, ts in times.items():
        out[v] = {"median_ms": round(statistics.median(ts), 3), "min_ms": round(min(ts), 3), "max_ms": round(max(ts), 3),
                  "img_per_s": round(args.batch / statistics.median(ts) * 1e3, 1)}
        print(f"{v:12s} median {out[v]['median_ms']:.3f} ms  min {out[v]['min_ms']:.3f}  max {out[v]['max_ms']:.3f}  "
              f"{out[v]['img_per_s']:.0f} img/s", flush=True)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
