"""Rank program for Collective-mode ResNet-50 (what a worker of
examples/resnet.yaml runs; the reference's deploy/examples/resnet.yaml runs
Paddle's train_fleet.py in its place).  Reads rank/world from the PaddleJob
environment (PADDLE_TRAINER_ID / PADDLE_TRAINERS_NUM, controllers/
paddlejob_helper.go:264,319-322) or from torchrun's."""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import torch


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--image", type=int, default=224)
    ap.add_argument("--tiny", action="store_true", help="small network / images (smoke runs)")
    ap.add_argument("--lr", type=float, default=0.1)
    ap.add_argument("--log-every", type=int, default=10)
    args = ap.parse_args()
    from . import require_gpu

    require_gpu("train_resnet")
    from ..parallel.symm import Communicator
    from .resnet import ResNetJob

    env = os.environ
    rank = int(env.get("RANK", env.get("PADDLE_TRAINER_ID", "0")))
    world = int(env.get("WORLD_SIZE", env.get("PADDLE_TRAINERS_NUM", "1")))
    dev = int(env.get("LOCAL_RANK", env.get("PDJ_LOCAL_DEVICE", "0")))
    torch.cuda.set_device(dev)
    comm = Communicator.from_env(384 << 20) if world > 1 else None
    job = ResNetJob(batch=args.batch, image=args.image, comm=comm, device=torch.device("cuda", dev), lr=args.lr,
                    tiny=args.tiny, num_classes=16 if args.tiny else 1000)
    job.prepare()
    g = torch.Generator().manual_seed(rank)
    ncls = 16 if args.tiny else 1000
    x = torch.randint(0, 256, (args.batch, args.image, args.image, 3), dtype=torch.uint8, generator=g).pin_memory()
    y = torch.randint(0, ncls, (args.batch,), generator=g).pin_memory()
    t0 = time.time()
    for step in range(1, args.steps + 1):
        loss = job.train_step(x, y)
        if step % args.log_every == 0 or step == args.steps:
            torch.cuda.current_stream().synchronize()
            dt = time.time() - t0
            # fp64 sum of the fp32 master weights: ranks that hold bit-identical parameters print the same digits
            wsum = float(job.trainer.flat.master.double().sum().item())
            print(json.dumps({"rank": rank, "world": world, "step": step, "loss": round(float(loss.item()), 4),
                              "img_per_s": round(step * args.batch * world / dt, 1),
                              "param_checksum": repr(wsum)}), flush=True)
    torch.cuda.synchronize()
    if comm is not None:
        comm.barrier()
        torch.cuda.synchronize()
        comm.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
