"""Rank program for PS-mode Wide&Deep (examples/wide_and_deep*.yaml; the
reference's deploy/examples/wide_and_deep.yaml runs Paddle's demo image).

TRAINING_ROLE (controllers/paddlejob_helper.go:323-330) selects the role:

TRAINER  trains; the table shards live in the trainers' GPU arenas and are pulled /
         pushed by the fused kernels (parallel/ps.py).
PSERVER  on one box the server is memory plus the owner-side half of those kernels,
         so the PSERVER process holds no table: it keeps the role alive for start
         ordering / status / clean-up and exits 0 once every trainer reports done
         (a marker file per trainer under the job's rendezvous directory).
HETER    same as PSERVER (no dedicated kernels are owed for heter workers).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import tempfile
import time


def _done_dir() -> str:
    d = os.path.join(tempfile.gettempdir(), f"pdj-{os.environ.get('PDJ_JOB_KEY', 'local')}", "done")
    os.makedirs(d, exist_ok=True)
    return d


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--batch", type=int, default=8192)
    ap.add_argument("--rows", type=int, default=1 << 20)
    ap.add_argument("--emb-dim", type=int, default=16)
    ap.add_argument("--async-sgd", action="store_true")
    ap.add_argument("--log-every", type=int, default=10)
    ap.add_argument("--ckpt-every", type=int, default=0, help="also checkpoint every N steps (always at the end)")
    ap.add_argument("--cpu-dry-run", action="store_true", help="no GPU: only exercise the role plumbing")
    args = ap.parse_args()
    env = os.environ
    role = env.get("TRAINING_ROLE", "TRAINER")
    rank = int(env.get("PADDLE_TRAINER_ID", env.get("RANK", "0")))
    world = int(env.get("PADDLE_TRAINERS_NUM", env.get("WORLD_SIZE", "1")))
    dry = args.cpu_dry_run or env.get("PDJ_CPU_DRY_RUN") == "1"
    if role != "TRAINER":
        deadline = time.time() + float(env.get("PDJ_PS_TIMEOUT", "3600"))
        while time.time() < deadline:
            if len([f for f in os.listdir(_done_dir()) if f.startswith("trainer-")]) >= world:
                print(json.dumps({"role": role, "rank": rank, "status": "trainers finished"}), flush=True)
                return 0
            time.sleep(0.1)
        return 1
    if not dry:
        from . import require_gpu

        require_gpu("train_wide_deep", "--cpu-dry-run")
        import torch

        from .wide_deep import WideDeepJob

        torch.cuda.set_device(int(env.get("LOCAL_RANK", env.get("PDJ_LOCAL_DEVICE", "0"))))
        job = WideDeepJob.from_env(batch=args.batch, rows=args.rows, emb_dim=args.emb_dim, async_sgd=args.async_sgd)
        job.prepare()
        from ..utils import checkpoint as ckpt

        root = ckpt.checkpoint_root()
        # checkpoints only where the job asked for a place to put them: a volume mounted at /checkpoint (the
        # reference's examples mount one) or PDJ_CHECKPOINT_DIR
        do_ckpt = root != "/tmp/checkpoint" or bool(env.get("PDJ_CHECKPOINT_DIR")) or args.ckpt_every > 0
        start = (job.load_checkpoint(root) or 0) if do_ckpt else 0
        if start:
            print(json.dumps({"rank": rank, "resumed_from_step": start, "shards": job.emb.shards_of(job.rank)}), flush=True)
        batches = [job.synthetic_batch(s) for s in range(4)]
        t0 = time.time()
        for step in range(start + 1, args.steps + 1):
            loss = job.train_step(*batches[step % 4])
            if step % args.log_every == 0 or step == args.steps:
                torch.cuda.current_stream().synchronize()
                print(json.dumps({"rank": rank, "world": world, "step": step, "loss": round(float(loss.item()), 4),
                                  "samples_per_s": round(step * args.batch * world / (time.time() - t0), 1)}), flush=True)
            if do_ckpt and args.ckpt_every > 0 and step % args.ckpt_every == 0 and step < args.steps:
                job.save_checkpoint(root, step)
        torch.cuda.synchronize()
        if do_ckpt:
            job.save_checkpoint(root, args.steps)
        if world > 1:
            job.comm.barrier()
            torch.cuda.synchronize()
    else:
        print(json.dumps({"role": role, "rank": rank, "dry_run": True,
                          "pservers": env.get("PADDLE_PSERVERS_IP_PORT_LIST", "")}), flush=True)
    open(os.path.join(_done_dir(), f"trainer-{rank}"), "w").close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
