"""Rank program for the elastic Collective-mode ResNet job
(examples/resnet_elastic.yaml; the reference's deploy/elastic/resnet.yaml runs
Paddle's train_fleet_dygraph_ckpt.py with /checkpoint on a shared host path).

Membership goes through the etcd-compatible store named by
PADDLE_ELASTIC_SERVER (runtime/elastic_agent.py).  Inside one *generation* the
ranks train in lock-step.  When the desired world size changes (`pdj scale`,
which the controller turns into pods created/deleted and an np update,
controllers/paddlejob_elastic.go:41-55) or a member disappears, the ranks agree
on a stop step, leave the collective together at that step, re-form the
symmetric-memory group at the new size and continue from the same step:

  * survivors keep parameters, momentum and BN statistics in HBM - nothing is
    reloaded; only the gradient buffer (it lives in the arena) moves and the step
    graph is re-captured;
  * ranks that join receive the state from a survivor GPU-to-GPU (K6 broadcast
    through the arena), not from disk;
  * the checkpoint under /checkpoint is for whole-job restarts (every member
    gone): the fresh generation's root loads it and broadcasts.

`--cpu-dry-run` replaces the GPU job by a counter so that the membership
protocol can be exercised with real processes on a machine without GPUs.
"""
from __future__ import annotations

import argparse
import json
import os
import signal
import sys
import time


class _DryJob:
    """Stands in for ResNetJob on CPU: the 'model state' is one float that every step increments."""

    def __init__(self, step_time: float):
        self.state = 0.0
        self.step_time = step_time
        self.world = 1

    def reform(self, comm_key, rank, world):
        self.world = world

    def broadcast_state(self, etcd, prefix, gen, is_root):
        key = f"{prefix}/dry-state/{gen}"
        if is_root:
            etcd.put(key, repr(self.state))
        deadline = time.time() + 30
        while time.time() < deadline:
            kv = etcd.get(key)
            if kv:
                self.state = float(kv[0][1])
                return
            time.sleep(0.01)
        raise TimeoutError("dry-run state broadcast")

    def step(self):
        time.sleep(self.step_time)
        self.state += 1.0
        return self.state

    def state_dict(self):
        return {"state": self.state}

    def load_state_dict(self, sd):
        self.state = sd["state"]


class _Reporter:
    """Progress breadcrumbs leave through a queue and a thread: an HTTP put to the store must not sit in the
    training loop."""

    def __init__(self, agent):
        import queue
        import threading

        self.agent, self.q = agent, queue.Queue()
        self.t = threading.Thread(target=self._run, name="elastic-reporter", daemon=True)
        self.t.start()

    def _run(self):
        while True:
            item = self.q.get()
            if item is None:
                return
            self.agent.report(*item)

    def put(self, key, value):
        self.q.put((key, value))

    def flush(self, timeout: float = 5.0):
        end = time.time() + timeout
        while not self.q.empty() and time.time() < end:
            time.sleep(0.005)


class _LaggedLoss:
    """Keeps the host at most `lag` steps ahead of the device without a per-step synchronisation: step k's loss is
    copied to a pinned slot behind an event, and the host waits for the event of step k - lag.  The value read is
    therefore `lag` steps old - good enough for a progress report, and the GPU never idles waiting for the host."""

    def __init__(self, lag: int):
        import torch

        self.lag = lag
        self.host = torch.zeros(lag + 1, dtype=torch.float32).pin_memory()
        self.events = [torch.cuda.Event() for _ in range(lag + 1)]
        self.n = 0
        self.last = float("nan")
        self.done = 0            # steps of this generation known to have finished on the device
        self.t_first = None      # host time at which the first of them was seen finished
        self.t_last = None

    def rate(self) -> float:
        """Steps per second between the first and the latest step seen finished (no start-up, no queue depth)."""
        if self.done < 2 or self.t_last is None or self.t_last <= self.t_first:
            return 0.0
        return (self.done - 1) / (self.t_last - self.t_first)

    def push(self, loss_dev) -> float:
        i = self.n % (self.lag + 1)
        self.host[i:i + 1].copy_(loss_dev.reshape(1), non_blocking=True)
        self.events[i].record()
        self.n += 1
        if self.n > self.lag:
            j = (self.n - 1 - self.lag) % (self.lag + 1)
            self.events[j].synchronize()
            self.last = float(self.host[j])
            self.done += 1
            self.t_last = time.time()
            if self.t_first is None:
                self.t_first = self.t_last
        return self.last

    def drain(self) -> float:
        import torch

        torch.cuda.synchronize()
        if self.n:
            self.last = float(self.host[(self.n - 1) % (self.lag + 1)])
        return self.last


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--run-ahead", type=int, default=4, help="steps the host may queue ahead of the device")
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--image", type=int, default=224)
    ap.add_argument("--tiny", action="store_true")
    ap.add_argument("--ckpt-every", type=int, default=50)
    ap.add_argument("--report-every", type=int, default=5)
    ap.add_argument("--cpu-dry-run", action="store_true")
    ap.add_argument("--dry-step-time", type=float, default=0.02)
    ap.add_argument("--fail-at-step", type=int, default=-1, help="fault injection: this rank SIGKILLs itself")
    ap.add_argument("--fail-rank", type=int, default=-1)
    args = ap.parse_args()
    env = os.environ
    dry = args.cpu_dry_run or env.get("PDJ_CPU_DRY_RUN") == "1"

    from ..runtime.elastic_agent import ElasticAgent
    from ..utils import checkpoint as ckpt

    agent = ElasticAgent.from_env(ttl=int(env.get("PDJ_ELASTIC_TTL", "6")))
    signal.signal(signal.SIGTERM, lambda *_: setattr(agent, "leaving", True))
    root_dir = ckpt.checkpoint_root()
    job_key = env.get("PDJ_JOB_KEY", agent.job_id)
    trainer_id = int(env.get("PADDLE_TRAINER_ID", "0"))

    t_boot = time.time()
    if dry:
        job = _DryJob(args.dry_step_time)
        comm = None
    else:
        from . import require_gpu

        require_gpu("train_elastic_resnet", "--cpu-dry-run")
        import torch

        from ..parallel.symm import Communicator
        from .resnet import ResNetJob

        dev = int(env.get("LOCAL_RANK", env.get("PDJ_LOCAL_DEVICE", "0")))
        torch.cuda.set_device(dev)
        device = torch.device("cuda", dev)
        # built for world 1 first; reform() binds it to each generation's communicator
        job = ResNetJob(batch=args.batch, image=args.image, comm=None, device=device, tiny=args.tiny,
                        num_classes=16 if args.tiny else 1000, use_graph=True)
        with torch.no_grad():
            if job.s2d:
                job._x_nhwc[:, 2:-1, 2:-1, :12].copy_(torch.randn((args.batch, args.image // 2, args.image // 2, 12),
                                                                  device=device).to(torch.bfloat16))
            else:
                job.x.copy_(torch.randn(job.x.shape, device=device).to(torch.bfloat16))
            job.y.copy_(torch.randint(0, 16 if args.tiny else 1000, (args.batch,), device=device))
        # warm cuDNN / allocator once so that later captures are quick
        job.trainer.use_graph = False
        job.trainer.step()
        job.trainer.use_graph = True
        torch.cuda.synchronize()
        comm = None
    boot_seconds = round(time.time() - t_boot, 3)
    # Warm standby (runtime/supervisor.py): the supervisor starts the ranks an elastic job may still grow into
    # (limits - replicas) ahead of time with PDJ_STANDBY_GATE set.  Such a process has paid for the interpreter, the
    # CUDA context, the model and the warm-up step above; it parks here, unregistered, until its pod is created and
    # the supervisor opens the gate - a scale-up then costs a re-form, not a process boot.
    gate = env.get("PDJ_STANDBY_GATE")
    if gate:
        print(json.dumps({"event": "standby", "boot_seconds": boot_seconds}), flush=True)
        t_park = time.time()
        while not os.path.exists(gate):
            if agent.leaving:
                return 0
            time.sleep(0.01)
        print(json.dumps({"event": "activated", "parked_seconds": round(time.time() - t_park, 3)}), flush=True)
    agent.report("boot_seconds", boot_seconds)
    agent.report("standby", bool(gate))

    saver = ckpt.AsyncSaver()
    reporter = _Reporter(agent)
    m = agent.join()
    step = 0
    while True:
        t_form = time.time()
        is_root = (m.root == agent.host) or (m.root is None and m.rank == 0)
        if dry:
            job.reform(f"{job_key}-g{m.generation}", m.rank, m.world)
            if m.fresh and is_root:
                sd = ckpt.load_dp(root_dir)
                if sd is not None:
                    job.load_state_dict(sd["trainer"])
                    step = int(sd["step"])
            job.broadcast_state(agent.etcd, agent.prefix, m.generation, is_root)
            # the step travels with the members record (live hand-over) or with the checkpoint (fresh)
        else:
            import torch

            # the flat buffers live in the arena: open the new group first, move them over, then close the old one
            old_comm = comm
            comm = None
            if m.world > 1:
                comm = Communicator.create(m.rank, m.world, dev, 448 << 20, key=f"{job_key}-g{m.generation}")
            job.trainer.reform(comm)
            if old_comm is not None:
                old_comm.close()
            job.comm, job.rank, job.world = comm, m.rank, m.world
            if m.fresh and is_root:
                sd = ckpt.load_dp(root_dir, map_location=device)
                if sd is not None:
                    job.trainer.load_state_dict(sd["trainer"])
                    step = int(sd["step"])
            root_rank = m.hosts.index(m.root) if m.root in m.hosts else 0
            job.trainer.broadcast_state(root=root_rank)
        # agree on the step this generation starts from
        if is_root:
            if m.step is not None:
                step = int(m.step)
            agent.publish_resume(m.generation, step, m.world)
        else:
            step = int(agent.wait_resume(m.generation)["step"])
        agent.report(f"formed/{m.generation}", {"world": m.world, "rank": m.rank, "step": step,
                                                  "form_seconds": round(time.time() - t_form, 3), "t": time.time()})
        print(json.dumps({"event": "generation", "gen": m.generation, "rank": m.rank, "world": m.world, "step": step,
                          "fresh": m.fresh, "form_seconds": round(time.time() - t_form, 3)}), flush=True)

        # ---- train inside this generation ------------------------------------------
        stop_at = None
        t0, s0 = time.time(), step
        lagged = None if dry else _LaggedLoss(max(1, args.run_ahead))
        # the stop step is proposed `margin` steps ahead of the proposer's host step; hosts run at most --run-ahead
        # steps ahead of their devices and their devices are in lock-step, so no host can already be past it
        margin = 4 if dry else 2 * max(1, args.run_ahead) + 4
        while step < args.steps:
            if args.fail_at_step == step + 1 and args.fail_rank == trainer_id and \
                    int(env.get("PDJ_RESTART_COUNT", "0")) == 0:
                os.kill(os.getpid(), signal.SIGKILL)  # fault injection
            if dry:
                loss = job.step()
            else:
                loss = lagged.push(job.step_device())  # no per-step sync: the host trails by --run-ahead steps
            step += 1
            if step % args.report_every == 0:
                # `done` = steps whose device work is known to be complete (the host may have queued more)
                if dry:
                    rate = (step - s0) * args.batch * m.world / max(1e-9, time.time() - t0)
                else:
                    rate = lagged.rate() * args.batch * m.world
                reporter.put("progress", {"gen": m.generation, "step": step, "world": m.world, "loss": loss,
                                          "img_per_s": round(rate, 1), "t": time.time()})
            if not dry and args.ckpt_every > 0 and step % args.ckpt_every == 0:
                job.trainer.sync_state()  # collective: momentum is kept by slice owners between checkpoints
            if m.rank == 0 and args.ckpt_every > 0 and step % args.ckpt_every == 0:
                saver.save_dp(root_dir, step, m.world, {"trainer": job.state_dict() if dry else job.trainer.state_dict(),
                                                        "step": step})
            stop_at = agent.should_stop(m, step, margin)
            if stop_at is not None:
                if step > stop_at:
                    raise RuntimeError(f"rank passed the agreed stop step ({step} > {stop_at})")
                if step == stop_at:
                    break
        if step >= args.steps and stop_at is None:
            break
        if stop_at is None or step < stop_at:
            break
        # ---- re-form ----------------------------------------------------------------------
        t_stop = time.time()
        if not dry:
            job.trainer.sync_state()  # every member of the old group holds the full momentum before it re-forms
            lagged.drain()
        reporter.flush()
        agent.report(f"stopped/{m.generation}", {"step": step, "t": t_stop})
        nm = agent.next_generation(m, step)
        if nm is None:
            print(json.dumps({"event": "leave", "gen": m.generation, "rank": m.rank, "step": step}), flush=True)
            if comm is not None:
                comm.close()
            return 0
        m = nm

    # normal completion
    if not dry and lagged is not None:
        lagged.drain()
    reporter.flush()
    saver.wait()
    if not dry:
        job.trainer.sync_state()
        import torch

        torch.cuda.synchronize()
    if m.rank == 0:
        ckpt.save_dp(root_dir, step, m.world, {"trainer": job.state_dict() if dry else job.trainer.state_dict(),
                                               "step": step})
    agent.report("done", {"step": step, "t": time.time()})
    print(json.dumps({"event": "done", "rank": m.rank, "step": step}), flush=True)
    if comm is not None:
        import torch

        torch.cuda.synchronize()
        comm.barrier()
        torch.cuda.synchronize()
        comm.close()
    agent.leave()
    return 0


if __name__ == "__main__":
    sys.exit(main())
