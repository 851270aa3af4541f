"""Rank checkpoints in the reference's layout.

The reference does not read or write checkpoints itself; it fixes where they
live: every worker of the elastic example mounts the same host directory
`/tmp/checkpoint` at `/checkpoint` (deploy/elastic/resnet.yaml:25-36), parameters
are saved "by overriding the previous save" (docs/design-arch.md:58) and a
restarted pserver loads "from its save path (inferred from its index)"
(docs/design-fault-tolerant.md:51).  So:

    <root>/step.json            {"step": N, "world": W, "time": ...}   (written last)
    <root>/model.pt             data-parallel state, written by rank 0
    <root>/ps-<idx>.pt          one file per table shard, written by its owner

Each file is written to a temporary name and renamed, so a reader never sees a
torn file; `step.json` is renamed last and names a complete checkpoint.
`<root>` is resolve_mount("/checkpoint"): the supervisor exports where that
container path lives on the host.
"""
from __future__ import annotations

import json
import os
import time
from typing import Optional


def checkpoint_root(default: str = "/tmp/checkpoint") -> str:
    from ..runtime.supervisor import resolve_mount

    return resolve_mount("/checkpoint", os.environ.get("PDJ_CHECKPOINT_DIR", default))


def _atomic_save(obj, path: str) -> None:
    import torch

    tmp = f"{path}.tmp.{os.getpid()}"
    torch.save(obj, tmp)
    os.replace(tmp, path)


def save_dp(root: str, step: int, world: int, state: dict, rank: int = 0) -> None:
    """Data-parallel checkpoint: rank 0 writes model.pt then step.json."""
    if rank != 0:
        return
    os.makedirs(root, exist_ok=True)
    _atomic_save(state, os.path.join(root, "model.pt"))
    _write_step(root, step, world)


class AsyncSaver:
    """Checkpoints without stalling the training loop.  The caller's thread only enqueues a device-side
    snapshot of the tensors on its current stream (a copy at HBM speed) and an event; a background thread waits
    for the event, brings the snapshot to the host on its own stream and writes the file.  The training stream
    never waits for a device-to-host copy.  At most one save is in flight; a save requested while another runs is
    skipped (the next interval will catch up), which keeps "overwrite the previous save" semantics."""

    def __init__(self):
        import threading

        self._thread = None
        self._lock = threading.Lock()
        self._stream = None

    @staticmethod
    def _map(obj, fn):
        import torch

        if torch.is_tensor(obj):
            return fn(obj)
        if isinstance(obj, dict):
            return {k: AsyncSaver._map(v, fn) for k, v in obj.items()}
        if isinstance(obj, (list, tuple)):
            return type(obj)(AsyncSaver._map(v, fn) for v in obj)
        return obj

    @staticmethod
    def _to_host(obj):
        return AsyncSaver._map(obj, lambda t: t.detach().to("cpu", copy=True))

    def _finish(self, root, step, world, snap, event, device) -> None:
        import torch

        if event is not None:
            torch.cuda.set_device(device)
            if self._stream is None:
                self._stream = torch.cuda.Stream(device)
            with torch.cuda.stream(self._stream):
                self._stream.wait_event(event)
                host = self._map(snap, lambda t: t.to("cpu", non_blocking=False) if t.is_cuda else t)
            snap = host
        save_dp(root, step, world, snap, 0)

    def save_dp(self, root: str, step: int, world: int, state: dict, rank: int = 0) -> bool:
        import threading

        import torch

        if rank != 0:
            return False
        with self._lock:
            if self._thread is not None and self._thread.is_alive():
                return False
            event, device = None, None
            cuda = [False]

            def snap_one(t):
                if t.is_cuda:
                    cuda[0] = True
                    return t.detach().clone()  # device-side snapshot, stream-ordered after the step that produced it
                return t.detach().clone()

            snap = self._map(state, snap_one)
            if cuda[0]:
                device = torch.cuda.current_device()
                event = torch.cuda.Event()
                event.record()
            self._thread = threading.Thread(target=self._finish, args=(root, step, world, snap, event, device),
                                            daemon=True)
            self._thread.start()
            return True

    def wait(self) -> None:
        t = self._thread
        if t is not None:
            t.join()


def save_ps_shard(root: str, idx: int, state: dict) -> None:
    os.makedirs(root, exist_ok=True)
    _atomic_save(state, os.path.join(root, f"ps-{idx}.pt"))


def _write_step(root: str, step: int, world: int) -> None:
    tmp = os.path.join(root, f"step.json.tmp.{os.getpid()}")
    with open(tmp, "w") as f:
        json.dump({"step": int(step), "world": int(world), "time": time.time()}, f)
    os.replace(tmp, os.path.join(root, "step.json"))


def commit(root: str, step: int, world: int) -> None:
    """Marks a multi-file (PS) checkpoint complete."""
    os.makedirs(root, exist_ok=True)
    _write_step(root, step, world)


def latest(root: str) -> Optional[dict]:
    try:
        with open(os.path.join(root, "step.json")) as f:
            return json.load(f)
    except (OSError, ValueError):
        return None


def load_dp(root: str, map_location=None) -> Optional[dict]:
    import torch

    if latest(root) is None:
        return None
    try:
        return torch.load(os.path.join(root, "model.pt"), map_location=map_location, weights_only=False)
    except OSError:
        return None


def load_ps_shard(root: str, idx: int, map_location=None) -> Optional[dict]:
    import torch

    try:
        return torch.load(os.path.join(root, f"ps-{idx}.pt"), map_location=map_location, weights_only=False)
    except OSError:
        return None


def write_job_status(root: str, job) -> None:
    """Job status JSON with exactly the PaddleJobStatus field names (api/v1/paddlejob_types.go:168-199)."""
    os.makedirs(root, exist_ok=True)
    tmp = os.path.join(root, f"status.json.tmp.{os.getpid()}")
    with open(tmp, "w") as f:
        json.dump(job.status.to_dict(), f, indent=1)
    os.replace(tmp, os.path.join(root, "status.json"))
