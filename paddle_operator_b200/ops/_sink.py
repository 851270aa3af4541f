"""Gradient sinks: kernels write parameter gradients where the trainer wants them.

A trainer that keeps its gradients in one flat buffer (parallel/collective.py, FlatParams) tags every parameter
with `_grad_sink`, a view of that buffer with the parameter's shape and strides.  An op whose backward produces a
parameter gradient with one of this repository's kernels asks `sink_of(param)`; when there is one, the kernel
writes straight into it and the op returns `adopt(sink)` - a fresh alias that autograd installs as `.grad` without
a copy.  The trainer's hook sees `.grad` already sitting at the sink's address and skips its copy, so no gradient
is written twice and no `_foreach_copy_` runs (PDJ_GRAD_SINK=0 restores the copy path; anything that does not
use a sink - library backward passes, accumulated gradients of shared weights - still takes it).

Reference parity: none (the reference has no compute code).
"""
from __future__ import annotations

import os
from typing import Optional

import torch

ENABLED = os.environ.get("PDJ_GRAD_SINK", "1") == "1"


def sink_of(p: Optional[torch.Tensor], dtype=torch.bfloat16) -> Optional[torch.Tensor]:
    if p is None or not ENABLED:
        return None
    s = getattr(p, "_grad_sink", None)
    if s is None or s.dtype != dtype or s.shape != p.shape or not s.is_cuda:
        return None
    return s


def adopt(sink: torch.Tensor) -> torch.Tensor:
    """A new tensor object over the sink's memory: autograd may take it as the gradient without copying."""
    return sink.detach()


def is_sunk(grad: Optional[torch.Tensor], sink: torch.Tensor) -> bool:
    return grad is not None and grad.data_ptr() == sink.data_ptr() and grad.dtype == sink.dtype


# ---- weight gradients on a side stream -----------------------------------------------------------------------------
# Inside a trainer step (CollectiveTrainer sets `overlap_scope`), an op may compute its PARAMETER gradients on a side
# stream so that the main stream can go on with the input-gradient chain and whatever follows it (in the Wide&Deep
# step: the embedding push / apply, NVLink- and DRAM-bound, while the tower's weight gradients use the tensor cores).
# The op records an event after its side-stream work and registers it here; the trainer makes every consumer of
# the gradients (bucket collective, optimizer, end of step) wait for the registered events, and keeps the tensors the
# side stream reads alive until the step is over (no allocator reuse while they are in flight).
WGRAD_OVERLAP = os.environ.get("PDJ_WGRAD_OVERLAP", "1") in ("1", "2")
WGRAD_OVERLAP_FORCE = os.environ.get("PDJ_WGRAD_OVERLAP", "1") == "2"
_scope = {"on": False, "events": [], "keep": [], "streams": {}}


def overlap_enabled() -> bool:
    return WGRAD_OVERLAP and _scope["on"]


def side_stream(device) -> "torch.cuda.Stream":
    key = (device.type, device.index)
    st = _scope["streams"].get(key)
    if st is None:
        st = _scope["streams"][key] = torch.cuda.Stream(device)
    return st


def register_side_work(event, *tensors) -> None:
    _scope["events"].append(event)
    _scope["keep"].extend(tensors)


def begin_step(overlap: bool = True) -> None:
    _scope["on"] = overlap
    _scope["events"].clear()
    _scope["keep"].clear()


def wait_side_work(stream) -> None:
    """Make `stream` wait for every side-stream gradient computation registered so far in this step."""
    for e in _scope["events"]:
        stream.wait_event(e)


def end_step() -> None:
    _scope["on"] = False
    _scope["events"].clear()
    _scope["keep"].clear()
