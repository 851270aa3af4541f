"""Tracing: spans around control-plane work and kernel launches.

The reference has no tracing (one log line per reconcile).  Here
`span("reconcile", job=...)` does two things when enabled:
  * pushes an NVTX range (visible in Nsight Systems / ncu) if CUDA is up;
  * records a Chrome-trace event; `PDJ_TRACE=/path/trace.json` turns recording on and the file is
    written at exit (open it in chrome://tracing or Perfetto).
Disabled (the default) it costs one attribute lookup.
"""
from __future__ import annotations

import atexit
import contextlib
import json
import os
import threading
import time

_enabled = bool(os.environ.get("PDJ_TRACE"))
_events: list = []
_lock = threading.Lock()
_nvtx = None


def enable(path: str = "") -> None:
    global _enabled
    _enabled = True
    if path:
        os.environ["PDJ_TRACE"] = path


def _get_nvtx():
    global _nvtx
    if _nvtx is None:
        try:
            import torch

            _nvtx = torch.cuda.nvtx if torch.cuda.is_available() else False
        except Exception:  # noqa: BLE001
            _nvtx = False
    return _nvtx


@contextlib.contextmanager
def span(name: str, **args):
    if not _enabled:
        yield
        return
    nv = _get_nvtx()
    if nv:
        nv.range_push(name)
    t0 = time.perf_counter_ns()
    try:
        yield
    finally:
        t1 = time.perf_counter_ns()
        if nv:
            nv.range_pop()
        with _lock:
            _events.append({"name": name, "ph": "X", "ts": t0 / 1e3, "dur": (t1 - t0) / 1e3, "pid": os.getpid(),
                            "tid": threading.get_ident() % 100000, "args": args})


def events() -> list:
    with _lock:
        return list(_events)


def flush(path: str = "") -> str:
    path = path or os.environ.get("PDJ_TRACE", "")
    if not path or path in ("1", "true"):
        return ""
    with _lock:
        data = {"traceEvents": list(_events)}
    with open(path, "w") as f:
        json.dump(data, f)
    return path


atexit.register(flush)
