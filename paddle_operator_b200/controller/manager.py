"""Controller manager: watches -> work-queue -> reconcile, the role
controller-runtime plays for the reference (`SetupWithManager`,
controllers/paddlejob_controller.go:535-571: For(PaddleJob).Owns(Pod, Service,
ConfigMap[, PodGroup]); one worker; requeue / requeue-after results).

Two ways to drive it:
  * `run_until_idle()` - synchronous and deterministic, for tests and for
    one-shot CLI runs: keeps reconciling while passes still change the store;
  * `start()` / `stop()` - a background worker with real timers, for the
    long-running manager process.
"""
from __future__ import annotations

import logging
import threading
import time
from typing import Optional

from ..api.types import KIND
from ..utils import trace
from .gang import GangScheduler
from .reconciler import PaddleJobReconciler, Result
from .store import ObjectStore

log = logging.getLogger("paddlejob-manager")

OWNED_KINDS = ("Pod", "Service", "ConfigMap", "PodGroup")


class ControllerManager:
    def __init__(self, store: ObjectStore, reconciler: PaddleJobReconciler, gang: Optional[GangScheduler] = None,
                 namespace: str = ""):
        self.store, self.reconciler, self.gang = store, reconciler, gang
        self.namespace = namespace  # "" = all namespaces (main.go:71)
        self._cv = threading.Condition()
        self._due: dict[tuple, float] = {}
        self._writes = 0
        self._stop = False
        self._thread: Optional[threading.Thread] = None
        self.reconcile_count = 0
        self.reconcile_errors = 0
        self._cancel = store.watch(self._on_event)

    # -- event -> key mapping ------------------------------------------------------
    def _on_event(self, typ: str, obj: dict) -> None:
        kind = obj.get("kind")
        if kind == "Event":
            return
        meta = obj.get("metadata") or {}
        ns = meta.get("namespace", "default")
        if self.namespace and ns != self.namespace:
            return
        key = None
        if kind == KIND:
            key = (ns, meta.get("name", ""))
        elif kind in OWNED_KINDS:
            for ref in meta.get("ownerReferences") or []:
                if ref.get("controller") and ref.get("kind") == KIND:
                    key = (ns, ref.get("name", ""))
        with self._cv:
            self._writes += 1
            if key is not None:
                self._enqueue_locked(key, 0.0)
            self._cv.notify_all()

    def _enqueue_locked(self, key: tuple, delay: float) -> None:
        due = time.monotonic() + delay
        if key not in self._due or due < self._due[key]:
            self._due[key] = due

    def enqueue(self, namespace: str, name: str, delay: float = 0.0) -> None:
        with self._cv:
            self._enqueue_locked((namespace, name), delay)
            self._cv.notify_all()

    def _pop(self, ignore_delay: bool) -> Optional[tuple]:
        with self._cv:
            if not self._due:
                return None
            key = min(self._due, key=lambda k: self._due[k])
            if not ignore_delay and self._due[key] > time.monotonic():
                return None
            del self._due[key]
            return key

    def _process(self, key: tuple) -> Result:
        self.reconcile_count += 1
        try:
            with trace.span("reconcile", namespace=key[0], job=key[1]):
                res = self.reconciler.reconcile(*key)
        except Exception as e:  # noqa: BLE001 - like controller-runtime: log, back off, retry
            self.reconcile_errors += 1
            log.exception("reconcile %s/%s failed: %s", key[0], key[1], e)
            res = Result(requeue_after=1.0)
        if res.wants_requeue:
            with self._cv:
                self._enqueue_locked(key, res.requeue_after)
        return res

    # -- synchronous driver ----------------------------------------------------------
    def run_until_idle(self, max_passes: int = 10000) -> int:
        """Reconciles until the queue is empty or only holds requeues that no
        longer change anything.  Returns the number of passes."""
        passes = 0
        stale = 0
        while passes < max_passes:
            if self.gang is not None:
                self.gang.schedule_once()
            key = self._pop(ignore_delay=True)
            if key is None:
                break
            before = self._writes
            self._process(key)
            passes += 1
            if self._writes == before:
                stale += 1
                with self._cv:
                    pending = len(self._due)
                if stale > pending + 1:  # every queued key had a no-op pass: waiting on the outside world
                    break
            else:
                stale = 0
        return passes

    # -- background driver -------------------------------------------------------------
    def start(self) -> None:
        if self._thread is not None:
            return
        self._stop = False
        for obj in self.store.list(KIND, self.namespace or None):
            self.enqueue(obj["metadata"]["namespace"], obj["metadata"]["name"])
        self._thread = threading.Thread(target=self._loop, name="paddlejob-controller", daemon=True)
        self._thread.start()

    def _loop(self) -> None:
        while not self._stop:
            if self.gang is not None:
                self.gang.schedule_once()
            key = self._pop(ignore_delay=False)
            if key is None:
                with self._cv:
                    if self._stop:
                        break
                    wait = 0.2
                    if self._due:
                        wait = max(0.0, min(0.2, min(self._due.values()) - time.monotonic()))
                    self._cv.wait(wait)
                continue
            self._process(key)

    def stop(self) -> None:
        self._stop = True
        with self._cv:
            self._cv.notify_all()
        if self._thread is not None:
            self._thread.join(5)
            self._thread = None

    def close(self) -> None:
        self.stop()
        self._cancel()

    def pending(self) -> int:
        with self._cv:
            return len(self._due)
