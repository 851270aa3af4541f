"""Durable PaddleJobs: the part of kube-apiserver + etcd the reference gets for free.

In the reference a PaddleJob and its status live in the API server (the status is written through the
status subresource, controllers/paddlejob_controller.go:125), so the operator itself is stateless: after a
restart it lists the jobs again and carries on.  `StorePersister` gives the in-memory ObjectStore the same
property on one box: every change to a persisted kind is written to

    <state_dir>/<kind plural>/<namespace>/<name>.json       (atomic: temp file + rename)

with exactly the object the API shows - `spec`, and `status` with the PaddleJobStatus field names
(api/v1/paddlejob_types.go:168-216) - and `load()` puts them back (ObjectStore.restore) before the
controllers start.  Children (pods, services, config maps) are deliberately NOT persisted: their processes
die with the supervisor, and the reconciler re-creates them from the job, just as it would after a node
loss; Completed / Failed jobs stay finished because those phases are sticky (controllers/paddlejob_helper.go:94-99).
"""
from __future__ import annotations

import json
import logging
import os
import tempfile
import threading
from typing import Callable, Iterable, Optional

from ..api.types import KIND
from .store import AlreadyExists, ObjectStore

log = logging.getLogger("pdjob-persist")

PLURALS = {KIND: "paddlejobs"}


class StorePersister:
    def __init__(self, store: ObjectStore, state_dir: str, kinds: Iterable[str] = (KIND,)):
        self.store, self.state_dir = store, state_dir
        self.kinds = tuple(kinds)
        self._lock = threading.Lock()
        self._cancel: Optional[Callable[[], None]] = None
        os.makedirs(state_dir, exist_ok=True)

    # -- paths ----------------------------------------------------------------------
    def _dir(self, kind: str, namespace: str) -> str:
        return os.path.join(self.state_dir, PLURALS.get(kind, kind.lower() + "s"), namespace)

    def path_of(self, kind: str, namespace: str, name: str) -> str:
        return os.path.join(self._dir(kind, namespace), name + ".json")

    # -- store -> disk ------------------------------------------------------------------
    def _on_event(self, typ: str, obj: dict) -> None:
        kind = obj.get("kind")
        if kind not in self.kinds:
            return
        meta = obj.get("metadata") or {}
        ns, name = meta.get("namespace") or "default", meta.get("name", "")
        path = self.path_of(kind, ns, name)
        with self._lock:
            try:
                if typ == "DELETED":
                    if os.path.exists(path):
                        os.unlink(path)
                    return
                os.makedirs(os.path.dirname(path), exist_ok=True)
                fd, tmp = tempfile.mkstemp(prefix="." + name + ".", dir=os.path.dirname(path))
                with os.fdopen(fd, "w") as f:
                    json.dump(obj, f, indent=1, sort_keys=True)
                os.replace(tmp, path)
            except OSError as e:  # a full or read-only disk must not take the controller down
                log.error("could not persist %s %s/%s: %s", kind, ns, name, e)

    def start(self) -> "StorePersister":
        if self._cancel is None:
            self._cancel = self.store.watch(self._on_event)
        return self

    def stop(self) -> None:
        if self._cancel is not None:
            self._cancel()
            self._cancel = None

    # -- disk -> store ------------------------------------------------------------------
    def load(self) -> int:
        """Restores every persisted object that the store does not hold yet; returns how many."""
        n = 0
        for kind in self.kinds:
            root = os.path.join(self.state_dir, PLURALS.get(kind, kind.lower() + "s"))
            if not os.path.isdir(root):
                continue
            for ns in sorted(os.listdir(root)):
                d = os.path.join(root, ns)
                if not os.path.isdir(d):
                    continue
                for fn in sorted(os.listdir(d)):
                    if not fn.endswith(".json") or fn.startswith("."):
                        continue
                    try:
                        with open(os.path.join(d, fn)) as f:
                            obj = json.load(f)
                        if obj.get("kind") != kind:
                            continue
                        self.store.restore(obj)
                        n += 1
                    except AlreadyExists:
                        pass
                    except (OSError, ValueError) as e:
                        log.error("skipping unreadable state file %s: %s", os.path.join(d, fn), e)
        return n
