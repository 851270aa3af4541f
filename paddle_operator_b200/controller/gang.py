"""Gang admission: the part of Volcano the reference depends on.

With `--scheduling=volcano` the reference creates one PodGroup per job
(minMember, minResources, queue, priorityClassName;
controllers/paddlejob_helper.go:478-507) and only creates pods while the
PodGroup's phase is Running or Inqueue (controllers/paddlejob_controller.go:
133-157).  On one box the "cluster" is the machine's GPUs (plus whatever other
resources are declared), so this scheduler admits PodGroups all-or-nothing
against that pool: ordered by priority class then creation time, per queue; a
group that fits moves Pending -> Inqueue (resources reserved) -> Running (once
its pods exist); deleting the PodGroup releases the reservation.
"""
from __future__ import annotations

import threading
from typing import Optional

from .quantity import Quantity
from .store import NotFound, ObjectStore

DEFAULT_QUEUE = "default"


class GangScheduler:
    def __init__(self, store: ObjectStore, capacity: Optional[dict] = None, priorities: Optional[dict] = None,
                 queues: Optional[dict] = None):
        """capacity: resource name -> quantity (e.g. {"nvidia.com/gpu": 8}); resources not
        listed are unlimited.  priorities: priorityClassName -> int (higher first).
        queues: queue name -> {"weight": int, "capability": {res: qty}} (optional caps)."""
        self.store = store
        self.capacity = {k: Quantity.parse(v) for k, v in (capacity or {"nvidia.com/gpu": 8}).items()}
        self.priorities = priorities or {}
        self.queues = queues or {DEFAULT_QUEUE: {}}
        self._reserved: dict[tuple, dict] = {}  # (ns, name) -> resources
        self._lock = threading.Lock()

    def _fits(self, need: dict, queue: str) -> bool:
        used: dict = {}
        used_q: dict = {}
        for (ns, name), (res, q) in self._reserved.items():
            for k, v in res.items():
                used[k] = used.get(k, Quantity.parse(0)) + v
                if q == queue:
                    used_q[k] = used_q.get(k, Quantity.parse(0)) + v
        for k, v in need.items():
            if k in self.capacity and not (used.get(k, Quantity.parse(0)) + v <= self.capacity[k]):
                return False
            cap = (self.queues.get(queue) or {}).get("capability") or {}
            if k in cap and not (used_q.get(k, Quantity.parse(0)) + v <= Quantity.parse(cap[k])):
                return False
        return True

    def schedule_once(self) -> list[str]:
        """One scheduling cycle.  Returns the names of groups whose phase changed."""
        changed = []
        with self._lock:
            groups = self.store.list("PodGroup")
            live = {(g["metadata"]["namespace"], g["metadata"]["name"]) for g in groups}
            for key in list(self._reserved):
                if key not in live:
                    del self._reserved[key]
            groups.sort(key=lambda g: (-self.priorities.get((g.get("spec") or {}).get("priorityClassName", ""), 0),
                                       g["metadata"].get("creationTimestamp", ""),
                                       int(g["metadata"].get("resourceVersion", "0") or 0)))
            for g in groups:
                ns, name = g["metadata"]["namespace"], g["metadata"]["name"]
                spec = g.get("spec") or {}
                phase = (g.get("status") or {}).get("phase", "") or "Pending"
                queue = spec.get("queue") or DEFAULT_QUEUE
                new_phase = phase
                if phase in ("", "Pending"):
                    if queue not in self.queues:
                        new_phase = "Pending"  # unknown queue: never admitted (Volcano rejects such groups)
                    else:
                        need = {k: Quantity.parse(v) for k, v in (spec.get("minResources") or {}).items()}
                        if self._fits(need, queue):
                            self._reserved[(ns, name)] = (need, queue)
                            new_phase = "Inqueue"
                elif phase == "Inqueue":
                    if (ns, name) not in self._reserved:  # e.g. after a manager restart
                        need = {k: Quantity.parse(v) for k, v in (spec.get("minResources") or {}).items()}
                        self._reserved[(ns, name)] = (need, queue)
                    pods = [p for p in self.store.list("Pod", ns)
                            if (p["metadata"].get("annotations") or {}).get("scheduling.k8s.io/group-name") == name]
                    if len(pods) >= int(spec.get("minMember", 0) or 0) and pods:
                        new_phase = "Running"
                elif phase == "Running" and (ns, name) not in self._reserved:
                    need = {k: Quantity.parse(v) for k, v in (spec.get("minResources") or {}).items()}
                    self._reserved[(ns, name)] = (need, queue)
                if new_phase != phase or not (g.get("status") or {}).get("phase"):
                    try:
                        self.store.patch_status("PodGroup", ns, name, lambda st, p=new_phase: st.update({"phase": p}))
                        changed.append(name)
                    except NotFound:
                        self._reserved.pop((ns, name), None)
        return changed

    def reserved(self) -> dict:
        with self._lock:
            return {k: {r: str(q) for r, q in v[0].items()} for k, v in self._reserved.items()}
