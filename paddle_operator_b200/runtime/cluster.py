"""LocalCluster: one object that wires the whole control plane together -
object store, event recorder, reconciler, controller manager, gang scheduler,
rank supervisor and (optionally) the etcd-compatible store - the single-box
equivalent of "a Kubernetes cluster with the operator installed"
(deploy/v1/operator.yaml in the reference).  The CLI and the tests use it.
"""
from __future__ import annotations

import time
from typing import Iterable, Optional, Union

from ..api import loader
from ..api.types import KIND, PaddleJob
from ..controller.gang import GangScheduler
from ..controller.manager import ControllerManager
from ..controller.reconciler import PaddleJobReconciler
from ..controller.store import EventRecorder, NotFound, ObjectStore
from .supervisor import Supervisor


class LocalCluster:
    def __init__(self, scheduling: str = "", init_image: str = "docker.io/library/busybox:1",
                 port_range: str = "35000,65000", etcd=None, namespace: str = "", gpus: Optional[list] = None,
                 work_root: str = "/tmp/pdjob", batch: bool = True, isolate_gpus: bool = True,
                 extra_env: Optional[dict] = None, supervisor: bool = True, gang_capacity: Optional[dict] = None,
                 priorities: Optional[dict] = None, state_dir: Optional[str] = None):
        self.store = ObjectStore()
        # durable PaddleJobs (spec + status), reloaded when a manager starts on the same directory
        self.persister = None
        if state_dir:
            from ..controller.persist import StorePersister

            self.persister = StorePersister(self.store, state_dir)
            self.restored = self.persister.load()
            self.persister.start()
        self.recorder = EventRecorder(self.store)
        self.supervisor = Supervisor(self.store, work_root=work_root, gpus=gpus, isolate_gpus=isolate_gpus,
                                     extra_env=extra_env) if supervisor else None
        start, end = (int(x) for x in port_range.split(","))
        self.reconciler = PaddleJobReconciler(self.store, self.recorder, scheduling=scheduling, init_image=init_image,
                                              host_port_range=(start, end), etcd=etcd, batch=batch,
                                              exec_in_pod=self.supervisor.exec_in_pod if self.supervisor else None)
        if self.supervisor is not None:
            self.supervisor.pod_factory = self.reconciler.new_pod  # warm standbys are the reconciler's own pods
        self.gang = None
        if scheduling == "volcano":
            cap = gang_capacity
            if cap is None and self.supervisor is not None:
                cap = {"nvidia.com/gpu": len(self.supervisor.gpu_pool)}
            self.gang = GangScheduler(self.store, cap, priorities)
        self.manager = ControllerManager(self.store, self.reconciler, self.gang, namespace=namespace)
        self.etcd = etcd
        self._started = False

    # -- lifecycle ----------------------------------------------------------------
    def start(self) -> "LocalCluster":
        if not self._started:
            if self.supervisor:
                self.supervisor.start()
            self.manager.start()
            self._started = True
        return self

    def stop(self) -> None:
        self.manager.close()
        if self.supervisor:
            self.supervisor.stop()
        if self.persister:
            self.persister.stop()  # the jobs stay on disk for the next manager
        self._started = False

    def __enter__(self):
        return self.start()

    def __exit__(self, *a):
        self.stop()

    # -- kubectl-like verbs ---------------------------------------------------------
    def apply(self, src: Union[str, dict, PaddleJob]) -> list[dict]:
        """Create-or-update every document in `src` (path, YAML text, dict or PaddleJob)."""
        docs: Iterable
        if isinstance(src, PaddleJob):
            docs = [src]
        elif isinstance(src, dict):
            docs = [loader.paddlejob_from_doc(src) if loader.is_paddlejob(src) else src]
        else:
            docs = loader.load_documents(src)
        out = []
        for d in docs:
            obj = d.to_dict() if isinstance(d, PaddleJob) else d
            if not isinstance(obj, dict) or "kind" not in obj:
                continue
            if isinstance(d, PaddleJob):
                obj.pop("status", None)
            meta = obj.setdefault("metadata", {})
            cur = self.store.try_get(obj["kind"], meta.get("namespace") or "default", meta.get("name", ""))
            if cur is None:
                out.append(self.store.create(obj))
            else:
                merged = dict(cur)
                merged["spec"] = obj.get("spec", cur.get("spec"))
                for k in ("labels", "annotations"):
                    if meta.get(k):
                        merged["metadata"].setdefault(k, {}).update(meta[k])
                out.append(self.store.update(merged))
        return out

    def get(self, name: str, namespace: str = "default") -> PaddleJob:
        return PaddleJob.from_dict(self.store.get(KIND, namespace, name))

    def list_jobs(self, namespace: Optional[str] = None) -> list[PaddleJob]:
        return [PaddleJob.from_dict(o) for o in self.store.list(KIND, namespace)]

    def delete(self, name: str, namespace: str = "default") -> None:
        self.store.delete(KIND, namespace, name)

    def scale(self, name: str, replicas: int, resource: str = "worker", namespace: str = "default") -> None:
        for _ in range(10):
            obj = self.store.get(KIND, namespace, name)
            if obj["spec"].get(resource) is None:
                raise ValueError(f"job {name} has no {resource} resource")
            obj["spec"][resource]["replicas"] = int(replicas)
            try:
                self.store.update(obj)
                return
            except Exception:  # Conflict: retry on the fresh object
                time.sleep(0.01)
        raise RuntimeError("could not update the job (conflicts)")

    def wait(self, name: str, phases=("Completed", "Failed"), timeout: float = 60.0, namespace: str = "default",
             poll: float = 0.02) -> str:
        deadline = time.monotonic() + timeout
        phase = ""
        while time.monotonic() < deadline:
            try:
                phase = self.get(name, namespace).status.phase
            except NotFound:
                phase = "<deleted>"
            if phase in phases:
                return phase
            time.sleep(poll)
        raise TimeoutError(f"job {name} is {phase!r}, wanted one of {phases} within {timeout}s")

    def wait_deleted(self, name: str, timeout: float = 30.0, namespace: str = "default") -> None:
        deadline = time.monotonic() + timeout
        while time.monotonic() < deadline:
            if self.store.try_get(KIND, namespace, name) is None:
                return
            time.sleep(0.02)
        raise TimeoutError(f"job {name} still exists")

    def events(self, name: str, namespace: str = "default") -> list[dict]:
        return self.recorder.events_for(KIND, namespace, name)

    def pods(self, name: Optional[str] = None, namespace: str = "default") -> list[dict]:
        return self.store.list("Pod", namespace, owner=name) if name else self.store.list("Pod", namespace)

    def logs(self, pod: str, namespace: str = "default") -> str:
        return self.supervisor.logs(namespace, pod) if self.supervisor else ""
