"""The PaddleJob reconcile loop.

Level-triggered and idempotent like the reference's `Reconcile`
(controllers/paddlejob_controller.go:101-333): each call looks at the job and
the objects it owns in the store and performs the next missing mutation.  The
thirteen steps are numbered as in SURVEY.md section 3.1.  Where the reference
relies on Kubernetes components (scheduler, kubelet, Volcano) the store objects
are driven by `runtime.supervisor` / `controller.gang` instead.

Deliberate differences from the reference:
  * status is written only when it changed (the reference compares a pointer
    with a value and always writes);
  * per-pod counts are taken first and phase/mode/times derived from them in
    the same pass (the reference derives the phase from the previous counts);
  * `batch=True` performs all creations/deletions of a step in one pass, `batch=False` keeps one mutation per
    pass as the reference does.  This class defaults to False (the reference's pacing, which the scenario tests
    step through); `LocalCluster` - what `pdj manager` and every CLI path build - passes batch=True, so the
    object a user gets batches;
  * status.observedGeneration is set to metadata.generation (the reference
    zeroes it and then reads it for the volcano job-version annotation).
"""
from __future__ import annotations

import logging
from dataclasses import dataclass
from typing import Callable, Optional

from ..api.types import (RESOURCE_ANNOTATION, RESOURCE_ORDER, CleanPodPolicy, Intranet, PaddleJob, PaddleJobPhase,
                         ResourceStatus)
from . import helpers as H
from .elastic import sync_np
from .hostport import HostPortAllocator
from .store import AlreadyExists, Conflict, EventRecorder, NotFound, ObjectStore

log = logging.getLogger("paddlejob-controller")

ExecFn = Callable[[str, str, str, list], Optional[int]]  # (namespace, pod, container, command) -> rc


@dataclass
class Result:
    requeue: bool = False
    requeue_after: float = 0.0

    @property
    def wants_requeue(self) -> bool:
        return self.requeue or self.requeue_after > 0


class PaddleJobReconciler:
    """Fields mirror PaddleJobReconciler in the reference (:65-76): Scheduling,
    InitImage, HostPortMap, EtcdCli, and the pods/exec client."""

    def __init__(self, store: ObjectStore, recorder: Optional[EventRecorder] = None, scheduling: str = "",
                 init_image: str = "docker.io/library/busybox:1", host_port_range: tuple = (35000, 65000),
                 etcd=None, exec_in_pod: Optional[ExecFn] = None, batch: bool = False):
        self.store = store
        self.recorder = recorder or EventRecorder(store)
        self.scheduling = scheduling
        self.init_image = init_image
        self.ports = HostPortAllocator(*host_port_range)
        self.etcd = etcd
        self.exec_in_pod = exec_in_pod
        self.batch = batch

    # ------------------------------------------------------------------ events
    def _create(self, pdj_obj: dict, obj: dict) -> bool:
        kind, name = obj.get("kind", ""), obj["metadata"]["name"]
        try:
            self.store.create(obj)
        except AlreadyExists:
            return False
        except Exception:
            self.recorder.event(pdj_obj, "Warning", "Create", f"create failed {kind} {name}")
            raise
        self.recorder.event(pdj_obj, "Normal", "Created", f"created {kind} {name}")
        return True

    def _delete(self, pdj_obj: dict, obj: dict) -> bool:
        if (obj.get("metadata") or {}).get("deletionTimestamp"):
            return False
        kind, ns, name = obj.get("kind", ""), obj["metadata"].get("namespace", "default"), obj["metadata"]["name"]
        try:
            self.store.delete(kind, ns, name)
        except NotFound:
            return False
        except Exception:
            self.recorder.event(pdj_obj, "Warning", "Delete", f"delete failed {kind} {name}")
            raise
        self.recorder.event(pdj_obj, "Normal", "Deleted", f"deleted {kind} {name}")
        return True

    # ------------------------------------------------------------------ steps
    def _finalize(self, pdj: PaddleJob, obj: dict) -> Optional[Result]:
        """Step 2 (:460-489).  Returns a Result when the pass must end here."""
        if not pdj.metadata.deletion_timestamp:
            if H.FINALIZER_NAME not in pdj.metadata.finalizers:
                obj["metadata"].setdefault("finalizers", []).append(H.FINALIZER_NAME)
                try:
                    new = self.store.update(obj)
                except Conflict:
                    return Result(requeue_after=1.0)
                obj.clear()
                obj.update(new)
                pdj.metadata.finalizers.append(H.FINALIZER_NAME)
                pdj.metadata.resource_version = new["metadata"]["resourceVersion"]
            return None
        if H.FINALIZER_NAME in pdj.metadata.finalizers:
            if pdj.spec.intranet == Intranet.HOST.value:
                port = pdj.metadata.annotations.get(H.HOST_PORT_ANNOTATION)
                if port is not None:
                    self.ports.release(port)
            obj["metadata"]["finalizers"] = [f for f in obj["metadata"].get("finalizers", []) if f != H.FINALIZER_NAME]
            try:
                self.store.update(obj)  # removes the object: owner references cascade to the children
            except (Conflict, NotFound):
                return Result(requeue_after=1.0)
        return Result()

    def sync_current_status(self, pdj: PaddleJob, child_pods: list) -> None:
        """Step 4 (:335-381), counts first, then phase."""
        statuses: dict[str, ResourceStatus] = {}
        for pod in child_pods:
            res_type = (pod["metadata"].get("annotations") or {}).get(RESOURCE_ANNOTATION, "")
            ss = statuses.setdefault(res_type, ResourceStatus())
            phase = H.pod_phase(pod)
            if phase == "Pending":
                if H.is_coord_container_running(pod):
                    ss.starting += 1
                else:
                    ss.pending += 1
            elif phase == "Running":
                if H.is_pod_real_running(pod):
                    ss.running += 1
                else:
                    ss.starting += 1
            elif phase == "Failed":
                ss.failed += 1
            elif phase == "Succeeded":
                ss.succeeded += 1
            ss.refs.append(H.object_reference(pod))
        old = pdj.status
        pdj.status = type(old)(phase=old.phase, start_time=old.start_time, completion_time=old.completion_time)
        for res_type, st in statuses.items():
            pdj.set_status(res_type, st)
        pdj.status.phase = H.get_paddlejob_phase(pdj)
        pdj.status.mode = H.get_paddlejob_mode(pdj)
        pdj.status.start_time = H.get_paddlejob_start_time(pdj)
        pdj.status.completion_time = H.get_paddlejob_complete_time(pdj)
        pdj.status.observed_generation = pdj.metadata.generation

    def _uses_volcano(self, pdj: PaddleJob) -> bool:
        return self.scheduling == H.SCHEDULER_NAME_VOLCANO and not H.without_volcano(pdj)

    def new_pod(self, pdj: PaddleJob, res_type: str, idx: int) -> dict:
        """constructPod + the decorations of createPod (:234-275)."""
        pod = H.construct_pod(pdj, res_type, idx)
        if self.init_image != "":
            pod["spec"].setdefault("initContainers", []).append(H.gen_coordinate_init_container(self.init_image))
        if self._uses_volcano(pdj):
            pod["spec"]["schedulerName"] = H.SCHEDULER_NAME_VOLCANO
            ann = pod["metadata"]["annotations"]
            ann[H.SCHEDULING_PODGROUP_ANNOTATION] = pdj.name
            ann[H.VOLCANO_TASK_SPEC_KEY] = res_type
            ann[H.VOLCANO_JOB_NAME_KEY] = pdj.name
            ann[H.VOLCANO_JOB_VERSION] = str(pdj.status.observed_generation)
            ann[H.VOLCANO_QUEUE_NAME_KEY] = pdj.spec.scheduling_policy.queue if pdj.spec.scheduling_policy else ""
        if pdj.spec.elastic is not None and self.etcd is not None:
            pod["spec"]["containers"][0]["env"].append(
                {"name": "PADDLE_ELASTIC_SERVER", "value": ",".join(self.etcd.endpoints())})
        pod["metadata"]["ownerReferences"] = [H.owner_reference(pdj)]
        return pod

    # ------------------------------------------------------------------ main
    def reconcile(self, namespace: str, name: str) -> Result:
        # 1. fetch
        obj = self.store.try_get("PaddleJob", namespace, name)
        if obj is None:
            return Result()
        pdj = PaddleJob.from_dict(obj)
        log.debug("Reconcile %s/%s version=%s phase=%s delete=%s", namespace, name, pdj.metadata.resource_version,
                  pdj.status.phase, pdj.metadata.deletion_timestamp)
        # 2. finalizer
        r = self._finalize(pdj, obj)
        if r is not None:
            return r
        # 3. owned pods
        child_pods = self.store.list("Pod", namespace, owner=name)
        # 4. status
        old_status = pdj.status.to_dict()
        self.sync_current_status(pdj, child_pods)
        if pdj.status.to_dict() != old_status:
            obj["status"] = pdj.status.to_dict()
            try:
                new = self.store.update_status(obj)
            except Conflict:
                return Result(requeue_after=1.0)
            except NotFound:
                return Result()
            obj["metadata"]["resourceVersion"] = new["metadata"]["resourceVersion"]
        phase = pdj.status.phase
        finished = phase in (PaddleJobPhase.FAILED.value, PaddleJobPhase.COMPLETED.value)

        # 5. gang scheduling through a PodGroup
        if self._uses_volcano(pdj):
            pg = self.store.try_get("PodGroup", namespace, name)
            if finished:
                if pg is not None:
                    self._delete(obj, pg)
                    return Result(requeue=True)
            elif pg is None:
                pg = H.construct_podgroup(pdj)
                pg["metadata"]["ownerReferences"] = [H.owner_reference(pdj)]
                self._create(obj, pg)
                return Result(requeue=True)
            elif (pg.get("status") or {}).get("phase") not in ("Running", "Inqueue"):
                return Result(requeue=True)

        specs = pdj.get_specs()
        # 6. scale down: pods whose index is beyond the desired replicas
        extra = []
        for pod in child_pods:
            res_type, idx = H.extract_name_index(pod["metadata"]["name"])
            if specs.get(res_type) is not None and idx >= specs[res_type].replicas:
                extra.append(pod)
        if extra:
            for pod in (extra if self.batch else extra[:1]):
                self._delete(obj, pod)
            return Result(requeue=True)

        # 7. Service mode: one headless service per pod
        svcs: list = []
        if pdj.spec.intranet == Intranet.SERVICE.value:
            svcs = self.store.list("Service", namespace, owner=name)
            have = {s["metadata"]["name"] for s in svcs}
            made = False
            for pod in child_pods:
                if pod["metadata"]["name"] in have:
                    continue
                svc = H.construct_service_for_pod(pod)
                svc["metadata"]["ownerReferences"] = [H.owner_reference(pdj)]
                made = self._create(obj, svc) or made
                if not self.batch:
                    return Result()
            if made:
                return Result(requeue=True)
        # 8. Host mode: a block of host ports recorded as an annotation
        if pdj.spec.intranet == Intranet.HOST.value:
            if self._alloc_host_port(pdj, obj):
                return Result(requeue_after=1.0)

        # 9. elastic: publish the desired world size
        if pdj.spec.elastic is not None and self.etcd is not None:
            try:
                np = sync_np(self.etcd, pdj)
            except Exception as e:  # noqa: BLE001 - any client failure means "try again"
                log.error("sync np failed: %s", e)
                return Result(requeue=True)
            if np is not None:
                self.recorder.event(obj, "Normal", "Scaled", f"scaled replicas to {np}")
                return Result(requeue=True)

        # 10. clean-pod policy once the job is over
        pol = pdj.spec.clean_pod_policy
        clean = (phase == PaddleJobPhase.FAILED.value and pol in (CleanPodPolicy.ALWAYS.value,
                                                                  CleanPodPolicy.ON_FAILURE.value)) or \
                (phase == PaddleJobPhase.COMPLETED.value and pol in ("", CleanPodPolicy.ALWAYS.value,
                                                                     CleanPodPolicy.ON_COMPLETION.value))
        if clean:
            victims = [p for p in child_pods if not p["metadata"].get("deletionTimestamp")] or \
                      [s for s in svcs if not s["metadata"].get("deletionTimestamp")]
            if self.batch:
                victims = [p for p in child_pods] + [s for s in svcs]
            for v in (victims if self.batch else victims[:1]):
                self._delete(obj, v)
            return Result()
        if finished:
            # Nothing is created for a finished job.  (In the reference the
            # sticky phase plus surviving pods has the same effect; without this
            # guard a cleaned-up Completed job would get fresh pods.)
            return Result()

        # 11. create missing pods
        statuses = pdj.get_statuses()
        existing = {p["metadata"]["name"] for p in child_pods}
        created = False
        for res in RESOURCE_ORDER:
            if H.is_pod_created(specs[res], statuses[res]):
                continue
            for i in range(specs[res].replicas):
                if H.gen_paddle_res_name(name, res, i) in existing:
                    continue
                if self._create(obj, self.new_pod(pdj, res, i)):
                    created = True
                    if not self.batch:
                        return Result()
        if created:
            return Result()

        # 12. the global environment, once every pod has an address
        if pdj.spec.elastic is None and H.is_all_pods_ready(pdj, child_pods):
            if self.store.try_get("ConfigMap", namespace, name) is None:
                cm = H.construct_configmap(pdj, child_pods)
                if cm is None:
                    return Result(requeue=True)
                cm["metadata"]["ownerReferences"] = [H.owner_reference(pdj)]
                self._create(obj, cm)
                return Result()

        # 13. start gate: release ps, then worker, then heter
        if phase == PaddleJobPhase.STARTING.value and self.init_image != "":
            for i, res in enumerate(RESOURCE_ORDER):
                st = statuses[res]
                if st is not None and st.running < specs[res].replicas:
                    if i == 0 and st.running == 0 and not H.is_all_coord_container_running(child_pods):
                        return Result(requeue_after=1.0)
                    self._run_resource(namespace, res, child_pods)
                    return Result(requeue_after=1.0)
        return Result()

    def _run_resource(self, namespace: str, res_type: str, child_pods: list) -> None:
        for pod in child_pods:
            if (pod["metadata"].get("annotations") or {}).get(RESOURCE_ANNOTATION) != res_type:
                continue
            if H.is_coord_container_running(pod) and self.exec_in_pod is not None:
                try:
                    self.exec_in_pod(namespace, pod["metadata"]["name"], H.COORD_CONTAINER_NAME,
                                     list(H.COORD_RELEASE_CMD))
                except Exception as e:  # noqa: BLE001
                    log.error("exec in %s failed: %s", pod["metadata"]["name"], e)

    def _alloc_host_port(self, pdj: PaddleJob, obj: dict) -> bool:
        """allocHostPortForJob (:407-435).  True means "annotation written or
        allocator state rebuilt: come back in a second"."""
        port = pdj.metadata.annotations.get(H.HOST_PORT_ANNOTATION)
        if port is not None:
            if self.ports.is_allocated(port):
                return False
            if not pdj.metadata.deletion_timestamp:
                self.ports.mark(port)  # after a manager restart: rebuild from the annotation
                return True
            return False
        new_port = self.ports.allocate()
        obj["metadata"].setdefault("annotations", {})[H.HOST_PORT_ANNOTATION] = str(new_port)
        try:
            self.store.update(obj)
        except (Conflict, NotFound):
            self.ports.release(str(new_port))
        return True
