"""Pure construction and state-derivation helpers (no side effects).

One function per helper of controllers/paddlejob_helper.go in the reference;
each docstring cites the lines it mirrors.  Objects are Kubernetes-shaped dicts
(what the object store holds), the PaddleJob is the typed object of api/types.
"""
from __future__ import annotations

import copy
import datetime as _dt
from typing import Iterable, Optional

from ..api.types import (API_VERSION, KIND, RESOURCE_ANNOTATION, RESOURCE_HETER, RESOURCE_NAME, RESOURCE_ORDER,
                         RESOURCE_PS, RESOURCE_TYPE, RESOURCE_WORKER, TRAINING_ROLE, Intranet, PaddleJob,
                         PaddleJobMode, PaddleJobPhase, ResourceSpec, ResourceStatus)
from .quantity import Quantity, add_resource_lists, resource_list_to_strings

# controllers/paddlejob_controller.go:48-62
HOST_PORT_NUM = 20
PADDLE_PORT = 2379
FINALIZER_NAME = "finalizers.paddlepaddle.org"
HOST_PORT_ANNOTATION = "host-port"

# controllers/paddlejob_helper.go:30-41
SCHEDULER_NAME_VOLCANO = "volcano"
SCHEDULING_PODGROUP_ANNOTATION = "scheduling.k8s.io/group-name"
COORD_CONTAINER_NAME = "coord-paddle"
COORD_CONTAINER_CPU = "10m"
COORD_CONTAINER_MEM = "10m"
COORD_CONTAINER_CMD = ["sh", "-c", "while true; do if [ -f goon ]; then exit 0; else sleep 0.1; fi; done"]
COORD_RELEASE_CMD = ["touch", "goon"]  # controllers/paddlejob_controller.go:313

# volcano batch annotation keys (volcano.sh/apis/pkg/apis/batch/v1alpha1)
VOLCANO_TASK_SPEC_KEY = "volcano.sh/task-spec"
VOLCANO_JOB_NAME_KEY = "volcano.sh/job-name"
VOLCANO_JOB_VERSION = "volcano.sh/job-version"
VOLCANO_QUEUE_NAME_KEY = "volcano.sh/queue-name"
PODGROUP_API_VERSION = "scheduling.volcano.sh/v1beta1"


def now_rfc3339() -> str:
    return _dt.datetime.now(_dt.timezone.utc).strftime("%Y-%m-%dT%H:%M:%SZ")


# --------------------------------------------------------------------------
# naming (helper.go:202-213)
def gen_paddle_res_name(name: str, res_type: str, idx: int) -> str:
    """`<job>-<resType>-<idx>`, the identifier of pod and service."""
    return f"{name}-{res_type}-{idx}"


def extract_name_index(name: str) -> tuple[str, int]:
    """Inverse of gen_paddle_res_name: the last two '-' tokens.  Returns
    ("", 0) when the last token is not an integer."""
    s = name.split("-")
    try:
        i = int(s[-1])
    except ValueError:
        return "", 0
    return (s[-2] if len(s) >= 2 else ""), i


# --------------------------------------------------------------------------
# pod status predicates (helper.go:43-90,134-173)
def pod_phase(pod: dict) -> str:
    return (pod.get("status") or {}).get("phase", "") or ""


def pod_ip(pod: dict) -> str:
    return (pod.get("status") or {}).get("podIP", "") or ""


def is_pod_created(spec: Optional[ResourceSpec], status: Optional[ResourceStatus]) -> bool:
    if spec is None:
        return True
    return status is not None and len(status.refs) == spec.replicas


def is_all_pods_created(pdj: PaddleJob) -> bool:
    specs, statuses = pdj.get_specs(), pdj.get_statuses()
    return all(is_pod_created(specs[k], statuses[k]) for k in specs)


def is_all_pods_ready(pdj: PaddleJob, child_pods: Iterable[dict]) -> bool:
    """All pods exist and every one has an IP (helper.go:43-53)."""
    if not is_all_pods_created(pdj):
        return False
    return all(pod_ip(p) != "" for p in child_pods)


def _is_failed(s): return s is not None and s.failed > 0
def _is_pending(s): return s is not None and s.pending > 0
def _is_starting(s): return s is not None and s.starting > 0
def _is_running(spec, s): return spec is None or (s is not None and spec.replicas == s.running)
def _is_completed(spec, s): return spec is None or (s is not None and spec.replicas == s.succeeded)


def is_pod_real_running(pod: dict) -> bool:
    """PodRunning with every init container Ready and every container Ready
    and in state Running (helper.go:134-151)."""
    st = pod.get("status") or {}
    if st.get("phase") != "Running":
        return False
    for c in st.get("initContainerStatuses") or []:
        if not c.get("ready"):
            return False
    for c in st.get("containerStatuses") or []:
        if not c.get("ready") or (c.get("state") or {}).get("running") is None:
            return False
    return True


def is_coord_container_running(pod: dict) -> bool:
    """PodPending with the coord-paddle init container running (helper.go:162-173)."""
    st = pod.get("status") or {}
    if st.get("phase") != "Pending":
        return False
    for c in st.get("initContainerStatuses") or []:
        if c.get("name") == COORD_CONTAINER_NAME and (c.get("state") or {}).get("running") is not None:
            return True
    return False


def is_all_coord_container_running(child_pods: Iterable[dict]) -> bool:
    return all(is_coord_container_running(p) for p in child_pods)


# --------------------------------------------------------------------------
# job state derivation (helper.go:92-132,175-199)
def get_paddlejob_phase(pdj: PaddleJob) -> str:
    """Completed/Failed are sticky; then, scanning ps, worker, heter in that
    order: any failed -> Failed, any starting -> Starting, any pending ->
    Pending (first hit wins, per resource kind as in the reference's loop);
    all running -> Running; all succeeded -> Completed; "" -> Pending; else
    unchanged.  (The reference ranges over a Go map so its cross-kind order is
    random; the fixed ps->worker->heter order here is one of its outcomes.)"""
    if pdj.status.phase == PaddleJobPhase.COMPLETED.value:
        return PaddleJobPhase.COMPLETED.value
    if pdj.status.phase == PaddleJobPhase.FAILED.value:
        return PaddleJobPhase.FAILED.value
    specs, statuses = pdj.get_specs(), pdj.get_statuses()
    for k in RESOURCE_ORDER:
        s = statuses[k]
        if _is_failed(s):
            return PaddleJobPhase.FAILED.value
        if _is_starting(s):
            return PaddleJobPhase.STARTING.value
        if _is_pending(s):
            return PaddleJobPhase.PENDING.value
    if all(_is_running(specs[k], statuses[k]) for k in RESOURCE_ORDER):
        return PaddleJobPhase.RUNNING.value
    if all(_is_completed(specs[k], statuses[k]) for k in RESOURCE_ORDER):
        return PaddleJobPhase.COMPLETED.value
    if pdj.status.phase == "":
        return PaddleJobPhase.PENDING.value
    return pdj.status.phase


def get_paddlejob_mode(pdj: PaddleJob) -> str:
    """ps set -> PS; worker.replicas > 1 -> Collective; else Single (helper.go:191-199)."""
    if pdj.spec.ps is not None:
        return PaddleJobMode.PS.value
    if pdj.spec.worker is not None and pdj.spec.worker.replicas > 1:
        return PaddleJobMode.COLLECTIVE.value
    return PaddleJobMode.SINGLE.value


def get_paddlejob_start_time(pdj: PaddleJob) -> Optional[str]:
    if not pdj.status.start_time and pdj.status.phase == PaddleJobPhase.RUNNING.value:
        return now_rfc3339()
    return pdj.status.start_time


def get_paddlejob_complete_time(pdj: PaddleJob) -> Optional[str]:
    if not pdj.status.completion_time and pdj.status.phase in (PaddleJobPhase.COMPLETED.value,
                                                               PaddleJobPhase.FAILED.value):
        return now_rfc3339()
    return pdj.status.completion_time


def object_reference(obj: dict) -> dict:
    m = obj.get("metadata") or {}
    return {k: v for k, v in {"kind": obj.get("kind", ""), "namespace": m.get("namespace", ""),
                              "name": m.get("name", ""), "uid": m.get("uid", ""),
                              "apiVersion": obj.get("apiVersion", ""),
                              "resourceVersion": str(m.get("resourceVersion", "") or "")}.items() if v}


def owner_reference(pdj: PaddleJob) -> dict:
    """ctrl.SetControllerReference equivalent."""
    return {"apiVersion": API_VERSION, "kind": KIND, "name": pdj.name, "uid": pdj.metadata.uid,
            "controller": True, "blockOwnerDeletion": True}


def controller_of(obj: dict) -> Optional[dict]:
    for ref in (obj.get("metadata") or {}).get("ownerReferences") or []:
        if ref.get("controller"):
            return ref
    return None


# --------------------------------------------------------------------------
# endpoint lists / ConfigMap (helper.go:215-279,396-403)
def endpoints_to_hosts(eps: list[str]) -> str:
    return ",".join(ep.split(":")[0] for ep in eps)


def gen_endpoints(name: str, res_type: str, num: int, port: int) -> str:
    """Present but unused in the reference (helper.go:405-412)."""
    return ",".join(f"{gen_paddle_res_name(name, res_type, i)}:{port}" for i in range(num))


def construct_configmap(pdj: PaddleJob, child_pods: Iterable[dict]) -> Optional[dict]:
    """The rendezvous object: one ConfigMap named after the job whose data
    is the global PADDLE_* environment.  Returns None (-> requeue) while any
    pod IP is not a dotted quad."""
    resources: dict[str, list[str]] = {}
    specs = pdj.get_specs()
    for res_type, spec in specs.items():
        if spec is not None:
            resources[res_type] = [""] * spec.replicas
    for pod in child_pods:
        ip = pod_ip(pod)
        if len(ip.split(".")) != 4:
            return None
        name = pod["metadata"]["name"]
        res_type, idx = extract_name_index(name)
        if res_type not in resources or idx >= len(resources[res_type]):
            continue  # a pod beyond the desired replicas; the reference would index out of range here
        host = name if pdj.spec.intranet == Intranet.SERVICE.value else ip
        resources[res_type][idx] = f"{host}:{PADDLE_PORT}"

    if pdj.spec.intranet == Intranet.HOST.value:
        paddle_port = pdj.metadata.annotations.get(HOST_PORT_ANNOTATION, "")
    else:
        paddle_port = str(PADDLE_PORT)
    data = {"TRAINER_PORTS_NUM": str(HOST_PORT_NUM), "PADDLE_PORT": paddle_port}
    if pdj.spec.ps is not None:
        data["PADDLE_PSERVERS_IP_PORT_LIST"] = ",".join(resources[RESOURCE_PS])
    if pdj.spec.worker is not None:
        data["PADDLE_TRAINER_ENDPOINTS"] = ",".join(resources[RESOURCE_WORKER])
        data["PADDLE_TRAINERS"] = endpoints_to_hosts(resources[RESOURCE_WORKER])
        data["PADDLE_TRAINERS_NUM"] = str(pdj.spec.worker.replicas)
    if pdj.spec.heter is not None:
        data["PADDLE_HETER_ENDPOINTS"] = ",".join(resources[RESOURCE_HETER])
    if pdj.spec.with_gloo is not None and pdj.spec.with_gloo > 0 and len(resources.get(RESOURCE_PS, [])) > 0:
        data["PADDLE_WITH_GLOO"] = str(pdj.spec.with_gloo)
        data["PADDLE_GLOO_RENDEZVOUS"] = "3"
        data["PADDLE_GLOO_HTTP_ENDPOINT"] = resources[RESOURCE_PS][0].replace(
            f":{PADDLE_PORT}", f":{PADDLE_PORT + HOST_PORT_NUM - 2}", 1)
    return {
        "apiVersion": "v1",
        "kind": "ConfigMap",
        "metadata": {"name": pdj.name, "namespace": pdj.namespace, "labels": {RESOURCE_NAME: pdj.name},
                     "annotations": {}},
        "data": data,
    }


# --------------------------------------------------------------------------
# pod (helper.go:281-377)
def construct_pod(pdj: PaddleJob, res_type: str, idx: int) -> dict:
    """Builds rank `idx` of kind `res_type` from the resource's pod template:
    labels/annotation, hostname/subdomain, POD_IP / PADDLE_TRAINER_ID /
    TRAINING_ROLE / PADDLE_TRAINING_ROLE env on containers[0], the elastic env
    or the envFrom ConfigMap reference, ports / hostNetwork per intranet mode,
    and the default restart policy."""
    name = gen_paddle_res_name(pdj.name, res_type, idx)
    spec = pdj.get_specs()[res_type]
    tmpl = spec.template or {}
    meta = copy.deepcopy(tmpl.get("metadata") or {})
    pspec = copy.deepcopy(tmpl.get("spec") or {})
    labels = dict(meta.get("labels") or {})
    labels[RESOURCE_NAME] = name
    labels[RESOURCE_TYPE] = res_type
    annotations = dict(meta.get("annotations") or {})
    annotations[RESOURCE_ANNOTATION] = res_type
    meta.update({"labels": labels, "annotations": annotations, "name": name, "namespace": pdj.namespace})
    pspec["hostname"] = name
    pspec["subdomain"] = name

    containers = pspec.setdefault("containers", [])
    if not containers:
        raise ValueError(f"spec.{res_type}.template.spec.containers must have at least one container")
    c0 = containers[0]
    env = list(c0.get("env") or [])
    if pdj.spec.intranet == Intranet.SERVICE.value:
        env.append({"name": "POD_IP", "value": name})
    else:
        env.append({"name": "POD_IP", "valueFrom": {"fieldRef": {"fieldPath": "status.podIP"}}})
    role = TRAINING_ROLE[res_type]
    env.append({"name": "PADDLE_TRAINER_ID", "value": str(idx)})
    env.append({"name": "TRAINING_ROLE", "value": role})
    env.append({"name": "PADDLE_TRAINING_ROLE", "value": role})
    if pdj.spec.elastic is not None:
        env.append({"name": "PADDLE_ELASTIC_JOB_ID", "value": f"{pdj.namespace}-{pdj.name}"})
        env.append({"name": "PADDLE_ELASTIC_NP", "value": str(pdj.spec.worker.replicas if pdj.spec.worker else 0)})
        env.append({"name": "PADDLE_ELASTIC_TIMEOUT", "value": "60"})
    else:
        env_from = list(c0.get("envFrom") or [])
        env_from.append({"configMapRef": {"name": pdj.name}})
        c0["envFrom"] = env_from
    c0["env"] = env

    if pdj.spec.intranet == Intranet.SERVICE.value:
        ports = list(c0.get("ports") or [])
        ports.append({"containerPort": PADDLE_PORT})
        c0["ports"] = ports
    elif pdj.spec.intranet == Intranet.HOST.value:
        pspec["hostNetwork"] = True

    if pdj.spec.elastic is not None:
        pspec["restartPolicy"] = "OnFailure"
    elif not pspec.get("restartPolicy"):
        if res_type == RESOURCE_WORKER and pdj.spec.intranet == Intranet.SERVICE.value:
            pspec["restartPolicy"] = "OnFailure"
        else:
            pspec["restartPolicy"] = "Never"
    return {"apiVersion": "v1", "kind": "Pod", "metadata": meta, "spec": pspec, "status": {}}


def gen_coordinate_init_container(image: str) -> dict:
    """The start gate: spins until a file named `goon` appears (helper.go:379-394)."""
    return {
        "name": COORD_CONTAINER_NAME,
        "image": image,
        "imagePullPolicy": "IfNotPresent",
        "command": list(COORD_CONTAINER_CMD),
        "resources": {"requests": {"cpu": COORD_CONTAINER_CPU, "memory": COORD_CONTAINER_MEM}},
    }


def construct_service_for_pod(pod: dict) -> dict:
    """Headless service per pod exposing ports 2379..2398 (helper.go:432-455)."""
    name = pod["metadata"]["name"]
    ports = [{"name": f"p-{i}", "port": PADDLE_PORT + i} for i in range(HOST_PORT_NUM)]
    return {
        "apiVersion": "v1",
        "kind": "Service",
        "metadata": {"name": name, "namespace": pod["metadata"].get("namespace", "default"), "labels": {}},
        "spec": {"ports": ports, "selector": {RESOURCE_NAME: name}, "clusterIP": "None"},
    }


# --------------------------------------------------------------------------
# volcano / gang scheduling (helper.go:457-549)
def without_volcano(pdj: PaddleJob) -> bool:
    """True when any template names a scheduler other than volcano: the job
    opts out of gang scheduling."""
    for spec in pdj.get_specs().values():
        if spec is not None and spec.scheduler_name not in ("", SCHEDULER_NAME_VOLCANO):
            return True
    return False


def get_total_replicas(pdj: PaddleJob) -> int:
    return sum(s.replicas for s in pdj.get_specs().values() if s is not None)


def get_pg_min_resource(pdj: PaddleJob) -> dict:
    """Sum over replicas x containers of requests (or limits when a container
    has no requests)."""
    total: dict = {}
    for spec in pdj.get_specs().values():
        if spec is None:
            continue
        for _ in range(spec.replicas):
            for c in (spec.template.get("spec") or {}).get("containers") or []:
                res = c.get("resources") or {}
                add_resource_lists(total, res.get("requests") if res.get("requests") is not None else res.get("limits"))
    return resource_list_to_strings(total)


def construct_podgroup(pdj: PaddleJob) -> dict:
    pg_spec: dict = {"minMember": get_total_replicas(pdj), "minResources": get_pg_min_resource(pdj)}
    sp = pdj.spec.scheduling_policy
    if sp is not None:
        if sp.min_available is not None:
            pg_spec["minMember"] = sp.min_available
        if sp.queue:
            pg_spec["queue"] = sp.queue
        if sp.priority_class:
            pg_spec["priorityClassName"] = sp.priority_class
        if sp.min_resources is not None:
            pg_spec["minResources"] = resource_list_to_strings(sp.min_resources)
    return {"apiVersion": PODGROUP_API_VERSION, "kind": "PodGroup",
            "metadata": {"name": pdj.name, "namespace": pdj.namespace}, "spec": pg_spec, "status": {}}


def gpus_requested(pod_or_template_spec: dict) -> int:
    """nvidia.com/gpu over all containers (limits, else requests)."""
    n = 0
    for c in pod_or_template_spec.get("containers") or []:
        res = c.get("resources") or {}
        q = (res.get("limits") or {}).get("nvidia.com/gpu", (res.get("requests") or {}).get("nvidia.com/gpu"))
        if q is not None:
            n += int(Quantity.parse(q))
    return n
