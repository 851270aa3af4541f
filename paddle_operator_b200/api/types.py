"""PaddleJob API types (batch.paddlepaddle.org/v1).

Python dataclasses with the field names, JSON keys, enum values and omit-empty
behaviour of the reference's Go types (api/v1/paddlejob_types.go:26-268,
api/v1/groupversion_info.go:27-33).  PodTemplateSpec is carried through as a
plain dict (the reference inlines the whole core/v1 schema, ~9,000 lines of
generated YAML per copy); typed accessors cover the parts the controller reads.
"""
from __future__ import annotations

import copy
import enum
from dataclasses import dataclass, field
from typing import Any, Optional

GROUP = "batch.paddlepaddle.org"
VERSION = "v1"
API_VERSION = f"{GROUP}/{VERSION}"
KIND = "PaddleJob"  # paddlejob_types.go:26
SHORT_NAME = "pdj"  # paddlejob_types.go:219

# label / annotation keys, paddlejob_types.go:31-35
RESOURCE_NAME = "paddle-res-name"
RESOURCE_TYPE = "paddle-res-type"
RESOURCE_ANNOTATION = "paddle-resource"

# resource kinds, paddlejob_types.go:37-41
RESOURCE_PS = "ps"
RESOURCE_WORKER = "worker"
RESOURCE_HETER = "heter"
RESOURCE_ORDER = (RESOURCE_PS, RESOURCE_WORKER, RESOURCE_HETER)  # GetResourceOrder, :250-256

# paddlejob_types.go:44-48
TRAINING_ROLE = {RESOURCE_PS: "PSERVER", RESOURCE_WORKER: "TRAINER", RESOURCE_HETER: "HETER"}


class PaddleJobMode(str, enum.Enum):  # :53-59
    PS = "PS"
    COLLECTIVE = "Collective"
    SINGLE = "Single"


class PaddleJobPhase(str, enum.Enum):  # :64-79
    STARTING = "Starting"
    PENDING = "Pending"
    SCALING = "Scaling"
    ABORTING = "Aborting"
    ABORTED = "Aborted"
    RUNNING = "Running"
    RESTARTING = "Restarting"
    COMPLETING = "Completing"
    COMPLETED = "Completed"
    TERMINATING = "Terminating"
    TERMINATED = "Terminated"
    FAILED = "Failed"
    SUCCEED = "Succeed"
    UNKNOWN = "Unknown"


class CleanPodPolicy(str, enum.Enum):  # :83-92
    ALWAYS = "Always"
    NEVER = "Never"
    ON_FAILURE = "OnFailure"
    ON_COMPLETION = "OnCompletion"


class ElasticStatus(str, enum.Enum):  # :97-102
    NONE = "NONE"
    DOING = "DOING"
    DONE = "DONE"
    ERROR = "ERROR"


class Intranet(str, enum.Enum):  # :106-110
    POD_IP = "PodIP"
    SERVICE = "Service"
    HOST = "Host"


def _drop_empty(d: dict) -> dict:
    """Go `omitempty`: drop None, "", 0, empty list/dict."""
    return {k: v for k, v in d.items() if v not in (None, "", 0, [], {})}


@dataclass
class SchedulingPolicy:  # :113-120
    min_available: Optional[int] = None
    queue: str = ""
    priority_class: str = ""
    min_resources: Optional[dict] = None

    @classmethod
    def from_dict(cls, d: Optional[dict]) -> Optional["SchedulingPolicy"]:
        if d is None:
            return None
        return cls(d.get("minAvailable"), d.get("queue", "") or "", d.get("priorityClass", "") or "",
                   copy.deepcopy(d.get("minResources")))

    def to_dict(self) -> dict:
        return _drop_empty({"minAvailable": self.min_available, "queue": self.queue,
                            "priorityClass": self.priority_class, "minResources": self.min_resources})


@dataclass
class ResourceSpec:  # :153-165
    replicas: int = 0
    requests: Optional[int] = None  # declared but never read by the reference controller
    limits: Optional[int] = None
    template: dict = field(default_factory=dict)  # core/v1 PodTemplateSpec, opaque

    @classmethod
    def from_dict(cls, d: Optional[dict]) -> Optional["ResourceSpec"]:
        if d is None:
            return None
        if "replicas" not in d:
            raise ValueError("spec.<resource>.replicas is required")  # CRD: required
        return cls(int(d["replicas"]), d.get("requests"), d.get("limits"), copy.deepcopy(d.get("template") or {}))

    def to_dict(self) -> dict:
        out = {"replicas": self.replicas}
        if self.requests is not None:
            out["requests"] = self.requests
        if self.limits is not None:
            out["limits"] = self.limits
        out["template"] = copy.deepcopy(self.template)
        return out

    # typed access into the opaque template ---------------------------------
    @property
    def pod_spec(self) -> dict:
        return self.template.setdefault("spec", {})

    @property
    def containers(self) -> list:
        return self.pod_spec.setdefault("containers", [])

    @property
    def scheduler_name(self) -> str:
        return self.pod_spec.get("schedulerName", "") or ""


@dataclass
class PaddleJobSpec:  # :123-151
    clean_pod_policy: str = ""  # an untyped string in the CRD: any value is accepted
    scheduling_policy: Optional[SchedulingPolicy] = None
    intranet: str = ""
    with_gloo: Optional[int] = None
    ps: Optional[ResourceSpec] = None
    worker: Optional[ResourceSpec] = None
    heter: Optional[ResourceSpec] = None
    elastic: Optional[int] = None

    KNOWN = ("cleanPodPolicy", "schedulingPolicy", "intranet", "withGloo", "ps", "worker", "heter", "elastic")

    @classmethod
    def from_dict(cls, d: Optional[dict]) -> "PaddleJobSpec":
        d = d or {}
        return cls(
            clean_pod_policy=d.get("cleanPodPolicy", "") or "",
            scheduling_policy=SchedulingPolicy.from_dict(d.get("schedulingPolicy")),
            intranet=d.get("intranet", "") or "",
            with_gloo=d.get("withGloo"),
            ps=ResourceSpec.from_dict(d.get("ps")),
            worker=ResourceSpec.from_dict(d.get("worker")),
            heter=ResourceSpec.from_dict(d.get("heter")),
            elastic=d.get("elastic"),
        )

    def to_dict(self) -> dict:
        out: dict[str, Any] = {}
        if self.clean_pod_policy:
            out["cleanPodPolicy"] = self.clean_pod_policy
        if self.scheduling_policy is not None:
            out["schedulingPolicy"] = self.scheduling_policy.to_dict()
        if self.intranet:
            out["intranet"] = self.intranet
        if self.with_gloo is not None:
            out["withGloo"] = self.with_gloo
        for k in RESOURCE_ORDER:
            rs = getattr(self, k)
            if rs is not None:
                out[k] = rs.to_dict()
        if self.elastic is not None:
            out["elastic"] = self.elastic
        return out


@dataclass
class ResourceStatus:  # :201-216
    pending: int = 0
    starting: int = 0
    running: int = 0
    failed: int = 0
    succeeded: int = 0
    unknown: int = 0
    refs: list = field(default_factory=list)  # core/v1 ObjectReference dicts

    @classmethod
    def from_dict(cls, d: Optional[dict]) -> Optional["ResourceStatus"]:
        if d is None:
            return None
        return cls(d.get("pending", 0), d.get("starting", 0), d.get("running", 0), d.get("failed", 0),
                   d.get("succeeded", 0), d.get("unknown", 0), copy.deepcopy(d.get("refs") or []))

    def to_dict(self) -> dict:
        return _drop_empty({"pending": self.pending, "starting": self.starting, "running": self.running,
                            "failed": self.failed, "succeeded": self.succeeded, "unknown": self.unknown,
                            "refs": copy.deepcopy(self.refs)})


@dataclass
class PaddleJobStatus:  # :168-199
    phase: str = ""
    mode: str = ""
    ps: Optional[ResourceStatus] = None
    worker: Optional[ResourceStatus] = None
    heter: Optional[ResourceStatus] = None
    elastic: Optional[str] = None
    start_time: Optional[str] = None  # RFC 3339, like metav1.Time
    completion_time: Optional[str] = None
    observed_generation: int = 0

    @classmethod
    def from_dict(cls, d: Optional[dict]) -> "PaddleJobStatus":
        d = d or {}
        return cls(d.get("phase", "") or "", d.get("mode", "") or "", ResourceStatus.from_dict(d.get("ps")),
                   ResourceStatus.from_dict(d.get("worker")), ResourceStatus.from_dict(d.get("heter")),
                   d.get("elastic"), d.get("startTime"), d.get("completionTime"), d.get("observedGeneration", 0))

    def to_dict(self) -> dict:
        out = _drop_empty({"phase": self.phase, "mode": self.mode, "elastic": self.elastic,
                           "startTime": self.start_time, "completionTime": self.completion_time,
                           "observedGeneration": self.observed_generation})
        for k in RESOURCE_ORDER:
            rs = getattr(self, k)
            if rs is not None:
                out[k] = rs.to_dict()
        return out


@dataclass
class ObjectMeta:
    name: str = ""
    namespace: str = "default"
    labels: dict = field(default_factory=dict)
    annotations: dict = field(default_factory=dict)
    finalizers: list = field(default_factory=list)
    owner_references: list = field(default_factory=list)
    uid: str = ""
    resource_version: str = ""
    generation: int = 0
    creation_timestamp: Optional[str] = None
    deletion_timestamp: Optional[str] = None

    @classmethod
    def from_dict(cls, d: Optional[dict]) -> "ObjectMeta":
        d = d or {}
        return cls(d.get("name", ""), d.get("namespace") or "default", dict(d.get("labels") or {}),
                   dict(d.get("annotations") or {}), list(d.get("finalizers") or []),
                   copy.deepcopy(d.get("ownerReferences") or []), d.get("uid", ""),
                   str(d.get("resourceVersion", "") or ""), d.get("generation", 0) or 0,
                   d.get("creationTimestamp"), d.get("deletionTimestamp"))

    def to_dict(self) -> dict:
        return _drop_empty({"name": self.name, "namespace": self.namespace, "labels": dict(self.labels),
                            "annotations": dict(self.annotations), "finalizers": list(self.finalizers),
                            "ownerReferences": copy.deepcopy(self.owner_references), "uid": self.uid,
                            "resourceVersion": self.resource_version, "generation": self.generation,
                            "creationTimestamp": self.creation_timestamp,
                            "deletionTimestamp": self.deletion_timestamp})


@dataclass
class PaddleJob:  # :226-232
    metadata: ObjectMeta = field(default_factory=ObjectMeta)
    spec: PaddleJobSpec = field(default_factory=PaddleJobSpec)
    status: PaddleJobStatus = field(default_factory=PaddleJobStatus)
    api_version: str = API_VERSION
    kind: str = KIND

    # -- accessors of the reference (:234-268) -------------------------------
    def get_specs(self) -> dict[str, Optional[ResourceSpec]]:
        return {RESOURCE_PS: self.spec.ps, RESOURCE_WORKER: self.spec.worker, RESOURCE_HETER: self.spec.heter}

    def get_statuses(self) -> dict[str, Optional[ResourceStatus]]:
        return {RESOURCE_PS: self.status.ps, RESOURCE_WORKER: self.status.worker, RESOURCE_HETER: self.status.heter}

    @staticmethod
    def get_resource_order() -> tuple:
        return RESOURCE_ORDER

    def set_status(self, res_type: str, status: Optional[ResourceStatus]) -> None:
        if res_type in RESOURCE_ORDER:
            setattr(self.status, res_type, status)

    # -- (de)serialisation -------------------------------------------------------
    @property
    def name(self) -> str:
        return self.metadata.name

    @property
    def namespace(self) -> str:
        return self.metadata.namespace

    @classmethod
    def from_dict(cls, d: dict) -> "PaddleJob":
        if d.get("kind") != KIND:
            raise ValueError(f"kind must be {KIND}, got {d.get('kind')!r}")
        api = d.get("apiVersion", "")
        if api.split("/")[0] != GROUP:
            raise ValueError(f"apiVersion must be {GROUP}/<version>, got {api!r}")
        return cls(ObjectMeta.from_dict(d.get("metadata")), PaddleJobSpec.from_dict(d.get("spec")),
                   PaddleJobStatus.from_dict(d.get("status")), api or API_VERSION, KIND)

    def to_dict(self) -> dict:
        return {"apiVersion": self.api_version, "kind": self.kind, "metadata": self.metadata.to_dict(),
                "spec": self.spec.to_dict(), "status": self.status.to_dict()}

    def deepcopy(self) -> "PaddleJob":
        return PaddleJob.from_dict(copy.deepcopy(self.to_dict()))


def pruned_spec_fields(doc: dict) -> list[str]:
    """Top-level spec keys a CRD-validating apiserver would prune (e.g. the
    misspelt `cleanPolicy` of docs/user-guide.md:312 in the reference)."""
    return sorted(k for k in (doc.get("spec") or {}) if k not in PaddleJobSpec.KNOWN)
