"""YAML / JSON loader for PaddleJob documents and their companions.

Accepts the reference's example manifests verbatim (deploy/examples/*.yaml,
deploy/elastic/*.yaml, config/samples/batch_v1_paddlejob.yaml), including
multi-document files.  PaddleJob documents become `PaddleJob` objects; anything
else (e.g. the etcd Deployment/Service of deploy/elastic/etcd.yaml) is returned
as a plain dict so callers can decide what to do with it.

Validation mirrors what the generated CRD enforces
(config/crd/bases/batch.paddlepaddle.org_paddlejobs.yaml: `replicas` required
in each of ps/worker/heter, unknown spec fields pruned) and nothing more - the
reference accepts e.g. `cleanPodPolicy: OnFailed` (deploy/examples/deepfm.yaml:6)
because that field is an untyped string in the schema.
"""
from __future__ import annotations

import io
import json
import os
from typing import Any, Iterable, Union

import yaml

from .types import API_VERSION, GROUP, KIND, PaddleJob, pruned_spec_fields


class ValidationError(ValueError):
    pass


def _validate(doc: dict) -> None:
    meta = doc.get("metadata") or {}
    if not meta.get("name"):
        raise ValidationError("metadata.name is required")
    # the API server of the reference rejects these with 422 before the controller ever sees the job
    from ..controller.store import Invalid, validate_name

    try:
        validate_name(meta["name"])
        if len(meta["name"]) > 63 or "." in meta["name"]:
            # pods, services and the `<job>-worker-0` host names derive from it: a single DNS label
            validate_name(meta["name"], namespace=True)
        if meta.get("namespace") is not None:
            validate_name(meta["namespace"], "metadata.namespace", namespace=True)
    except Invalid as e:
        raise ValidationError(str(e)) from None
    spec = doc.get("spec")
    if spec is not None and not isinstance(spec, dict):
        raise ValidationError("spec must be a mapping")
    for res in ("ps", "worker", "heter"):
        rs = (spec or {}).get(res)
        if rs is None:
            continue
        if not isinstance(rs, dict):
            raise ValidationError(f"spec.{res} must be a mapping")
        if "replicas" not in rs:
            raise ValidationError(f"spec.{res}.replicas: Required value")
        if not isinstance(rs["replicas"], int) or isinstance(rs["replicas"], bool):
            raise ValidationError(f"spec.{res}.replicas must be an integer")
        for k in ("requests", "limits"):
            if k in rs and rs[k] is not None and not isinstance(rs[k], int):
                raise ValidationError(f"spec.{res}.{k} must be an integer")
        tmpl = rs.get("template")
        if tmpl is not None and not isinstance(tmpl, dict):
            raise ValidationError(f"spec.{res}.template must be a mapping")
        # requests / limits are the minimum / maximum replica count of the kind (api/v1/paddlejob_types.go:
        # 155-161).  The reference never reads them; here an ELASTIC job may only be scaled inside that window -
        # non-elastic jobs keep the reference's behaviour (stored, not enforced).
        lo, hi = rs.get("requests"), rs.get("limits")
        if lo is not None and hi is not None and lo > hi:
            raise ValidationError(f"spec.{res}.requests ({lo}) is larger than spec.{res}.limits ({hi})")
        if (spec or {}).get("elastic") is not None:
            if lo is not None and rs["replicas"] < lo:
                raise ValidationError(f"spec.{res}.replicas ({rs['replicas']}) is below spec.{res}.requests ({lo}) of this elastic job")
            if hi is not None and rs["replicas"] > hi:
                raise ValidationError(f"spec.{res}.replicas ({rs['replicas']}) is above spec.{res}.limits ({hi}) of this elastic job")
    for k in ("withGloo", "elastic"):
        v = (spec or {}).get(k)
        if v is not None and (not isinstance(v, int) or isinstance(v, bool)):
            raise ValidationError(f"spec.{k} must be an integer")


def is_paddlejob(doc: Any) -> bool:
    return isinstance(doc, dict) and doc.get("kind") == KIND and str(doc.get("apiVersion", "")).startswith(GROUP + "/")


def paddlejob_from_doc(doc: dict) -> PaddleJob:
    _validate(doc)
    try:
        return PaddleJob.from_dict(doc)
    except ValueError as e:
        raise ValidationError(str(e)) from e


def load_documents(src: Union[str, os.PathLike, io.IOBase]) -> list[Any]:
    """Parses YAML (or JSON, a YAML subset) text, a path, or a file object into
    a list of PaddleJob objects and raw dicts, in document order."""
    if hasattr(src, "read"):
        text = src.read()
    else:
        s = os.fspath(src)
        if "\n" not in s and os.path.exists(s):
            with open(s) as f:
                text = f.read()
        else:
            text = s
    out: list[Any] = []
    for doc in yaml.safe_load_all(text):
        if doc is None:
            continue
        if isinstance(doc, dict) and doc.get("kind") == "List" and isinstance(doc.get("items"), list):
            docs: Iterable = doc["items"]
        else:
            docs = [doc]
        for d in docs:
            out.append(paddlejob_from_doc(d) if is_paddlejob(d) else d)
    return out


def load_paddlejobs(src) -> list[PaddleJob]:
    return [d for d in load_documents(src) if isinstance(d, PaddleJob)]


def load_paddlejob(src) -> PaddleJob:
    jobs = load_paddlejobs(src)
    if len(jobs) != 1:
        raise ValidationError(f"expected exactly one PaddleJob document, found {len(jobs)}")
    return jobs[0]


def warnings_for(doc_or_job: Union[dict, PaddleJob]) -> list[str]:
    """Human-readable warnings an apiserver would not give but a user wants:
    pruned fields and clean-pod policies that match no branch of the controller
    (controllers/paddlejob_controller.go:221-232)."""
    d = doc_or_job.to_dict() if isinstance(doc_or_job, PaddleJob) else doc_or_job
    w = [f"unknown field spec.{k} is pruned" for k in pruned_spec_fields(d)]
    pol = (d.get("spec") or {}).get("cleanPodPolicy", "")
    if pol not in ("", "Always", "Never", "OnFailure", "OnCompletion"):
        w.append(f"cleanPodPolicy {pol!r} matches no policy; it behaves like Never")
    return w


def dump_yaml(obj: Union[PaddleJob, dict, list]) -> str:
    def conv(o):
        return o.to_dict() if isinstance(o, PaddleJob) else o

    if isinstance(obj, list):
        return yaml.safe_dump_all([conv(o) for o in obj], sort_keys=False)
    return yaml.safe_dump(conv(obj), sort_keys=False)


def dump_json(obj: Union[PaddleJob, dict]) -> str:
    return json.dumps(obj.to_dict() if isinstance(obj, PaddleJob) else obj, indent=2)


def openapi_schema() -> dict:
    """Top level of the CRD's OpenAPI v3 schema (PodTemplateSpec left open)."""
    res = {
        "type": "object",
        "required": ["replicas"],
        "properties": {
            "replicas": {"type": "integer"},
            "requests": {"type": "integer"},
            "limits": {"type": "integer"},
            "template": {"type": "object", "x-kubernetes-preserve-unknown-fields": True},
        },
    }
    rstat = {
        "type": "object",
        "properties": {**{k: {"type": "integer"} for k in ("pending", "starting", "running", "failed", "succeeded",
                                                            "unknown")},
                       "refs": {"type": "array", "items": {"type": "object"}}},
    }
    return {
        "type": "object",
        "properties": {
            "apiVersion": {"type": "string", "enum": [API_VERSION]},
            "kind": {"type": "string", "enum": [KIND]},
            "metadata": {"type": "object"},
            "spec": {
                "type": "object",
                "properties": {
                    "cleanPodPolicy": {"type": "string"},
                    "schedulingPolicy": {
                        "type": "object",
                        "properties": {"minAvailable": {"type": "integer", "format": "int32"},
                                       "queue": {"type": "string"}, "priorityClass": {"type": "string"},
                                       "minResources": {"type": "object"}},
                    },
                    "intranet": {"type": "string"},
                    "withGloo": {"type": "integer"},
                    "elastic": {"type": "integer"},
                    "ps": res, "worker": res, "heter": res,
                },
            },
            "status": {
                "type": "object",
                "properties": {"phase": {"type": "string"}, "mode": {"type": "string"}, "ps": rstat, "worker": rstat,
                               "heter": rstat, "elastic": {"type": "string"},
                               "startTime": {"type": "string", "format": "date-time"},
                               "completionTime": {"type": "string", "format": "date-time"},
                               "observedGeneration": {"type": "integer"}},
            },
        },
    }
