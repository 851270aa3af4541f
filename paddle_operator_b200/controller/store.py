"""In-memory object store: the stand-in for kube-apiserver + etcd.

Holds Kubernetes-shaped dict objects keyed by (kind, namespace, name) with the
semantics the reference controller relies on (controller-runtime client,
controllers/paddlejob_controller.go:105-131,383-405,460-489,520-571):

* resourceVersion bumped on every write; Update with a stale version -> Conflict
* a status "subresource": update_status() touches only .status, update() never does
* metadata.generation bumped when .spec changes
* finalizers: delete() of an object with finalizers only sets deletionTimestamp;
  the object goes away when the last finalizer is removed
* owner references: deleting an owner garbage-collects its dependents
  (background propagation)
* watches: every change is delivered to subscribers, which is what drives the
  reconcile work-queue
* an index of objects by controlling owner (the `.metadata.controller` field
  index of the reference, :520-532)
"""
from __future__ import annotations

import copy
import datetime as _dt
import re
import threading
import uuid
from typing import Callable, Optional


class StoreError(Exception):
    pass


class NotFound(StoreError):
    pass


class AlreadyExists(StoreError):
    pass


class Conflict(StoreError):
    pass


class Invalid(StoreError):
    """A name or namespace Kubernetes would reject (the API server answers 422)."""


# Kubernetes validates object names as RFC 1123 DNS subdomains (labels of [a-z0-9-] joined by single dots, at most
# 253 characters) and namespaces as single labels of at most 63.  The reference inherits that from the API server;
# here it also keeps names usable as path components (no "/", no "..") - the supervisor builds work directories
# and log paths from them.
_LABEL = r"[a-z0-9]([-a-z0-9]*[a-z0-9])?"
_NAME_RE = re.compile(rf"^{_LABEL}(\.{_LABEL})*$")
_NS_RE = re.compile(rf"^{_LABEL}$")


def validate_name(name, what: str = "metadata.name", namespace: bool = False) -> None:
    limit = 63 if namespace else 253
    if not isinstance(name, str) or not name or len(name) > limit or not (_NS_RE if namespace else _NAME_RE).match(name):
        raise Invalid(f"{what}: {name!r} is not a valid RFC 1123 {'label' if namespace else 'subdomain'} "
                      f"(lower-case alphanumerics and '-'{'' if namespace else ', labels joined by single dots'}, "
                      f"at most {limit} characters)")


def _now() -> str:
    return _dt.datetime.now(_dt.timezone.utc).strftime("%Y-%m-%dT%H:%M:%SZ")


Key = tuple  # (kind, namespace, name)
WatchFn = Callable[[str, dict], None]  # (event_type, object)


def _clone(o):
    """Deep copy of a JSON-shaped object (dict / list / scalars) - several times faster than copy.deepcopy, which
    the store would otherwise spend most of its time in (profiles/control_plane_bench.py)."""
    if isinstance(o, dict):
        return {k: _clone(v) for k, v in o.items()}
    if isinstance(o, list):
        return [_clone(v) for v in o]
    if o is None or isinstance(o, (str, int, float, bool)):
        return o
    return copy.deepcopy(o)


class ObjectStore:
    def __init__(self):
        self._lock = threading.RLock()
        self._objs: dict[Key, dict] = {}
        self._rv = 0
        self._watchers: list[WatchFn] = []
        # (owner kind, namespace, owner name) -> keys of the objects that carry an ownerReference to it: serves
        # list(owner=...) - the `.metadata.controller` field index of the reference - and garbage collection
        self._by_owner: dict[tuple, set] = {}

    # -- helpers ----------------------------------------------------------------
    @staticmethod
    def key_of(obj: dict) -> Key:
        m = obj.get("metadata") or {}
        return (obj.get("kind", ""), m.get("namespace") or "default", m.get("name", ""))

    def _index(self, key: Key, obj: dict, add: bool) -> None:
        for ref in obj["metadata"].get("ownerReferences") or []:
            okey = (ref.get("kind", ""), key[1], ref.get("name", ""))
            if add:
                self._by_owner.setdefault(okey, set()).add(key)
            else:
                bucket = self._by_owner.get(okey)
                if bucket is not None:
                    bucket.discard(key)
                    if not bucket:
                        del self._by_owner[okey]

    def _next_rv(self) -> str:
        self._rv += 1
        return str(self._rv)

    def _notify(self, typ: str, obj: dict) -> None:
        for w in list(self._watchers):
            w(typ, _clone(obj))

    def watch(self, fn: WatchFn) -> Callable[[], None]:
        with self._lock:
            self._watchers.append(fn)

        def cancel():
            with self._lock:
                if fn in self._watchers:
                    self._watchers.remove(fn)

        return cancel

    # -- CRUD ---------------------------------------------------------------------
    def create(self, obj: dict) -> dict:
        obj = _clone(obj)
        meta = obj.setdefault("metadata", {})
        meta.setdefault("namespace", "default")
        if not meta.get("name"):
            raise StoreError("metadata.name is required")
        validate_name(meta["name"])
        validate_name(meta["namespace"], "metadata.namespace", namespace=True)
        key = self.key_of(obj)
        with self._lock:
            if key in self._objs:
                raise AlreadyExists(f"{key[0]} {key[1]}/{key[2]} already exists")
            for ref in meta.get("ownerReferences") or []:
                if not ref.get("uid"):
                    okey = (ref.get("kind", ""), meta["namespace"], ref.get("name", ""))
                    owner = self._objs.get(okey)
                    if owner is not None:
                        ref["uid"] = owner["metadata"]["uid"]
            meta["uid"] = meta.get("uid") or str(uuid.uuid4())
            meta["resourceVersion"] = self._next_rv()
            meta["generation"] = 1
            meta["creationTimestamp"] = meta.get("creationTimestamp") or _now()
            meta.pop("deletionTimestamp", None)
            self._objs[key] = obj
            self._index(key, obj, True)
            self._notify("ADDED", obj)
            return _clone(obj)

    def restore(self, obj: dict) -> dict:
        """Re-inserts an object exactly as it was persisted (uid, generation, creation time, finalizers, a pending
        deletionTimestamp and its status are kept; only the resourceVersion is new).  Used by controller.persist
        when a manager starts on a state directory left by a previous one."""
        obj = _clone(obj)
        meta = obj.setdefault("metadata", {})
        meta.setdefault("namespace", "default")
        if not meta.get("name"):
            raise StoreError("metadata.name is required")
        validate_name(meta["name"])
        validate_name(meta["namespace"], "metadata.namespace", namespace=True)
        key = self.key_of(obj)
        with self._lock:
            if key in self._objs:
                raise AlreadyExists(f"{key[0]} {key[1]}/{key[2]} already exists")
            meta["uid"] = meta.get("uid") or str(uuid.uuid4())
            meta["resourceVersion"] = self._next_rv()
            meta["generation"] = int(meta.get("generation") or 1)
            meta["creationTimestamp"] = meta.get("creationTimestamp") or _now()
            self._objs[key] = obj
            self._index(key, obj, True)
            self._notify("ADDED", obj)
            return _clone(obj)

    def get(self, kind: str, namespace: str, name: str) -> dict:
        with self._lock:
            obj = self._objs.get((kind, namespace or "default", name))
            if obj is None:
                raise NotFound(f"{kind} {namespace}/{name} not found")
            return _clone(obj)

    def try_get(self, kind: str, namespace: str, name: str) -> Optional[dict]:
        try:
            return self.get(kind, namespace, name)
        except NotFound:
            return None

    def list(self, kind: str, namespace: Optional[str] = None, owner: Optional[str] = None,
             owner_kind: str = "PaddleJob", labels: Optional[dict] = None) -> list[dict]:
        """Objects of `kind`, optionally restricted to a namespace, to those
        controlled by the named owner (the field index of the reference) and to
        a label selector.  Sorted by name for determinism."""
        out = []
        with self._lock:
            if owner is not None and namespace is not None:
                # the owner index: only the owner's dependents are looked at, not every object of the kind
                cand = [(k, self._objs[k]) for k in self._by_owner.get((owner_kind, namespace, owner), ()) if k in self._objs]
            else:
                cand = self._objs.items()
            for (k, ns, _), obj in cand:
                if k != kind or (namespace is not None and ns != namespace):
                    continue  # (a key-tuple compare; cloning the matches is what costs)
                if owner is not None:
                    ctrl = next((r for r in obj["metadata"].get("ownerReferences") or [] if r.get("controller")), None)
                    if ctrl is None or ctrl.get("name") != owner or ctrl.get("kind") != owner_kind:
                        continue
                if labels:
                    have = obj["metadata"].get("labels") or {}
                    if any(have.get(a) != b for a, b in labels.items()):
                        continue
                out.append(_clone(obj))
        out.sort(key=lambda o: (o["metadata"].get("namespace", ""), o["metadata"]["name"]))
        return out

    def _check_rv(self, cur: dict, new: dict) -> None:
        rv = str((new.get("metadata") or {}).get("resourceVersion", "") or "")
        if rv and rv != cur["metadata"]["resourceVersion"]:
            raise Conflict(f"{cur.get('kind')} {cur['metadata']['name']}: the object has been modified "
                           f"(have {cur['metadata']['resourceVersion']}, update carries {rv})")

    def update(self, obj: dict) -> dict:
        """Replaces everything except .status.  Removes the object when it is
        being deleted and the update drops the last finalizer."""
        key = self.key_of(obj)
        with self._lock:
            cur = self._objs.get(key)
            if cur is None:
                raise NotFound(f"{key[0]} {key[1]}/{key[2]} not found")
            self._check_rv(cur, obj)
            new = _clone(obj)
            new["status"] = _clone(cur.get("status", {}))
            meta, cmeta = new.setdefault("metadata", {}), cur["metadata"]
            for k in ("uid", "creationTimestamp", "deletionTimestamp"):
                if k in cmeta:
                    meta[k] = cmeta[k]
                else:
                    meta.pop(k, None)
            meta["generation"] = cmeta.get("generation", 1) + (1 if new.get("spec") != cur.get("spec") else 0)
            meta["resourceVersion"] = self._next_rv()
            if meta.get("deletionTimestamp") and not meta.get("finalizers"):
                self._remove(key)
                return _clone(new)
            self._index(key, cur, False)
            self._objs[key] = new
            self._index(key, new, True)
            self._notify("MODIFIED", new)
            return _clone(new)

    def update_status(self, obj: dict) -> dict:
        key = self.key_of(obj)
        with self._lock:
            cur = self._objs.get(key)
            if cur is None:
                raise NotFound(f"{key[0]} {key[1]}/{key[2]} not found")
            self._check_rv(cur, obj)
            cur["status"] = _clone(obj.get("status", {}))
            cur["metadata"]["resourceVersion"] = self._next_rv()
            self._notify("MODIFIED", cur)
            return _clone(cur)

    def patch_status(self, kind: str, namespace: str, name: str, fn: Callable[[dict], None]) -> dict:
        """Read-modify-write of .status without a resourceVersion check (what a
        kubelet does to its pods)."""
        with self._lock:
            cur = self._objs.get((kind, namespace or "default", name))
            if cur is None:
                raise NotFound(f"{kind} {namespace}/{name} not found")
            st = cur.setdefault("status", {})
            fn(st)
            cur["metadata"]["resourceVersion"] = self._next_rv()
            self._notify("MODIFIED", cur)
            return _clone(cur)

    def delete(self, kind: str, namespace: str, name: str) -> None:
        key = (kind, namespace or "default", name)
        with self._lock:
            cur = self._objs.get(key)
            if cur is None:
                raise NotFound(f"{kind} {namespace}/{name} not found")
            if cur["metadata"].get("finalizers"):
                if not cur["metadata"].get("deletionTimestamp"):
                    cur["metadata"]["deletionTimestamp"] = _now()
                    cur["metadata"]["resourceVersion"] = self._next_rv()
                    self._notify("MODIFIED", cur)
                return
            self._remove(key)

    def _remove(self, key: Key) -> None:
        obj = self._objs.pop(key, None)
        if obj is None:
            return
        self._index(key, obj, False)
        self._notify("DELETED", obj)
        # garbage-collect dependents (background propagation); candidates come from the owner index
        uid = obj["metadata"].get("uid")
        deps = [k for k in list(self._by_owner.get((key[0], key[1], key[2]), ()))
                if k in self._objs and any(r.get("uid") == uid or (not r.get("uid") and r.get("name") == key[2]
                                                                      and r.get("kind") == key[0])
                                           for r in self._objs[k]["metadata"].get("ownerReferences") or [])]
        for k in deps:
            d = self._objs.get(k)
            if d is None:
                continue
            if d["metadata"].get("finalizers"):
                d["metadata"]["deletionTimestamp"] = d["metadata"].get("deletionTimestamp") or _now()
                self._notify("MODIFIED", d)
            else:
                self._remove(k)

    def kinds(self) -> set[str]:
        with self._lock:
            return {k[0] for k in self._objs}

    def dump(self) -> list[dict]:
        with self._lock:
            return [_clone(o) for o in self._objs.values()]


class EventRecorder:
    """k8s Events on the job (controllers/paddlejob_controller.go:383-405,215):
    Normal/Warning, reason, message.  Stored as Event objects so `pdj events`
    and the tests can read the audit trail.

    Kept bounded the way the API server keeps them bounded: a repeat of the latest event of an object (same type,
    reason and message) bumps its `count` / `lastTimestamp` instead of creating a new object, events older than
    `ttl_seconds` (Kubernetes: 1 h) are dropped when the next one is recorded, at most `max_per_object` are kept
    per involved object, and every event carries an ownerReference to its object so that deleting a job
    garbage-collects its events."""

    def __init__(self, store: ObjectStore, component: str = "paddlejob-controller", ttl_seconds: float = 3600.0,
                 max_per_object: int = 256):
        self.store, self.component = store, component
        self.ttl_seconds, self.max_per_object = ttl_seconds, max_per_object
        self._n = 0
        self._lock = threading.Lock()
        # (kind, namespace, name) of the involved object -> [(event name, monotonic time)], oldest first
        self._recent: dict[tuple, list] = {}

    def event(self, involved: dict, etype: str, reason: str, message: str) -> None:
        import time as _time

        m = involved.get("metadata") or {}
        ns, name, kind = m.get("namespace", "default"), m.get("name", ""), involved.get("kind", "")
        okey = (kind, ns, name)
        now = _time.monotonic()
        with self._lock:
            mine = self._recent.setdefault(okey, [])
            # expire by age and by count
            while mine and (now - mine[0][1] > self.ttl_seconds or len(mine) >= self.max_per_object):
                old, _ = mine.pop(0)
                try:
                    self.store.delete("Event", ns, old)
                except NotFound:
                    pass
            if mine:
                last = self.store.try_get("Event", ns, mine[-1][0])
                if last is None:
                    mine.clear()  # collected together with a previous incarnation of the object
                elif (last.get("type"), last.get("reason"), last.get("message")) == (etype, reason, message) and \
                        last["involvedObject"].get("uid", "") == m.get("uid", ""):
                    last["count"] = int(last.get("count", 1)) + 1
                    last["lastTimestamp"] = _now()
                    last["metadata"].pop("resourceVersion", None)
                    self.store.update(last)
                    mine[-1] = (mine[-1][0], now)
                    return
            self._n += 1
            ev_name = f"{name or 'obj'}.{self._n:08x}"
            mine.append((ev_name, now))
        owner = ([{"apiVersion": involved.get("apiVersion", ""), "kind": kind, "name": name, "uid": m["uid"]}]
                 if m.get("uid") else [])
        self.store.create({
            "apiVersion": "v1", "kind": "Event",
            "metadata": {"name": ev_name, "namespace": ns, "ownerReferences": owner},
            "involvedObject": {"kind": kind, "name": name, "namespace": ns, "uid": m.get("uid", "")},
            "type": etype, "reason": reason, "message": message,
            "source": {"component": self.component}, "firstTimestamp": _now(), "lastTimestamp": _now(), "count": 1,
        })

    def events_for(self, kind: str, namespace: str, name: str) -> list[dict]:
        evs = [e for e in self.store.list("Event", namespace)
               if e["involvedObject"]["kind"] == kind and e["involvedObject"]["name"] == name]
        evs.sort(key=lambda e: e["metadata"]["name"])
        return evs
