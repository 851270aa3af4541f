"""Elastic world-size sync: the controller's half of elastic training.

Mirror of controllers/paddlejob_elastic.go:27-55 in the reference: in
Collective mode the desired number of workers is written to the etcd key
`/paddle/<namespace>-<name>/np`, but only when that key already exists with a
different value (the workers create it; the controller never does).  The
client is anything with `get(key) -> list[(key, value)]` and `put(key, value)`,
e.g. runtime.etcd_client.EtcdClient or the in-process runtime.etcd_store.
"""
from __future__ import annotations

from typing import Optional

from ..api.types import PaddleJob, PaddleJobMode

NP_TIMEOUT_S = 3.0  # context.WithTimeout in the reference (:42)


def np_key(namespace: str, name: str) -> str:
    return f"/paddle/{namespace}-{name}/np"


def job_prefix(namespace: str, name: str) -> str:
    return f"/paddle/{namespace}-{name}"


def sync_etcd(etcd, path: str, np: str) -> bool:
    kvs = etcd.get(path)
    if len(kvs) != 1:
        return False
    value = kvs[0][1]
    if isinstance(value, bytes):
        value = value.decode()
    if value == np:
        return False
    etcd.put(path, np)
    return True


def sync_np(etcd, pdj: PaddleJob) -> Optional[str]:
    """Returns the new np (as a string) when the key was updated, else None."""
    if pdj.spec.worker is None or pdj.spec.ps is not None:
        return None
    # The reference syncs only while status.mode is Collective (paddlejob_elastic.go:45), and the
    # mode flips to Single when an elastic job is scaled to one worker (helper.go:191-199), which
    # leaves np stale at 2 and the last rank waiting for a peer that was deleted.  An elastic job
    # without ps is collective by construction, so the sync also covers the one-worker case here.
    if pdj.status.mode != PaddleJobMode.COLLECTIVE.value and pdj.spec.elastic is None:
        return None
    np = str(pdj.spec.worker.replicas)
    return np if sync_etcd(etcd, np_key(pdj.namespace, pdj.name), np) else None
