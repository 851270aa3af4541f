"""An etcd-v3-compatible key-value store, in process.

The reference needs a real etcd for elastic jobs (`--etcd-server`, main.go:
114-125; deploy/elastic/etcd.yaml runs a single-replica bitnami/etcd).  No etcd
binary exists on this box, so the runtime brings its own store with the subset
of the v3 semantics that the controller (controllers/paddlejob_elastic.go:27-39:
Get / Put) and the launcher side of Paddle's elastic manager (leases with TTL
and keep-alive for node registration, prefix watches on `/paddle/<job>/`,
create-if-absent transactions) rely on:

  * a single monotonically increasing revision; per-key create_revision,
    mod_revision, version;
  * Range (key, [key, range_end), prefix), Put (optional lease, prev_kv),
    DeleteRange, Txn (compare on create/mod/version/value; put/range/delete ops);
  * leases: grant / keep-alive / revoke / time-to-live; expiry deletes the keys;
  * watches from a revision, on a key or a range, with PUT / DELETE events.

`runtime.etcd_server` exposes it over etcd's JSON gateway (`/v3/kv/range`, ...)
so that the same client also talks to a real etcd.
"""
from __future__ import annotations

import threading
import time
from dataclasses import dataclass, field
from typing import Callable, Optional


def prefix_range_end(prefix: bytes) -> bytes:
    """etcd's convention: the smallest key greater than every key with `prefix`."""
    b = bytearray(prefix)
    for i in range(len(b) - 1, -1, -1):
        if b[i] < 0xFF:
            b[i] += 1
            return bytes(b[: i + 1])
    return b"\0"


def _b(x) -> bytes:
    return x if isinstance(x, bytes) else str(x).encode()


@dataclass
class KeyValue:
    key: bytes
    value: bytes
    create_revision: int
    mod_revision: int
    version: int
    lease: int = 0


@dataclass
class Event:
    type: str  # "PUT" | "DELETE"
    kv: KeyValue
    prev_kv: Optional[KeyValue] = None


@dataclass
class _Lease:
    id: int
    ttl: int
    expiry: float
    keys: set = field(default_factory=set)


@dataclass
class _Watcher:
    id: int
    key: bytes
    range_end: Optional[bytes]
    fn: Callable[[list], None]

    def matches(self, k: bytes) -> bool:
        if not self.range_end:
            return k == self.key
        if self.range_end == b"\0":
            return k >= self.key
        return self.key <= k < self.range_end


class EtcdStore:
    def __init__(self, reap_interval: float = 0.1):
        self._lock = threading.RLock()
        self._kv: dict[bytes, KeyValue] = {}
        self._rev = 1
        self._leases: dict[int, _Lease] = {}
        self._next_lease = 0x1000
        self._watchers: dict[int, _Watcher] = {}
        self._next_watch = 1
        self._history: list[tuple[int, Event]] = []  # (revision, event), bounded
        self._history_cap = 100000
        self._stop = threading.Event()
        self._reaper = threading.Thread(target=self._reap_loop, args=(reap_interval,), daemon=True)
        self._reaper.start()

    # -- internals -------------------------------------------------------------
    @property
    def revision(self) -> int:
        with self._lock:
            return self._rev

    def _in_range(self, k: bytes, key: bytes, range_end: Optional[bytes]) -> bool:
        if not range_end:
            return k == key
        if range_end == b"\0":
            return k >= key
        return key <= k < range_end

    def _emit(self, events: list) -> None:
        for ev in events:
            self._history.append((ev.kv.mod_revision, ev))
        if len(self._history) > self._history_cap:
            del self._history[: len(self._history) - self._history_cap]
        for w in list(self._watchers.values()):
            mine = [e for e in events if w.matches(e.kv.key)]
            if mine:
                try:
                    w.fn(mine)
                except Exception:  # noqa: BLE001 - a broken watcher must not break writers
                    pass

    def _reap_loop(self, interval: float) -> None:
        while not self._stop.wait(interval):
            now = time.monotonic()
            with self._lock:
                dead = [l.id for l in self._leases.values() if l.expiry <= now]
            for lid in dead:
                self.lease_revoke(lid)

    def close(self) -> None:
        self._stop.set()

    # -- KV -------------------------------------------------------------------------
    def range(self, key, range_end=None, limit: int = 0, keys_only: bool = False, count_only: bool = False,
              revision: int = 0) -> tuple[list[KeyValue], int, int]:
        """Returns (kvs sorted by key, count, store revision)."""
        key, range_end = _b(key), (_b(range_end) if range_end else None)
        with self._lock:
            ks = sorted(k for k in self._kv if self._in_range(k, key, range_end))
            count = len(ks)
            if count_only:
                return [], count, self._rev
            if limit:
                ks = ks[:limit]
            out = []
            for k in ks:
                kv = self._kv[k]
                out.append(KeyValue(kv.key, b"" if keys_only else kv.value, kv.create_revision, kv.mod_revision,
                                    kv.version, kv.lease))
            return out, count, self._rev

    def put(self, key, value, lease: int = 0, prev_kv: bool = False) -> tuple[Optional[KeyValue], int]:
        key, value = _b(key), _b(value)
        with self._lock:
            if lease and lease not in self._leases:
                raise KeyError(f"etcdserver: requested lease not found ({lease})")
            ev, prev = self._put_locked(key, value, lease)
            self._emit([ev])
            return (prev if prev_kv else None), self._rev

    def _put_locked(self, key: bytes, value: bytes, lease: int):
        self._rev += 1
        prev = self._kv.get(key)
        if prev is not None and prev.lease and prev.lease != lease and prev.lease in self._leases:
            self._leases[prev.lease].keys.discard(key)
        kv = KeyValue(key, value, prev.create_revision if prev else self._rev, self._rev,
                      (prev.version + 1) if prev else 1, lease)
        self._kv[key] = kv
        if lease:
            self._leases[lease].keys.add(key)
        return Event("PUT", kv, prev), prev

    def delete_range(self, key, range_end=None, prev_kv: bool = False) -> tuple[int, list[KeyValue], int]:
        key, range_end = _b(key), (_b(range_end) if range_end else None)
        with self._lock:
            events = self._delete_locked(key, range_end)
            self._emit(events)
            return len(events), ([e.prev_kv for e in events] if prev_kv else []), self._rev

    def _delete_locked(self, key: bytes, range_end: Optional[bytes]) -> list:
        ks = sorted(k for k in self._kv if self._in_range(k, key, range_end))
        if not ks:
            return []
        self._rev += 1
        events = []
        for k in ks:
            prev = self._kv.pop(k)
            if prev.lease in self._leases:
                self._leases[prev.lease].keys.discard(k)
            events.append(Event("DELETE", KeyValue(k, b"", 0, self._rev, 0, 0), prev))
        return events

    def txn(self, compare: list, success: list, failure: list) -> tuple[bool, list, int]:
        """compare: [(key, target, op, value)] with target in create/mod/version/value/lease and op in
        =,!=,<,>.  success/failure: [("put", key, value, lease) | ("range", key, range_end) |
        ("delete", key, range_end)].  Atomic."""
        with self._lock:
            ok = all(self._cmp(*c) for c in compare)
            results, events = [], []
            for op in (success if ok else failure):
                if op[0] == "put":
                    lease = op[3] if len(op) > 3 else 0
                    if lease and lease not in self._leases:
                        raise KeyError(f"etcdserver: requested lease not found ({lease})")
                    ev, _ = self._put_locked(_b(op[1]), _b(op[2]), lease)
                    events.append(ev)
                    results.append(("put", None))
                elif op[0] == "range":
                    kvs, count, _ = self.range(op[1], op[2] if len(op) > 2 else None)
                    results.append(("range", kvs))
                elif op[0] == "delete":
                    evs = self._delete_locked(_b(op[1]), _b(op[2]) if len(op) > 2 and op[2] else None)
                    events.extend(evs)
                    results.append(("delete", len(evs)))
                else:
                    raise ValueError(f"unknown txn op {op[0]!r}")
            self._emit(events)
            return ok, results, self._rev

    def _cmp(self, key, target: str, op: str, value) -> bool:
        kv = self._kv.get(_b(key))
        target = target.lower()
        if target == "value":
            have, want = (kv.value if kv else b""), _b(value)
        else:
            attr = {"create": "create_revision", "mod": "mod_revision", "version": "version", "lease": "lease"}[target]
            have, want = (getattr(kv, attr) if kv else 0), int(value)
        return {"=": have == want, "!=": have != want, ">": have > want, "<": have < want}[op]

    # -- leases --------------------------------------------------------------------------
    def lease_grant(self, ttl: int, lease_id: int = 0) -> tuple[int, int]:
        with self._lock:
            if not lease_id:
                self._next_lease += 1
                lease_id = self._next_lease
            if lease_id in self._leases:
                raise KeyError("etcdserver: lease already exists")
            ttl = max(1, int(ttl))
            self._leases[lease_id] = _Lease(lease_id, ttl, time.monotonic() + ttl)
            return lease_id, ttl

    def lease_keepalive(self, lease_id: int) -> int:
        """Returns the refreshed TTL, or 0 when the lease does not exist (as etcd does)."""
        with self._lock:
            l = self._leases.get(lease_id)
            if l is None:
                return 0
            l.expiry = time.monotonic() + l.ttl
            return l.ttl

    def lease_revoke(self, lease_id: int) -> bool:
        with self._lock:
            l = self._leases.pop(lease_id, None)
            if l is None:
                return False
            events = []
            for k in sorted(l.keys):
                events.extend(self._delete_locked(k, None))
            self._emit(events)
            return True

    def lease_ttl(self, lease_id: int) -> tuple[int, int, list]:
        """(remaining seconds or -1, granted ttl, keys)."""
        with self._lock:
            l = self._leases.get(lease_id)
            if l is None:
                return -1, 0, []
            return max(0, int(l.expiry - time.monotonic())), l.ttl, sorted(l.keys)

    def leases(self) -> list[int]:
        with self._lock:
            return sorted(self._leases)

    # -- watch -------------------------------------------------------------------------------
    def watch(self, key, range_end=None, start_revision: int = 0, fn: Optional[Callable[[list], None]] = None) -> int:
        """Registers fn(events).  With start_revision, past events >= it are replayed first."""
        key, range_end = _b(key), (_b(range_end) if range_end else None)
        with self._lock:
            wid = self._next_watch
            self._next_watch += 1
            w = _Watcher(wid, key, range_end, fn or (lambda evs: None))
            if start_revision:
                past = [e for r, e in self._history if r >= start_revision and w.matches(e.kv.key)]
                if past:
                    w.fn(past)
            self._watchers[wid] = w
            return wid

    def cancel_watch(self, wid: int) -> None:
        with self._lock:
            self._watchers.pop(wid, None)


class InProcessEtcd:
    """The client interface (see etcd_client.EtcdClient) straight over an EtcdStore, no sockets."""

    def __init__(self, store: Optional[EtcdStore] = None, endpoints: Optional[list] = None):
        self.store = store or EtcdStore()
        self._endpoints = endpoints or ["inproc://etcd"]

    def endpoints(self) -> list:
        return list(self._endpoints)

    def get(self, key) -> list:
        kvs, _, _ = self.store.range(key)
        return [(kv.key.decode(), kv.value.decode()) for kv in kvs]

    def get_prefix(self, prefix) -> list:
        kvs, _, _ = self.store.range(prefix, prefix_range_end(_b(prefix)))
        return [(kv.key.decode(), kv.value.decode()) for kv in kvs]

    def put(self, key, value, lease: int = 0) -> None:
        self.store.put(key, value, lease)

    def delete(self, key) -> int:
        return self.store.delete_range(key)[0]

    def delete_prefix(self, prefix) -> int:
        return self.store.delete_range(prefix, prefix_range_end(_b(prefix)))[0]

    def put_if_absent(self, key, value, lease: int = 0) -> bool:
        ok, _, _ = self.store.txn([(key, "create", "=", 0)], [("put", key, value, lease)], [])
        return ok

    def lease_grant(self, ttl: int) -> int:
        return self.store.lease_grant(ttl)[0]

    def lease_keepalive(self, lease_id: int) -> int:
        return self.store.lease_keepalive(lease_id)

    def lease_revoke(self, lease_id: int) -> None:
        self.store.lease_revoke(lease_id)

    def revision(self) -> int:
        return self.store.revision

    def watch_prefix(self, prefix, fn, start_revision: int = 0):
        """fn(list of (type, key, value)).  Returns a cancel function."""

        def cb(events):
            fn([(e.type, e.kv.key.decode(), e.kv.value.decode()) for e in events])

        wid = self.store.watch(prefix, prefix_range_end(_b(prefix)), start_revision, cb)
        return lambda: self.store.cancel_watch(wid)

    def close(self) -> None:
        pass
