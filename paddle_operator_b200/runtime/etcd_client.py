"""Client for etcd's v3 JSON gateway (works against runtime.etcd_server and
against a real etcd >= 3.4).  Interface shared with etcd_store.InProcessEtcd:

    get(key) / get_prefix(p) -> [(key, value)]      put(key, value, lease=0)
    delete(key) / delete_prefix(p)                  put_if_absent(key, value, lease=0)
    lease_grant(ttl) / lease_keepalive(id) / lease_revoke(id)
    watch_prefix(prefix, fn, start_revision=0) -> cancel()
    endpoints()

The reference's manager dials etcd with a 2 s timeout (main.go:117-120) and the
np sync uses a 3 s deadline (controllers/paddlejob_elastic.go:42); the default
timeouts here match.
"""
from __future__ import annotations

import base64
import http.client
import json
import threading
import time
from typing import Callable, Optional

from .etcd_store import prefix_range_end


def _e(x) -> str:
    return base64.b64encode(x if isinstance(x, bytes) else str(x).encode()).decode()


def _d(s: Optional[str]) -> str:
    return base64.b64decode(s).decode() if s else ""


class EtcdError(RuntimeError):
    pass


class EtcdClient:
    def __init__(self, endpoints, dial_timeout: float = 2.0, op_timeout: float = 3.0):
        if isinstance(endpoints, str):
            endpoints = [e for e in endpoints.split(",") if e]
        self._endpoints = [e.replace("http://", "").rstrip("/") for e in endpoints]
        if not self._endpoints:
            raise ValueError("no etcd endpoints")
        self.dial_timeout, self.op_timeout = dial_timeout, op_timeout
        self._watch_threads: list = []

    def endpoints(self) -> list:
        return list(self._endpoints)

    def _post(self, path: str, body: dict, retries: int = 3) -> dict:
        last: Optional[Exception] = None
        for attempt in range(retries):
            try:
                return self._post_once(path, body)
            except EtcdError as e:
                if "no etcd endpoint reachable" not in str(e):
                    raise
                last = e
                time.sleep(0.05 * (attempt + 1))
        raise last  # type: ignore[misc]

    def _post_once(self, path: str, body: dict) -> dict:
        last: Optional[Exception] = None
        for ep in self._endpoints:
            host, _, port = ep.partition(":")
            try:
                conn = http.client.HTTPConnection(host, int(port or 2379), timeout=self.op_timeout)
                conn.request("POST", path, json.dumps(body), {"Content-Type": "application/json"})
                resp = conn.getresponse()
                data = resp.read()
                conn.close()
                out = json.loads(data) if data.strip() else {}
                if resp.status >= 400:
                    raise EtcdError(out.get("error") or out.get("message") or f"HTTP {resp.status}")
                return out
            except (OSError, http.client.HTTPException) as e:
                last = e
        raise EtcdError(f"no etcd endpoint reachable: {last}")

    def ping(self) -> bool:
        try:
            self._post("/v3/maintenance/status", {})
            return True
        except EtcdError:
            return False

    # -- kv -----------------------------------------------------------------------
    def _range(self, key, range_end=None) -> list:
        body = {"key": _e(key)}
        if range_end:
            body["range_end"] = _e(range_end)
        out = self._post("/v3/kv/range", body)
        return [(_d(kv.get("key")), _d(kv.get("value"))) for kv in out.get("kvs", [])]

    def get(self, key) -> list:
        return self._range(key)

    def get_prefix(self, prefix) -> list:
        return self._range(prefix, prefix_range_end(prefix.encode() if isinstance(prefix, str) else prefix))

    def put(self, key, value, lease: int = 0) -> None:
        body = {"key": _e(key), "value": _e(value)}
        if lease:
            body["lease"] = str(lease)
        self._post("/v3/kv/put", body)

    def delete(self, key) -> int:
        return int(self._post("/v3/kv/deleterange", {"key": _e(key)}).get("deleted", 0))

    def delete_prefix(self, prefix) -> int:
        p = prefix.encode() if isinstance(prefix, str) else prefix
        return int(self._post("/v3/kv/deleterange", {"key": _e(p), "range_end": _e(prefix_range_end(p))}).get("deleted", 0))

    def put_if_absent(self, key, value, lease: int = 0) -> bool:
        put = {"key": _e(key), "value": _e(value)}
        if lease:
            put["lease"] = str(lease)
        out = self._post("/v3/kv/txn", {"compare": [{"key": _e(key), "target": "CREATE", "result": "EQUAL",
                                                     "create_revision": "0"}],
                                        "success": [{"request_put": put}]})
        return bool(out.get("succeeded"))

    def revision(self) -> int:
        return int(self._post("/v3/kv/range", {"key": _e("\0"), "count_only": True})["header"]["revision"])

    # -- leases ---------------------------------------------------------------------
    def lease_grant(self, ttl: int) -> int:
        return int(self._post("/v3/lease/grant", {"TTL": str(int(ttl))})["ID"])

    def lease_keepalive(self, lease_id: int) -> int:
        out = self._post("/v3/lease/keepalive", {"ID": str(lease_id)})
        return int((out.get("result") or {}).get("TTL", 0))

    def lease_revoke(self, lease_id: int) -> None:
        try:
            self._post("/v3/lease/revoke", {"ID": str(lease_id)})
        except EtcdError:
            pass

    # -- watch ------------------------------------------------------------------------
    def watch_prefix(self, prefix, fn: Callable[[list], None], start_revision: int = 0):
        """Calls fn([(type, key, value), ...]) from a background thread.  Returns cancel()."""
        p = prefix.encode() if isinstance(prefix, str) else prefix
        stop = threading.Event()
        ready = threading.Event()
        holder: dict = {}

        def run():
            host, _, port = self._endpoints[0].partition(":")
            try:
                conn = http.client.HTTPConnection(host, int(port or 2379), timeout=5.0)
                cr = {"key": _e(p), "range_end": _e(prefix_range_end(p))}
                if start_revision:
                    cr["start_revision"] = str(start_revision)
                conn.request("POST", "/v3/watch", json.dumps({"create_request": cr}),
                             {"Content-Type": "application/json"})
                resp = conn.getresponse()
                holder["conn"] = conn
                while not stop.is_set():
                    line = resp.readline()
                    if not line:
                        break
                    line = line.strip()
                    if not line:
                        continue
                    res = (json.loads(line).get("result") or {})
                    if res.get("created"):
                        ready.set()
                    evs = [(ev.get("type", "PUT"), _d(ev["kv"].get("key")), _d(ev["kv"].get("value")))
                           for ev in res.get("events", [])]
                    if evs and not stop.is_set():
                        fn(evs)
            except (OSError, http.client.HTTPException, ValueError):
                pass
            finally:
                ready.set()

        th = threading.Thread(target=run, name="etcd-watch", daemon=True)
        th.start()
        ready.wait(self.dial_timeout)

        def cancel():
            stop.set()
            c = holder.get("conn")
            if c is not None:
                try:
                    c.sock and c.sock.close()
                except OSError:
                    pass

        return cancel

    def close(self) -> None:
        pass
