"""etcd v3 JSON gateway over an EtcdStore.

Speaks the grpc-gateway dialect that a real etcd serves next to gRPC: POST with
JSON bodies, byte strings base64-encoded, 64-bit integers as decimal strings.
Endpoints: /v3/kv/range, /v3/kv/put, /v3/kv/deleterange, /v3/kv/txn,
/v3/lease/grant, /v3/lease/keepalive, /v3/lease/revoke (and the older
/v3/kv/lease/revoke), /v3/lease/timetolive, /v3/watch (a chunked stream of
newline-delimited JSON), /v3/maintenance/status, /version, /health.

`python -m paddle_operator_b200.runtime.etcd_server --listen 127.0.0.1:2379`
is the stand-in for deploy/elastic/etcd.yaml of the reference (a single-replica
etcd behind the Service `paddle-elastic-etcd:2379`).
"""
from __future__ import annotations

import argparse
import base64
import json
import queue
import threading
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from typing import Optional

from .etcd_store import EtcdStore, KeyValue

CLUSTER_ID = "14841639068965178418"
MEMBER_ID = "10276657743932975437"


def b64e(b: bytes) -> str:
    return base64.b64encode(b).decode()


def b64d(s: Optional[str]) -> bytes:
    return base64.b64decode(s) if s else b""


def _kv_json(kv: KeyValue) -> dict:
    d = {"key": b64e(kv.key), "create_revision": str(kv.create_revision), "mod_revision": str(kv.mod_revision),
         "version": str(kv.version)}
    if kv.value:
        d["value"] = b64e(kv.value)
    if kv.lease:
        d["lease"] = str(kv.lease)
    return d


def _header(store: EtcdStore, rev: Optional[int] = None) -> dict:
    return {"cluster_id": CLUSTER_ID, "member_id": MEMBER_ID, "revision": str(rev if rev is not None else store.revision),
            "raft_term": "2"}


_CMP_TARGET = {"VERSION": "version", "CREATE": "create", "MOD": "mod", "VALUE": "value", "LEASE": "lease",
               0: "version", 1: "create", 2: "mod", 3: "value", 4: "lease"}
_CMP_RESULT = {"EQUAL": "=", "GREATER": ">", "LESS": "<", "NOT_EQUAL": "!=", 0: "=", 1: ">", 2: "<", 3: "!="}


class _Handler(BaseHTTPRequestHandler):
    protocol_version = "HTTP/1.1"
    store: EtcdStore = None  # set by serve()

    def log_message(self, fmt, *args):  # quiet
        pass

    def _body(self) -> dict:
        n = int(self.headers.get("Content-Length") or 0)
        raw = self.rfile.read(n) if n else b""
        return json.loads(raw) if raw.strip() else {}

    def _send(self, obj: dict, code: int = 200) -> None:
        data = json.dumps(obj).encode()
        self.send_response(code)
        self.send_header("Content-Type", "application/json")
        self.send_header("Content-Length", str(len(data)))
        self.end_headers()
        self.wfile.write(data)

    def _error(self, msg: str, code: int = 400, grpc_code: int = 3) -> None:
        self._send({"error": msg, "code": grpc_code, "message": msg}, code)

    def do_GET(self):
        if self.path == "/version":
            self._send({"etcdserver": "3.5.0-pdjob", "etcdcluster": "3.5.0"})
        elif self.path == "/health":
            self._send({"health": "true"})
        else:
            self._error("not found", 404, 5)

    def do_POST(self):
        s = self.store
        path = self.path.split("?")[0]
        try:
            body = self._body()
            if path == "/v3/kv/range":
                kvs, count, rev = s.range(b64d(body.get("key")), b64d(body.get("range_end")) or None,
                                          int(body.get("limit", 0) or 0), bool(body.get("keys_only")),
                                          bool(body.get("count_only")))
                out = {"header": _header(s, rev)}
                if kvs:
                    out["kvs"] = [_kv_json(k) for k in kvs]
                if count:
                    out["count"] = str(count)
                self._send(out)
            elif path == "/v3/kv/put":
                prev, rev = s.put(b64d(body.get("key")), b64d(body.get("value")), int(body.get("lease", 0) or 0),
                                  bool(body.get("prev_kv")))
                out = {"header": _header(s, rev)}
                if prev is not None:
                    out["prev_kv"] = _kv_json(prev)
                self._send(out)
            elif path == "/v3/kv/deleterange":
                n, prevs, rev = s.delete_range(b64d(body.get("key")), b64d(body.get("range_end")) or None,
                                               bool(body.get("prev_kv")))
                out = {"header": _header(s, rev)}
                if n:
                    out["deleted"] = str(n)
                if prevs:
                    out["prev_kvs"] = [_kv_json(k) for k in prevs]
                self._send(out)
            elif path == "/v3/kv/txn":
                self._txn(body)
            elif path == "/v3/lease/grant":
                lid, ttl = s.lease_grant(int(body.get("TTL", 0) or 0), int(body.get("ID", 0) or 0))
                self._send({"header": _header(s), "ID": str(lid), "TTL": str(ttl)})
            elif path == "/v3/lease/keepalive":
                lid = int(body.get("ID", 0) or 0)
                ttl = s.lease_keepalive(lid)
                res = {"header": _header(s), "ID": str(lid)}
                if ttl:
                    res["TTL"] = str(ttl)
                self._send({"result": res})
            elif path in ("/v3/lease/revoke", "/v3/kv/lease/revoke"):
                if not s.lease_revoke(int(body.get("ID", 0) or 0)):
                    return self._error("etcdserver: requested lease not found", 404, 5)
                self._send({"header": _header(s)})
            elif path in ("/v3/lease/timetolive", "/v3/kv/lease/timetolive"):
                lid = int(body.get("ID", 0) or 0)
                rem, ttl, keys = s.lease_ttl(lid)
                out = {"header": _header(s), "ID": str(lid), "TTL": str(rem)}
                if rem >= 0:
                    out["grantedTTL"] = str(ttl)
                    if body.get("keys"):
                        out["keys"] = [b64e(k) for k in keys]
                self._send(out)
            elif path == "/v3/maintenance/status":
                self._send({"header": _header(s), "version": "3.5.0-pdjob", "leader": MEMBER_ID, "raftTerm": "2"})
            elif path == "/v3/watch":
                self._watch(body)
            else:
                self._error("not found", 404, 5)
        except KeyError as e:
            self._error(str(e).strip("'\""), 404, 5)
        except (ValueError, TypeError) as e:
            self._error(f"bad request: {e}")
        except (BrokenPipeError, ConnectionResetError):
            pass

    def _txn(self, body: dict) -> None:
        s = self.store
        compare = []
        for c in body.get("compare") or []:
            target = _CMP_TARGET[c.get("target", "VERSION")]
            op = _CMP_RESULT[c.get("result", "EQUAL")]
            if target == "value":
                val = b64d(c.get("value"))
            else:
                field = {"version": "version", "create": "create_revision", "mod": "mod_revision", "lease": "lease"}[target]
                val = int(c.get(field, 0) or 0)
            compare.append((b64d(c.get("key")), target, op, val))

        def ops(lst):
            out = []
            for o in lst or []:
                if "request_put" in o or "requestPut" in o:
                    r = o.get("request_put") or o.get("requestPut")
                    out.append(("put", b64d(r.get("key")), b64d(r.get("value")), int(r.get("lease", 0) or 0)))
                elif "request_range" in o or "requestRange" in o:
                    r = o.get("request_range") or o.get("requestRange")
                    out.append(("range", b64d(r.get("key")), b64d(r.get("range_end")) or None))
                elif "request_delete_range" in o or "requestDeleteRange" in o:
                    r = o.get("request_delete_range") or o.get("requestDeleteRange")
                    out.append(("delete", b64d(r.get("key")), b64d(r.get("range_end")) or None))
                else:
                    raise ValueError("unsupported txn request")
            return out

        ok, results, rev = s.txn(compare, ops(body.get("success")), ops(body.get("failure")))
        responses = []
        for kind, val in results:
            if kind == "put":
                responses.append({"response_put": {"header": _header(s, rev)}})
            elif kind == "range":
                r = {"header": _header(s, rev)}
                if val:
                    r["kvs"] = [_kv_json(k) for k in val]
                    r["count"] = str(len(val))
                responses.append({"response_range": r})
            else:
                r = {"header": _header(s, rev)}
                if val:
                    r["deleted"] = str(val)
                responses.append({"response_delete_range": r})
        out = {"header": _header(s, rev)}
        if ok:
            out["succeeded"] = True
        if responses:
            out["responses"] = responses
        self._send(out)

    def _watch(self, body: dict) -> None:
        s = self.store
        cr = body.get("create_request") or body.get("createRequest") or {}
        q: "queue.Queue[list]" = queue.Queue()
        self.send_response(200)
        self.send_header("Content-Type", "application/json")
        self.send_header("Transfer-Encoding", "chunked")
        self.end_headers()

        def chunk(obj: dict) -> None:
            data = (json.dumps(obj) + "\n").encode()
            self.wfile.write(f"{len(data):x}\r\n".encode() + data + b"\r\n")
            self.wfile.flush()

        wid = s.watch(b64d(cr.get("key")), b64d(cr.get("range_end")) or None, int(cr.get("start_revision", 0) or 0),
                      q.put)
        try:
            chunk({"result": {"header": _header(s), "created": True, "watch_id": str(wid)}})
            while True:
                try:
                    events = q.get(timeout=1.0)
                except queue.Empty:
                    # an empty progress notification doubles as a liveness probe of the connection
                    chunk({"result": {"header": _header(s), "watch_id": str(wid)}})
                    continue
                evs = []
                for e in events:
                    d = {"kv": _kv_json(e.kv)}
                    if e.type == "DELETE":
                        d["type"] = "DELETE"
                    if e.prev_kv is not None and cr.get("prev_kv"):
                        d["prev_kv"] = _kv_json(e.prev_kv)
                    evs.append(d)
                chunk({"result": {"header": _header(s, events[-1].kv.mod_revision), "watch_id": str(wid), "events": evs}})
        except (BrokenPipeError, ConnectionResetError, OSError):
            pass
        finally:
            s.cancel_watch(wid)
            self.close_connection = True


class EtcdServer:
    def __init__(self, host: str = "127.0.0.1", port: int = 2379, store: Optional[EtcdStore] = None):
        self.store = store or EtcdStore()
        handler = type("Handler", (_Handler,), {"store": self.store})
        # eight ranks keep-alive, watch and report at once: the default backlog of 5 resets connections
        server_cls = type("Server", (ThreadingHTTPServer,), {"request_queue_size": 256})
        self.httpd = server_cls((host, port), handler)
        self.httpd.daemon_threads = True
        self.host, self.port = self.httpd.server_address[:2]
        self._thread: Optional[threading.Thread] = None

    @property
    def endpoint(self) -> str:
        return f"{self.host}:{self.port}"

    def start(self) -> "EtcdServer":
        self._thread = threading.Thread(target=self.httpd.serve_forever, name="etcd-gateway", daemon=True)
        self._thread.start()
        return self

    def stop(self) -> None:
        self.httpd.shutdown()
        self.httpd.server_close()
        self.store.close()

    def __enter__(self):
        return self.start()

    def __exit__(self, *a):
        self.stop()


def main() -> int:
    ap = argparse.ArgumentParser(description="etcd-v3-compatible store (JSON gateway)")
    ap.add_argument("--listen", default="127.0.0.1:2379")
    args = ap.parse_args()
    host, port = args.listen.rsplit(":", 1)
    srv = EtcdServer(host, int(port))
    print(f"etcd-compatible store listening on {srv.endpoint}", flush=True)
    try:
        srv.httpd.serve_forever()
    except KeyboardInterrupt:
        pass
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
