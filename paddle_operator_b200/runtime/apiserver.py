"""HTTP face of the manager process: a Kubernetes-shaped REST API over the object
store (what `kubectl` talks to in the reference's world), plus the manager's
health and metrics endpoints.

  /apis/batch.paddlepaddle.org/v1/namespaces/{ns}/paddlejobs[/{name}[/status|/scale]]
  /api/v1/namespaces/{ns}/{pods|services|configmaps|events|podgroups}[/{name}]
  /api/v1/namespaces/{ns}/pods/{name}/log
  /healthz  /readyz          (main.go:158-165 of the reference: healthz.Ping on :8081)
  /metrics                   (main.go:70,90: controller-runtime's registry on :8080)

Verbs: GET (one / list), POST (create), PUT (replace spec, with resourceVersion
conflict detection), PATCH (JSON merge of .spec), DELETE.
"""
from __future__ import annotations

import json
import threading
import time
from http.server import BaseHTTPRequestHandler, ThreadingHTTPServer
from typing import Optional

from ..api import loader
from ..api.types import GROUP, KIND
from ..controller.store import AlreadyExists, Conflict, NotFound, ObjectStore, Invalid

PLURALS = {"paddlejobs": KIND, "pdj": KIND, "pods": "Pod", "services": "Service", "configmaps": "ConfigMap",
           "events": "Event", "podgroups": "PodGroup"}


def _merge(dst: dict, patch: dict) -> dict:
    for k, v in patch.items():
        if v is None:
            dst.pop(k, None)
        elif isinstance(v, dict) and isinstance(dst.get(k), dict):
            _merge(dst[k], v)
        else:
            dst[k] = v
    return dst


class Metrics:
    def __init__(self, cluster):
        self.cluster = cluster
        self.started = time.time()

    def render(self) -> str:
        c = self.cluster
        lines = [
            "# HELP controller_runtime_reconcile_total Total number of reconciliations per controller",
            "# TYPE controller_runtime_reconcile_total counter",
            f'controller_runtime_reconcile_total{{controller="paddlejob"}} {c.manager.reconcile_count}',
            "# HELP controller_runtime_reconcile_errors_total Total number of reconciliation errors per controller",
            "# TYPE controller_runtime_reconcile_errors_total counter",
            f'controller_runtime_reconcile_errors_total{{controller="paddlejob"}} {c.manager.reconcile_errors}',
            "# HELP workqueue_depth Current depth of workqueue",
            "# TYPE workqueue_depth gauge",
            f'workqueue_depth{{name="paddlejob"}} {c.manager.pending()}',
            "# HELP paddlejob_jobs Number of PaddleJobs by phase",
            "# TYPE paddlejob_jobs gauge",
        ]
        phases: dict = {}
        for o in c.store.list(KIND):
            ph = (o.get("status") or {}).get("phase", "") or "Unknown"
            phases[ph] = phases.get(ph, 0) + 1
        for ph, n in sorted(phases.items()):
            lines.append(f'paddlejob_jobs{{phase="{ph}"}} {n}')
        pods: dict = {}
        for o in c.store.list("Pod"):
            ph = (o.get("status") or {}).get("phase", "") or "Unknown"
            pods[ph] = pods.get(ph, 0) + 1
        lines += ["# HELP paddlejob_ranks Number of rank processes (pods) by phase", "# TYPE paddlejob_ranks gauge"]
        for ph, n in sorted(pods.items()):
            lines.append(f'paddlejob_ranks{{phase="{ph}"}} {n}')
        if c.supervisor is not None:
            lines += ["# HELP paddlejob_gpus_total GPUs in the local pool", "# TYPE paddlejob_gpus_total gauge",
                      f"paddlejob_gpus_total {len(c.supervisor.gpu_pool)}",
                      "# HELP paddlejob_gpus_allocated GPUs assigned to ranks", "# TYPE paddlejob_gpus_allocated gauge",
                      f"paddlejob_gpus_allocated {len(c.supervisor._gpu_owner)}"]
        lines += ["# HELP process_uptime_seconds Seconds since the manager started", "# TYPE process_uptime_seconds gauge",
                  f"process_uptime_seconds {time.time() - self.started:.1f}"]
        return "\n".join(lines) + "\n"


def make_handler(cluster, serve_api: bool, serve_health: bool, serve_metrics: bool):
    store: ObjectStore = cluster.store
    metrics = Metrics(cluster)

    class Handler(BaseHTTPRequestHandler):
        protocol_version = "HTTP/1.1"

        def log_message(self, fmt, *args):
            pass

        def _send(self, code: int, body, ctype: str = "application/json") -> None:
            data = body.encode() if isinstance(body, str) else json.dumps(body).encode()
            self.send_response(code)
            self.send_header("Content-Type", ctype)
            self.send_header("Content-Length", str(len(data)))
            self.end_headers()
            self.wfile.write(data)

        def _status(self, code: int, reason: str, message: str) -> None:
            self._send(code, {"kind": "Status", "apiVersion": "v1", "status": "Failure", "message": message,
                              "reason": reason, "code": code})

        def _body(self) -> dict:
            n = int(self.headers.get("Content-Length") or 0)
            raw = self.rfile.read(n) if n else b""
            return json.loads(raw) if raw.strip() else {}

        def _route(self):
            """-> (kind, namespace|None, name|None, subresource|None) or None"""
            parts = [p for p in self.path.split("?")[0].split("/") if p]
            if parts[:2] == ["apis", GROUP] and len(parts) >= 3:
                rest = parts[3:]
            elif parts[:2] == ["api", "v1"]:
                rest = parts[2:]
            elif parts[:2] == ["apis", "scheduling.volcano.sh"] and len(parts) >= 3:
                rest = parts[3:]
            else:
                return None
            ns = None
            if rest[:1] == ["namespaces"] and len(rest) >= 2:
                ns, rest = rest[1], rest[2:]
            if not rest or rest[0] not in PLURALS:
                return None
            return PLURALS[rest[0]], ns, (rest[1] if len(rest) > 1 else None), (rest[2] if len(rest) > 2 else None)

        def do_GET(self):
            path = self.path.split("?")[0]
            if path in ("/healthz", "/readyz") and serve_health:
                return self._send(200, "ok", "text/plain")
            if path == "/metrics" and serve_metrics:
                return self._send(200, metrics.render(), "text/plain; version=0.0.4")
            if not serve_api:
                return self._status(404, "NotFound", "not found")
            r = self._route()
            if r is None:
                return self._status(404, "NotFound", f"no route for {path}")
            kind, ns, name, sub = r
            try:
                if name is None:
                    items = store.list(kind, ns)
                    return self._send(200, {"kind": kind + "List", "apiVersion": "v1", "items": items})
                if kind == "Pod" and sub == "log":
                    return self._send(200, cluster.logs(name, ns or "default"), "text/plain")
                obj = store.get(kind, ns or "default", name)
                if sub == "status":
                    return self._send(200, obj)
                return self._send(200, obj)
            except NotFound as e:
                return self._status(404, "NotFound", str(e))

        def do_POST(self):
            if not serve_api:
                return self._status(404, "NotFound", "not found")
            r = self._route()
            if r is None:
                return self._status(404, "NotFound", "no route")
            kind, ns, _, _ = r
            try:
                body = self._body()
                body.setdefault("metadata", {}).setdefault("namespace", ns or "default")
                if kind == KIND:
                    job = loader.paddlejob_from_doc(body)  # schema validation, pruning
                    body = job.to_dict()
                    body.pop("status", None)
                body["kind"] = kind
                return self._send(201, store.create(body))
            except (loader.ValidationError, Invalid) as e:
                return self._status(422, "Invalid", str(e))
            except AlreadyExists as e:
                return self._status(409, "AlreadyExists", str(e))
            except (ValueError, KeyError) as e:
                return self._status(400, "BadRequest", str(e))

        def do_PUT(self):
            if not serve_api:
                return self._status(404, "NotFound", "not found")
            r = self._route()
            if r is None or r[2] is None:
                return self._status(404, "NotFound", "no route")
            kind, ns, name, sub = r
            try:
                body = self._body()
                body.setdefault("metadata", {}).update({"namespace": ns or "default", "name": name})
                body["kind"] = kind
                if kind == KIND and sub != "status":
                    loader.paddlejob_from_doc(body)
                out = store.update_status(body) if sub == "status" else store.update(body)
                return self._send(200, out)
            except loader.ValidationError as e:
                return self._status(422, "Invalid", str(e))
            except Conflict as e:
                return self._status(409, "Conflict", str(e))
            except NotFound as e:
                return self._status(404, "NotFound", str(e))

        def do_PATCH(self):
            if not serve_api:
                return self._status(404, "NotFound", "not found")
            r = self._route()
            if r is None or r[2] is None:
                return self._status(404, "NotFound", "no route")
            kind, ns, name, _ = r
            try:
                patch = self._body()
                for _ in range(10):
                    obj = store.get(kind, ns or "default", name)
                    _merge(obj, {k: v for k, v in patch.items() if k in ("spec", "metadata")})
                    obj["metadata"]["name"], obj["metadata"]["namespace"] = name, ns or "default"
                    if kind == KIND:
                        loader.paddlejob_from_doc(obj)
                    try:
                        return self._send(200, store.update(obj))
                    except Conflict:
                        continue
                return self._status(409, "Conflict", "too many conflicts")
            except loader.ValidationError as e:
                return self._status(422, "Invalid", str(e))
            except NotFound as e:
                return self._status(404, "NotFound", str(e))

        def do_DELETE(self):
            if not serve_api:
                return self._status(404, "NotFound", "not found")
            r = self._route()
            if r is None or r[2] is None:
                return self._status(404, "NotFound", "no route")
            kind, ns, name, _ = r
            try:
                store.delete(kind, ns or "default", name)
                return self._send(200, {"kind": "Status", "apiVersion": "v1", "status": "Success"})
            except NotFound as e:
                return self._status(404, "NotFound", str(e))

    return Handler


def parse_bind(addr: str, default_host: str = "0.0.0.0") -> tuple:
    """':8080' -> ('0.0.0.0', 8080); '127.0.0.1:0' -> ('127.0.0.1', 0); '0' -> disabled (None)."""
    if addr in ("0", "", None):
        return None
    host, _, port = addr.rpartition(":")
    return (host or default_host, int(port))


class HttpEndpoint:
    def __init__(self, cluster, bind: tuple, api: bool = False, health: bool = False, metrics: bool = False):
        self.httpd = ThreadingHTTPServer(bind, make_handler(cluster, api, health, metrics))
        self.httpd.daemon_threads = True
        self.address = self.httpd.server_address[:2]
        self._t: Optional[threading.Thread] = None

    @property
    def url(self) -> str:
        host = "127.0.0.1" if self.address[0] in ("0.0.0.0", "") else self.address[0]
        return f"http://{host}:{self.address[1]}"

    def start(self) -> "HttpEndpoint":
        self._t = threading.Thread(target=self.httpd.serve_forever, daemon=True)
        self._t.start()
        return self

    def stop(self) -> None:
        self.httpd.shutdown()
        self.httpd.server_close()
