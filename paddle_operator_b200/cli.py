"""`pdj` - command line of the single-box PaddleJob runtime.

    pdj manager [flags]                 the long-running controller manager (reference: /manager, main.go)
    pdj run -f job.yaml                 one-shot: apply, reconcile + supervise in this process, wait, exit
    pdj apply|delete -f job.yaml        } client verbs against a running manager (kubectl's role in the
    pdj get pdj|pods|svc|cm|events ...  } reference: README.md:48-63, `kubectl get pdj` prints
    pdj describe pdj NAME               } NAME STATUS MODE AGE, api/v1/paddlejob_types.go:219-223)
    pdj scale pdj NAME --replicas N
    pdj logs POD
    pdj etcd --listen 127.0.0.1:2379    the etcd-compatible store (deploy/elastic/etcd.yaml's role)
    pdj client-demo                     create / get / list / delete through the API (client/client.go's role)

Manager flags are the reference's (main.go:69-83): --etcd-server, --metrics-bind-address,
--namespace, --health-probe-bind-address, --port-range, --leader-elect, --scheduling, --initImage,
and the zap logging flags (accepted; mapped onto Python logging).
"""
from __future__ import annotations

import argparse
import datetime as dt
import json
import logging
import os
import signal
import sys
import threading
import time
import urllib.error
import urllib.request
from typing import Optional

DEFAULT_SERVER = os.environ.get("PDJ_SERVER", "http://127.0.0.1:8079")
API = "/apis/batch.paddlepaddle.org/v1"


# ---------------------------------------------------------------------------
def age(ts: Optional[str]) -> str:
    if not ts:
        return "<unknown>"
    try:
        t = dt.datetime.strptime(ts, "%Y-%m-%dT%H:%M:%SZ").replace(tzinfo=dt.timezone.utc)
    except ValueError:
        return "<unknown>"
    s = int((dt.datetime.now(dt.timezone.utc) - t).total_seconds())
    if s < 120:
        return f"{max(s, 0)}s"
    if s < 7200:
        return f"{s // 60}m"
    if s < 172800:
        return f"{s // 3600}h"
    return f"{s // 86400}d"


def table(rows: list, header: list) -> str:
    rows = [header] + rows
    w = [max(len(str(r[i])) for r in rows) for i in range(len(header))]
    return "\n".join("   ".join(str(c).ljust(w[i]) for i, c in enumerate(r)).rstrip() for r in rows)


class Client:
    def __init__(self, server: str):
        self.server = server.rstrip("/")

    def req(self, method: str, path: str, body=None, raw: bool = False):
        data = json.dumps(body).encode() if body is not None else None
        r = urllib.request.Request(self.server + path, data=data, method=method,
                                   headers={"Content-Type": "application/json"})
        try:
            with urllib.request.urlopen(r, timeout=10) as resp:
                out = resp.read()
        except urllib.error.HTTPError as e:
            try:
                msg = json.loads(e.read()).get("message", str(e))
            except ValueError:
                msg = str(e)
            raise SystemExit(f"Error from server ({e.code}): {msg}")
        except urllib.error.URLError as e:
            raise SystemExit(f"cannot reach the manager at {self.server}: {e.reason} (start one with `pdj manager`)")
        return out.decode() if raw else (json.loads(out) if out.strip() else {})


RESOURCES = {"pdj": ("paddlejobs", API), "paddlejob": ("paddlejobs", API), "paddlejobs": ("paddlejobs", API),
             "pod": ("pods", "/api/v1"), "pods": ("pods", "/api/v1"), "po": ("pods", "/api/v1"),
             "svc": ("services", "/api/v1"), "service": ("services", "/api/v1"), "services": ("services", "/api/v1"),
             "cm": ("configmaps", "/api/v1"), "configmap": ("configmaps", "/api/v1"),
             "configmaps": ("configmaps", "/api/v1"), "events": ("events", "/api/v1"), "ev": ("events", "/api/v1"),
             "podgroup": ("podgroups", "/api/v1"), "podgroups": ("podgroups", "/api/v1"), "pg": ("podgroups", "/api/v1")}


def _path(res: str, ns: str, name: Optional[str] = None, sub: Optional[str] = None) -> str:
    plural, base = RESOURCES[res]
    p = f"{base}/namespaces/{ns}/{plural}"
    if name:
        p += f"/{name}"
    if sub:
        p += f"/{sub}"
    return p


def print_objects(res: str, items: list, output: str) -> None:
    import yaml

    if output == "yaml":
        print(yaml.safe_dump_all(items, sort_keys=False) if len(items) != 1 else yaml.safe_dump(items[0], sort_keys=False))
        return
    if output == "json":
        print(json.dumps(items if len(items) != 1 else items[0], indent=2))
        return
    plural = RESOURCES[res][0]
    if not items:
        print("No resources found.")
        return
    if plural == "paddlejobs":
        rows = [[o["metadata"]["name"], (o.get("status") or {}).get("phase", ""), (o.get("status") or {}).get("mode", ""),
                 age(o["metadata"].get("creationTimestamp"))] for o in items]
        hdr = ["NAME", "STATUS", "MODE", "AGE"]
        if output == "wide":
            hdr += ["PS", "WORKER", "HETER"]
            for r, o in zip(rows, items):
                for k in ("ps", "worker", "heter"):
                    st, sp = (o.get("status") or {}).get(k), (o.get("spec") or {}).get(k)
                    r.append(f"{(st or {}).get('running', 0)}/{sp['replicas']}" if sp else "-")
        print(table(rows, hdr))
    elif plural == "pods":
        rows = []
        for o in items:
            st = o.get("status") or {}
            cs = st.get("containerStatuses") or []
            ready = sum(1 for c in cs if c.get("ready"))
            total = len((o.get("spec") or {}).get("containers") or [])
            restarts = sum(c.get("restartCount", 0) for c in cs)
            status = st.get("phase", "") or "Pending"
            for c in cs:
                w = (c.get("state") or {}).get("waiting")
                if w:
                    status = w.get("reason", status)
            if status == "Pending" and any((c.get("state") or {}).get("running") for c in st.get("initContainerStatuses") or []):
                status = "Init:0/1"
            row = [o["metadata"]["name"], f"{ready}/{total}", status, restarts, age(o["metadata"].get("creationTimestamp"))]
            if output == "wide":
                row += [st.get("podIP", "<none>"), ",".join(str(g) for g in st.get("gpus") or []) or "<none>"]
            rows.append(row)
        print(table(rows, ["NAME", "READY", "STATUS", "RESTARTS", "AGE"] + (["IP", "GPUS"] if output == "wide" else [])))
    elif plural == "events":
        rows = [[age(o.get("firstTimestamp")), o.get("type", ""), o.get("reason", ""),
                 f"{o['involvedObject']['kind'].lower()}/{o['involvedObject']['name']}", o.get("message", "")] for o in items]
        print(table(rows, ["LAST SEEN", "TYPE", "REASON", "OBJECT", "MESSAGE"]))
    elif plural == "podgroups":
        rows = [[o["metadata"]["name"], (o.get("status") or {}).get("phase", "Pending"),
                 (o.get("spec") or {}).get("minMember", ""), (o.get("spec") or {}).get("queue", ""),
                 age(o["metadata"].get("creationTimestamp"))] for o in items]
        print(table(rows, ["NAME", "STATUS", "MINMEMBER", "QUEUE", "AGE"]))
    else:
        print(table([[o["metadata"]["name"], age(o["metadata"].get("creationTimestamp"))] for o in items], ["NAME", "AGE"]))


def describe(job: dict, pods: list, events: list) -> str:
    m, sp, st = job["metadata"], job.get("spec") or {}, job.get("status") or {}
    out = [f"Name:         {m['name']}", f"Namespace:    {m.get('namespace', 'default')}",
           f"API Version:  {job.get('apiVersion')}", f"Kind:         {job.get('kind')}",
           f"Annotations:  {json.dumps(m.get('annotations') or {})}", f"Finalizers:   {m.get('finalizers') or []}", "Spec:"]
    for k in ("cleanPodPolicy", "intranet", "withGloo", "elastic", "schedulingPolicy"):
        if k in sp:
            out.append(f"  {k}: {sp[k]}")
    for k in ("ps", "worker", "heter"):
        if sp.get(k):
            out.append(f"  {k}: replicas={sp[k]['replicas']}" + (f" requests={sp[k]['requests']}" if "requests" in sp[k] else "")
                       + (f" limits={sp[k]['limits']}" if "limits" in sp[k] else ""))
    out.append("Status:")
    for k in ("phase", "mode", "startTime", "completionTime", "observedGeneration"):
        if k in st:
            out.append(f"  {k}: {st[k]}")
    for k in ("ps", "worker", "heter"):
        if st.get(k):
            c = st[k]
            out.append(f"  {k}: " + " ".join(f"{n}={c.get(n, 0)}" for n in ("pending", "starting", "running", "failed",
                                                                         "succeeded")) + f" refs={len(c.get('refs') or [])}")
    out.append("Pods:")
    for p in pods:
        out.append(f"  {p['metadata']['name']}  {(p.get('status') or {}).get('phase', 'Pending')}  "
                   f"{(p.get('status') or {}).get('podIP', '')}")
    out.append("Events:")
    for e in events:
        out.append(f"  {e.get('type', ''):8} {e.get('reason', ''):10} {e.get('message', '')}")
    return "\n".join(out)


# ---------------------------------------------------------------------------
def build_cluster(args, extra_env=None):
    from .runtime.cluster import LocalCluster

    etcd = None
    if getattr(args, "etcd_server", ""):
        from .runtime.etcd_client import EtcdClient

        etcd = EtcdClient(args.etcd_server, dial_timeout=2.0)
        if not etcd.ping():
            raise SystemExit(f"etcd connect failed: {args.etcd_server}")
    gpus = None
    if getattr(args, "gpus", None) is not None:
        gpus = [int(x) for x in args.gpus.split(",") if x != ""]
    return LocalCluster(scheduling=args.scheduling, init_image=args.initImage, port_range=args.port_range, etcd=etcd,
                        namespace=args.namespace, gpus=gpus, work_root=args.work_root, extra_env=extra_env,
                        state_dir=getattr(args, "state_dir", "") or None)


def add_manager_flags(p: argparse.ArgumentParser) -> None:
    p.add_argument("--etcd-server", default="", help="The etcd server endpoints.")
    p.add_argument("--metrics-bind-address", default=":8080", help="The address the metric endpoint binds to.")
    p.add_argument("--namespace", default="", help="The namespace the controller binds to.")
    p.add_argument("--health-probe-bind-address", default=":8081", help="The address the probe endpoint binds to.")
    p.add_argument("--port-range", default="35000,65000", help="Host port range 'start,end' for intranet: Host.")
    p.add_argument("--leader-elect", action="store_true",
                   help="Enable leader election (a lock file on this box; only one active manager).")
    p.add_argument("--scheduling", default="", help="The scheduler to take, e.g. volcano")
    p.add_argument("--initImage", default="docker.io/library/busybox:1",
                   help="Image of the coordination init container; '' disables the start gate.")
    p.add_argument("--api-bind-address", default="127.0.0.1:8079", help="Address of the REST API the CLI talks to.")
    p.add_argument("--work-root", default="/tmp/pdjob", help="Where rank work directories and logs live.")
    p.add_argument("--gpus", default=None, help="Comma list of GPU indices in the pool (default: all visible).")
    p.add_argument("--state-dir", default="",
                   help="Keep PaddleJobs (spec and status) in this directory and reload them on start - what the API "
                        "server's storage does for the reference.  Empty: jobs live only as long as the manager.")
    for z in ("--zap-devel", "--zap-encoder", "--zap-log-level", "--zap-stacktrace-level"):
        p.add_argument(z, nargs="?", const="true", default=None, help="zap logging flag (accepted for compatibility)")


def cmd_manager(args) -> int:
    from .runtime.apiserver import HttpEndpoint, parse_bind

    level = {"debug": logging.DEBUG, "info": logging.INFO, "error": logging.ERROR}.get(str(args.zap_log_level).lower(),
                                                                                      logging.INFO)
    logging.basicConfig(level=level, format="%(asctime)s %(levelname)s %(name)s %(message)s")
    log = logging.getLogger("setup")
    lock = None
    if args.leader_elect:
        import fcntl

        lock = open(os.path.join(args.work_root if os.path.isdir(args.work_root) else "/tmp",
                                 "b2a304f2.paddlepaddle.org.lock"), "w")  # the reference's LeaderElectionID
        log.info("attempting to acquire leader lease")
        fcntl.flock(lock, fcntl.LOCK_EX)
        log.info("successfully acquired lease")
    cluster = build_cluster(args)
    endpoints = []
    binds = {}
    for name, addr, kw in (("api", args.api_bind_address, {"api": True}),
                           ("metrics", args.metrics_bind_address, {"metrics": True}),
                           ("health", args.health_probe_bind_address, {"health": True})):
        b = parse_bind(addr)
        if b is None:
            continue
        if b in binds:
            continue
        ep = HttpEndpoint(cluster, b, **{"api": kw.get("api", False), "health": kw.get("health", False),
                                         "metrics": kw.get("metrics", False)}).start()
        binds[b] = ep
        endpoints.append(ep)
        log.info("%s endpoint on %s", name, ep.url)
    if cluster.persister is not None:
        log.info("state directory %s: %d job(s) restored", cluster.persister.state_dir, cluster.restored)
    log.info("starting manager")
    cluster.start()
    stop = threading.Event()
    for s in (signal.SIGINT, signal.SIGTERM):
        signal.signal(s, lambda *_: stop.set())
    if args.ready_file:
        with open(args.ready_file, "w") as f:
            json.dump({"api": endpoints[0].url if endpoints else ""}, f)
    stop.wait()
    log.info("shutting down")
    cluster.stop()
    for ep in endpoints:
        ep.stop()
    return 0


def cmd_run(args) -> int:
    from .api import loader

    etcd_srv = None
    if args.etcd == "embedded":
        from .runtime.etcd_server import EtcdServer

        etcd_srv = EtcdServer("127.0.0.1", 0).start()
        args.etcd_server = etcd_srv.endpoint
    cluster = build_cluster(args)
    rc = 0
    with cluster:
        docs = loader.load_documents(args.filename)
        jobs = [d for d in docs if hasattr(d, "spec")]
        for j in jobs:
            for w in loader.warnings_for(j):
                print(f"Warning: {w}", file=sys.stderr)
        cluster.apply(args.filename)
        for j in jobs:
            print(f"paddlejob.batch.paddlepaddle.org/{j.name} created")
        try:
            for j in jobs:
                phase = cluster.wait(j.name, timeout=args.timeout, namespace=j.namespace)
                print(f"paddlejob.batch.paddlepaddle.org/{j.name} {phase}")
                if phase != "Completed":
                    rc = 1
        except TimeoutError as e:
            print(str(e), file=sys.stderr)
            rc = 2
        print_objects("pdj", cluster.store.list("PaddleJob"), "wide")
        if args.show_logs:
            for j in jobs:
                for ref in sum([(getattr(cluster.get(j.name, j.namespace).status, k) or type("x", (), {"refs": []})).refs
                                for k in ("ps", "worker", "heter")], []):
                    print(f"----- logs {ref['name']}")
                    print(cluster.logs(ref["name"], j.namespace).rstrip())
    if etcd_srv:
        etcd_srv.stop()
    return rc


def cmd_apply(args) -> int:
    from .api import loader

    c = Client(args.server)
    for d in loader.load_documents(args.filename):
        obj = d.to_dict() if hasattr(d, "to_dict") else d
        if not isinstance(obj, dict) or obj.get("kind") != "PaddleJob":
            print(f"skipped {obj.get('kind', '?')}/{(obj.get('metadata') or {}).get('name', '?')} (not a PaddleJob)")
            continue
        for w in loader.warnings_for(obj):
            print(f"Warning: {w}", file=sys.stderr)
        obj.pop("status", None)
        ns = args.namespace or obj["metadata"].get("namespace") or "default"
        name = obj["metadata"]["name"]
        try:
            cur = c.req("GET", _path("pdj", ns, name))
        except SystemExit:
            cur = None
        if cur is None:
            c.req("POST", _path("pdj", ns), obj)
            print(f"paddlejob.batch.paddlepaddle.org/{name} created")
        else:
            c.req("PATCH", _path("pdj", ns, name), {"spec": obj.get("spec", {})})
            print(f"paddlejob.batch.paddlepaddle.org/{name} configured")
    return 0


def cmd_get(args) -> int:
    c = Client(args.server)
    if args.resource not in RESOURCES:
        raise SystemExit(f"error: the server doesn't have a resource type {args.resource!r}")
    ns = args.namespace or "default"
    def fetch():
        if args.name:
            return [c.req("GET", _path(args.resource, ns, args.name))]
        return c.req("GET", _path(args.resource, ns)).get("items", [])

    items = fetch()
    print_objects(args.resource, items, args.output)
    if not getattr(args, "watch", False):
        return 0
    # -w: like `kubectl get -w`, print again whenever something observable changed (polling the API)
    def sig(objs):
        return [(o["metadata"]["name"], o["metadata"].get("resourceVersion"), json.dumps(o.get("status"), sort_keys=True))
                for o in objs]

    last = sig(items)
    deadline = time.monotonic() + args.watch_timeout if args.watch_timeout > 0 else None
    try:
        while deadline is None or time.monotonic() < deadline:
            time.sleep(args.watch_interval)
            items = fetch()
            cur = sig(items)
            if cur != last:
                print_objects(args.resource, items, args.output)
                sys.stdout.flush()
                last = cur
    except KeyboardInterrupt:
        pass
    return 0


def cmd_wait(args) -> int:
    """`pdj wait pdj NAME --for Completed[,Failed]`: block until the job reaches one of the phases
    (exit 0), or the timeout passes (exit 1) - the `kubectl wait --for=jsonpath=...` users script around."""
    c = Client(args.server)
    ns = args.namespace or "default"
    wanted = [p.strip() for p in args.for_phase.split(",") if p.strip()]
    deadline = time.monotonic() + args.timeout
    phase = ""
    while time.monotonic() < deadline:
        try:
            phase = (c.req("GET", _path("pdj", ns, args.name)).get("status") or {}).get("phase", "")
        except SystemExit:
            phase = "<deleted>"
        if phase in wanted:
            print(f"paddlejob.batch.paddlepaddle.org/{args.name} condition met: {phase}")
            return 0
        time.sleep(0.2)
    print(f"error: timed out waiting for {args.name} to be {' or '.join(wanted)} (it is {phase or 'unknown'})", file=sys.stderr)
    return 1


def cmd_describe(args) -> int:
    c = Client(args.server)
    ns = args.namespace or "default"
    job = c.req("GET", _path("pdj", ns, args.name))
    pods = [p for p in c.req("GET", _path("pods", ns)).get("items", [])
            if any(r.get("name") == args.name for r in p["metadata"].get("ownerReferences") or [])]
    events = [e for e in c.req("GET", _path("events", ns)).get("items", []) if e["involvedObject"]["name"] == args.name]
    print(describe(job, pods, events))
    return 0


def cmd_delete(args) -> int:
    from .api import loader

    c = Client(args.server)
    names = []
    if args.filename:
        names = [(d.name, d.namespace) for d in loader.load_paddlejobs(args.filename)]
    elif args.name:
        names = [(args.name, args.namespace or "default")]
    else:
        raise SystemExit("error: give a name or -f FILE")
    for name, ns in names:
        c.req("DELETE", _path("pdj", args.namespace or ns, name))
        print(f'paddlejob.batch.paddlepaddle.org "{name}" deleted')
    return 0


def cmd_scale(args) -> int:
    c = Client(args.server)
    ns = args.namespace or "default"
    c.req("PATCH", _path("pdj", ns, args.name), {"spec": {args.resource_kind: {"replicas": args.replicas}}})
    print(f"paddlejob.batch.paddlepaddle.org/{args.name} scaled")
    return 0


def cmd_logs(args) -> int:
    c = Client(args.server)
    ns = args.namespace or "default"
    text = c.req("GET", _path("pods", ns, args.pod, "log"), raw=True)
    sys.stdout.write(text)
    if not getattr(args, "follow", False):
        return 0
    # -f: keep printing what is appended until the pod has finished (polling the API)
    seen = len(text)
    try:
        while True:
            time.sleep(0.5)
            text = c.req("GET", _path("pods", ns, args.pod, "log"), raw=True)
            if len(text) > seen:
                sys.stdout.write(text[seen:])
                sys.stdout.flush()
                seen = len(text)
            phase = (c.req("GET", _path("pods", ns, args.pod)).get("status") or {}).get("phase", "")
            if phase in ("Succeeded", "Failed"):
                text = c.req("GET", _path("pods", ns, args.pod, "log"), raw=True)
                sys.stdout.write(text[seen:])
                return 0
    except KeyboardInterrupt:
        return 0


def cmd_etcd(args) -> int:
    from .runtime import etcd_server

    sys.argv = ["etcd_server", "--listen", args.listen]
    return etcd_server.main()


def cmd_crd(args) -> int:
    """The CustomResourceDefinition the reference generates (config/crd/bases/...paddlejobs.yaml), reduced to
    what is PaddleJob-specific: names, scope, printer columns, status subresource and the top-level schema
    (PodTemplateSpec is left open instead of being inlined three times)."""
    import yaml

    from .api import loader
    from .api.types import GROUP, KIND, SHORT_NAME, VERSION

    columns = [("Status", "string", ".status.phase"), ("Mode", "string", ".status.mode"),
               ("Age", "date", ".metadata.creationTimestamp")]
    names = {"kind": KIND, "listKind": KIND + "List", "plural": "paddlejobs", "singular": "paddlejob",
             "shortNames": [SHORT_NAME]}
    if getattr(args, "version", "v1") == "v1beta1":
        # the apiextensions.k8s.io/v1beta1 variant the reference ships for clusters older than 1.16
        # (deploy/v1beta1/crd.yaml, config/crd/v1beta1/): columns, subresources and the schema sit at the top of
        # the spec (`JSONPath`, `validation`), and the served version is a plain list entry
        crd = {"apiVersion": "apiextensions.k8s.io/v1beta1", "kind": "CustomResourceDefinition",
               "metadata": {"name": f"paddlejobs.{GROUP}"},
               "spec": {"additionalPrinterColumns": [{"JSONPath": p, "name": n, "type": t} for n, t, p in columns],
                        "group": GROUP, "names": names, "scope": "Namespaced", "subresources": {"status": {}},
                        "validation": {"openAPIV3Schema": loader.openapi_schema()},
                        "version": VERSION, "versions": [{"name": VERSION, "served": True, "storage": True}]}}
    else:
        crd = {"apiVersion": "apiextensions.k8s.io/v1", "kind": "CustomResourceDefinition",
               "metadata": {"name": f"paddlejobs.{GROUP}"},
               "spec": {"group": GROUP, "scope": "Namespaced", "names": names,
                        "versions": [{"name": VERSION, "served": True, "storage": True, "subresources": {"status": {}},
                                      "additionalPrinterColumns": [{"name": n, "type": t, "jsonPath": p} for n, t, p in columns],
                                      "schema": {"openAPIV3Schema": loader.openapi_schema()}}]}}
    print(yaml.safe_dump(crd, sort_keys=False))
    return 0


def cmd_client_demo(args) -> int:
    """Create, get, list, delete a PaddleJob through the API - what client/client.go of the
    reference demonstrates with controller-runtime's client ("not an SDK", client/README.MD:3)."""
    c = Client(args.server)
    tmpl = {"spec": {"containers": [{"name": "paddle", "image": "local/pdjob-b200:env-dump"}]}}
    job = {"apiVersion": "batch.paddlepaddle.org/v1", "kind": "PaddleJob", "metadata": {"name": "client-demo"},
           "spec": {"cleanPodPolicy": "Never", "worker": {"replicas": 2, "template": tmpl},
                    "ps": {"replicas": 1, "template": tmpl}}}
    c.req("POST", _path("pdj", "default"), job)
    print("created client-demo")
    got = c.req("GET", _path("pdj", "default", "client-demo"))
    print("got", got["metadata"]["name"], "mode", (got.get("status") or {}).get("mode", ""))
    time.sleep(1.0)
    items = c.req("GET", _path("pdj", "default")).get("items", [])
    print("list:", [o["metadata"]["name"] for o in items])
    c.req("DELETE", _path("pdj", "default", "client-demo"))
    print("deleted client-demo")
    return 0


def build_parser() -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser(prog="pdj", description=__doc__.split("\n\n")[0])
    sub = ap.add_subparsers(dest="cmd", required=True)

    p = sub.add_parser("manager", help="run the controller manager")
    add_manager_flags(p)
    p.add_argument("--ready-file", default="", help="write a JSON file with the API URL once serving")
    p.set_defaults(fn=cmd_manager)

    p = sub.add_parser("run", help="apply a manifest and run it to completion in this process")
    add_manager_flags(p)
    p.add_argument("-f", "--filename", required=True)
    p.add_argument("--timeout", type=float, default=3600.0)
    p.add_argument("--etcd", default="", choices=["", "embedded"], help="start an embedded etcd-compatible store")
    p.add_argument("--show-logs", action="store_true")
    p.set_defaults(fn=cmd_run)

    def client_parser(name, help_):
        q = sub.add_parser(name, help=help_)
        q.add_argument("--server", default=DEFAULT_SERVER)
        q.add_argument("-n", "--namespace", default="")
        return q

    p = client_parser("apply", "create or update PaddleJobs from a file")
    p.add_argument("-f", "--filename", required=True)
    p.set_defaults(fn=cmd_apply)
    p = client_parser("get", "list resources")
    p.add_argument("resource")
    p.add_argument("name", nargs="?")
    p.add_argument("-o", "--output", default="", choices=["", "wide", "yaml", "json"])
    p.add_argument("-w", "--watch", action="store_true", help="keep printing when something changes")
    p.add_argument("--watch-interval", type=float, default=0.5, help=argparse.SUPPRESS)
    p.add_argument("--watch-timeout", type=float, default=0.0, help="stop watching after this many seconds (0: never)")
    p.set_defaults(fn=cmd_get)
    p = client_parser("wait", "wait until a PaddleJob reaches a phase")
    p.add_argument("resource", choices=["pdj", "paddlejob", "paddlejobs"])
    p.add_argument("name")
    p.add_argument("--for", dest="for_phase", default="Completed,Failed", help="comma list of phases (default Completed,Failed)")
    p.add_argument("--timeout", type=float, default=300.0)
    p.set_defaults(fn=cmd_wait)
    p = client_parser("describe", "show a PaddleJob with its pods and events")
    p.add_argument("resource", choices=["pdj", "paddlejob", "paddlejobs"])
    p.add_argument("name")
    p.set_defaults(fn=cmd_describe)
    p = client_parser("delete", "delete PaddleJobs")
    p.add_argument("resource", nargs="?", choices=["pdj", "paddlejob", "paddlejobs"])
    p.add_argument("name", nargs="?")
    p.add_argument("-f", "--filename", default="")
    p.set_defaults(fn=cmd_delete)
    p = client_parser("scale", "change the replicas of one resource kind")
    p.add_argument("resource", choices=["pdj", "paddlejob", "paddlejobs"])
    p.add_argument("name")
    p.add_argument("--replicas", type=int, required=True)
    p.add_argument("--resource-kind", default="worker", choices=["ps", "worker", "heter"])
    p.set_defaults(fn=cmd_scale)
    p = client_parser("logs", "print a rank's log")
    p.add_argument("pod")
    p.add_argument("-f", "--follow", action="store_true", help="keep printing until the rank has finished")
    p.set_defaults(fn=cmd_logs)
    p = client_parser("events", "list events")
    p.set_defaults(fn=lambda a: cmd_get(argparse.Namespace(server=a.server, namespace=a.namespace, resource="events",
                                                           name=None, output="")))
    p = client_parser("client-demo", "CRUD round trip through the API")
    p.set_defaults(fn=cmd_client_demo)
    p = sub.add_parser("crd", help="print the PaddleJob schema (the generated CRD's top level) as YAML")
    p.add_argument("--version", default="v1", choices=["v1", "v1beta1"],
                   help="apiextensions.k8s.io version of the CustomResourceDefinition document (reference: deploy/v1 and deploy/v1beta1)")
    p.set_defaults(fn=cmd_crd)
    p = sub.add_parser("etcd", help="run the etcd-compatible store")
    p.add_argument("--listen", default="127.0.0.1:2379")
    p.set_defaults(fn=cmd_etcd)

    return ap


def main(argv=None) -> int:
    args = build_parser().parse_args(argv)
    return args.fn(args)


if __name__ == "__main__":
    sys.exit(main())
