#!/usr/bin/env python3
"""Installs (or, with --uninstall, removes) the manager and the etcd-compatible store as systemd services -
the single-box counterpart of `kubectl apply -f deploy/v1/crd.yaml -f deploy/v1/operator.yaml` / `helm install
paddle-operator charts/paddle-operator` in the reference.  Values come from deploy/values.env (the chart's
values.yaml); override any of them with KEY=VALUE arguments:

    deploy/install.py SCHEDULING=volcano GPUS=0,1,2,3              # render, install, enable, start
    deploy/install.py --dry-run --out /tmp/units NAMESPACE=team-a   # only render the unit files
    deploy/install.py --uninstall
"""
import argparse
import os
import shlex
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
UNITS = ("pdjob-etcd.service", "pdjob-manager.service")


def load_values(path: str) -> dict:
    vals = {}
    for line in open(path):
        line = line.strip()
        if not line or line.startswith("#") or "=" not in line:
            continue
        k, v = line.split("=", 1)
        vals[k.strip()] = v.strip().strip('"')
    return vals


def render(v: dict) -> dict:
    flags = ["--leader-elect", "--etcd-server", v["ETCD_LISTEN"], "--metrics-bind-address", v["METRICS_BIND"],
             "--health-probe-bind-address", v["HEALTH_BIND"], "--api-bind-address", v["API_BIND"],
             "--port-range", v["PORT_RANGE"], "--work-root", v["WORK_ROOT"], "--state-dir", v["STATE_DIR"],
             "--initImage", v["INIT_IMAGE"]]
    for key, flag in (("NAMESPACE", "--namespace"), ("SCHEDULING", "--scheduling"), ("GPUS", "--gpus")):
        if v.get(key):
            flags += [flag, v[key]]
    py = v["PYTHON"]
    etcd = f"""# rendered by deploy/install.py - the etcd-compatible store for elastic jobs (reference: deploy/elastic/etcd.yaml)
[Unit]
Description=etcd-compatible store for elastic PaddleJobs

[Service]
WorkingDirectory={v['REPO_DIR']}
ExecStart={py} -m paddle_operator_b200.cli etcd --listen {v['ETCD_LISTEN']}
Restart=on-failure

[Install]
WantedBy=multi-user.target
"""
    manager = f"""# rendered by deploy/install.py - the controller manager (reference: the Deployment of deploy/v1/operator.yaml:
# args --leader-elect, probes on :8081, metrics on :8080, limits 100m CPU / 300Mi)
[Unit]
Description=PaddleJob controller manager (b200-paddlejob-runtime)
After=network.target pdjob-etcd.service

[Service]
WorkingDirectory={v['REPO_DIR']}
ExecStartPre={py} -c "import __graft_entry__ as g; g.build()"
ExecStart={py} -m paddle_operator_b200.cli manager {' '.join(shlex.quote(f) for f in flags)}
Restart=on-failure
MemoryMax=300M
CPUQuota=100%

[Install]
WantedBy=multi-user.target
"""
    return {"pdjob-etcd.service": etcd, "pdjob-manager.service": manager}


def main() -> int:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--dry-run", action="store_true", help="render only; do not touch systemd")
    ap.add_argument("--uninstall", action="store_true")
    ap.add_argument("--out", default="", help="directory for the unit files (default: UNIT_DIR)")
    ap.add_argument("--values", default=os.path.join(HERE, "values.env"))
    ap.add_argument("overrides", nargs="*", help="KEY=VALUE")
    args = ap.parse_args()
    v = load_values(args.values)
    for o in args.overrides:
        if "=" not in o:
            ap.error(f"not KEY=VALUE: {o}")
        k, val = o.split("=", 1)
        if k not in v:
            ap.error(f"unknown value {k}; known: {', '.join(sorted(v))}")
        v[k] = val
    out = os.path.abspath(args.out or v["UNIT_DIR"])
    if args.uninstall:
        for u in UNITS:
            if not args.dry_run:
                subprocess.run(["systemctl", "disable", "--now", u], check=False)
            path = os.path.join(out, u)
            if os.path.isfile(path):
                os.unlink(path)
        if not args.dry_run:
            subprocess.run(["systemctl", "daemon-reload"], check=False)
        print(f"removed {' and '.join(UNITS)} from {out} (jobs stay in {v['STATE_DIR']})")
        return 0
    os.makedirs(out, exist_ok=True)
    for name, text in render(v).items():
        with open(os.path.join(out, name), "w") as f:
            f.write(text)
    print(f"rendered {', '.join(UNITS)} into {out}")
    if not args.dry_run:
        os.makedirs(v["WORK_ROOT"], exist_ok=True)
        os.makedirs(v["STATE_DIR"], exist_ok=True)
        subprocess.run(["systemctl", "daemon-reload"], check=True)
        subprocess.run(["systemctl", "enable", "--now", *UNITS], check=True)
        print("started; try: python -m paddle_operator_b200.cli get pdj")
    return 0


if __name__ == "__main__":
    sys.exit(main())
