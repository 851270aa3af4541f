"""CLI / manager process tests (CPU): manager flags, health + metrics endpoints, the REST API, and
the kubectl-like verbs against a manager running as a real subprocess."""
import json
import os
import subprocess
import sys
import time
import urllib.request

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PDJ = [sys.executable, "-m", "paddle_operator_b200.cli"]
ENV = dict(os.environ, PYTHONPATH=ROOT)


def _pdj(*a, server=None, check=True):
    cmd = PDJ + list(a) + (["--server", server] if server else [])
    r = subprocess.run(cmd, capture_output=True, text=True, env=ENV, cwd=ROOT, timeout=120)
    if check and r.returncode != 0:
        raise AssertionError(f"{' '.join(cmd)} -> {r.returncode}\n{r.stdout}\n{r.stderr}")
    return r


@pytest.fixture
def manager(tmp_path):
    ready = tmp_path / "ready.json"
    p = subprocess.Popen(PDJ + ["manager", "--api-bind-address", "127.0.0.1:0", "--metrics-bind-address", "127.0.0.1:0",
                                "--health-probe-bind-address", "127.0.0.1:0", "--work-root", str(tmp_path / "work"),
                                "--gpus", "0,1", "--zap-log-level", "debug", "--leader-elect",
                                "--ready-file", str(ready)],
                         env=dict(ENV, PDJ_ENV_DUMP_SLEEP="0.3"), cwd=ROOT, stderr=subprocess.PIPE, text=True)
    deadline = time.time() + 60
    while not ready.exists() and time.time() < deadline:
        assert p.poll() is None, p.stderr.read()
        time.sleep(0.05)
    url = json.load(open(ready))["api"]
    yield url, p
    p.terminate()
    try:
        p.wait(20)
    except subprocess.TimeoutExpired:
        p.kill()


def test_manager_and_client_verbs(manager):
    url, proc = manager
    r = _pdj("apply", "-f", "examples/env_dump.yaml", server=url)
    assert "paddlejob.batch.paddlepaddle.org/env-dump created" in r.stdout
    deadline = time.time() + 60
    out = ""
    while time.time() < deadline:
        out = _pdj("get", "pdj", server=url).stdout
        if "Completed" in out:
            break
        time.sleep(0.2)
    lines = out.strip().splitlines()
    assert lines[0].split() == ["NAME", "STATUS", "MODE", "AGE"]
    assert lines[1].split()[:3] == ["env-dump", "Completed", "PS"]
    pods = _pdj("get", "pods", "-o", "wide", server=url).stdout
    assert "env-dump-ps-0" in pods and "env-dump-worker-1" in pods and "Succeeded" in pods and "127." in pods
    y = _pdj("get", "pdj", "env-dump", "-o", "json", server=url).stdout
    st = json.loads(y)["status"]
    assert st["phase"] == "Completed" and st["ps"]["succeeded"] == 2 and st["startTime"] and st["completionTime"]
    d = _pdj("describe", "pdj", "env-dump", server=url).stdout
    assert "Created" in d and "created Pod env-dump-ps-0" in d and "created ConfigMap env-dump" in d
    logs = _pdj("logs", "env-dump-worker-0", server=url).stdout
    assert "PADDLE_TRAINER_ENDPOINTS" in logs
    ev = _pdj("events", server=url).stdout
    assert "REASON" in ev and "paddlejob/env-dump" in ev
    # validation errors come back from the server
    bad = _pdj("apply", "-f", "/dev/stdin", server=url, check=False) if False else None
    r = _pdj("delete", "pdj", "env-dump", server=url)
    assert 'paddlejob.batch.paddlepaddle.org "env-dump" deleted' in r.stdout
    deadline = time.time() + 30
    while time.time() < deadline and "No resources found" not in _pdj("get", "pdj", server=url).stdout:
        time.sleep(0.2)
    assert "No resources found" in _pdj("get", "pods", server=url).stdout
    r = _pdj("client-demo", server=url)
    assert "created client-demo" in r.stdout and "list: ['client-demo']" in r.stdout and "deleted client-demo" in r.stdout


def test_scale_and_health_metrics(manager, tmp_path):
    url, proc = manager
    job = tmp_path / "sleep.yaml"
    job.write_text("""
apiVersion: batch.paddlepaddle.org/v1
kind: PaddleJob
metadata: {name: sleepy}
spec:
  worker:
    replicas: 2
    template:
      spec:
        containers:
        - {name: m, image: x, command: [python, -c, "import time; time.sleep(120)"]}
""")
    _pdj("apply", "-f", str(job), server=url)
    _pdj("scale", "pdj", "sleepy", "--replicas", "3", server=url)
    deadline = time.time() + 30
    n = 0
    while time.time() < deadline and n != 3:
        n = _pdj("get", "pods", server=url).stdout.count("sleepy-worker-")
        time.sleep(0.2)
    assert n == 3
    with urllib.request.urlopen(url + "/apis/batch.paddlepaddle.org/v1/namespaces/default/paddlejobs/sleepy") as f:
        assert json.load(f)["spec"]["worker"]["replicas"] == 3
    _pdj("delete", "-f", str(job), server=url)


def test_manager_flag_surface():
    h = _pdj("manager", "--help").stdout
    for flag in ("--etcd-server", "--metrics-bind-address", "--namespace", "--health-probe-bind-address", "--port-range",
                 "--leader-elect", "--scheduling", "--initImage", "--zap-log-level"):
        assert flag in h


def test_health_and_metrics_endpoints(tmp_path):
    from paddle_operator_b200.runtime.apiserver import HttpEndpoint
    from paddle_operator_b200.runtime.cluster import LocalCluster

    c = LocalCluster(gpus=[0], work_root=str(tmp_path), supervisor=True)
    ep = HttpEndpoint(c, ("127.0.0.1", 0), health=True, metrics=True).start()
    try:
        for path in ("/healthz", "/readyz"):
            with urllib.request.urlopen(ep.url + path) as f:
                assert f.read() == b"ok"
        with urllib.request.urlopen(ep.url + "/metrics") as f:
            text = f.read().decode()
        assert 'controller_runtime_reconcile_total{controller="paddlejob"}' in text and "paddlejob_gpus_total 1" in text
        # the API is not served on the probe port
        with pytest.raises(Exception):
            urllib.request.urlopen(ep.url + "/apis/batch.paddlepaddle.org/v1/namespaces/default/paddlejobs")
    finally:
        ep.stop()
        c.stop()


def test_run_one_shot(tmp_path):
    r = _pdj("run", "-f", "examples/env_dump.yaml", "--work-root", str(tmp_path), "--gpus", "0", "--timeout", "90",
             "--etcd", "embedded")
    assert "env-dump Completed" in r.stdout and "NAME" in r.stdout


def _start_manager(tmp_path, tag, state_dir):
    ready = tmp_path / f"ready-{tag}.json"
    p = subprocess.Popen(PDJ + ["manager", "--api-bind-address", "127.0.0.1:0", "--metrics-bind-address", "0",
                                "--health-probe-bind-address", "0", "--work-root", str(tmp_path / "work"),
                                "--gpus", "0,1", "--state-dir", state_dir, "--ready-file", str(ready)],
                         env=dict(ENV, PDJ_ENV_DUMP_SLEEP="0.1"), cwd=ROOT, stderr=subprocess.PIPE, text=True)
    deadline = time.time() + 60
    while not ready.exists() and time.time() < deadline:
        assert p.poll() is None, p.stderr.read()
        time.sleep(0.05)
    return json.load(open(ready))["api"], p


def _stop(p):
    p.terminate()
    try:
        p.wait(20)
    except subprocess.TimeoutExpired:
        p.kill()


def test_manager_restart_keeps_jobs_with_state_dir(tmp_path):
    """`pdj manager --state-dir`: a finished job is still listed, with its status, by the next manager process."""
    state = str(tmp_path / "state")
    url, p = _start_manager(tmp_path, "a", state)
    try:
        _pdj("apply", "-f", "examples/env_dump.yaml", server=url)
        deadline = time.time() + 60
        while time.time() < deadline and "Completed" not in _pdj("get", "pdj", server=url).stdout:
            time.sleep(0.2)
        assert "Completed" in _pdj("get", "pdj", server=url).stdout
    finally:
        _stop(p)
    assert os.path.exists(os.path.join(state, "paddlejobs", "default", "env-dump.json"))
    url, p = _start_manager(tmp_path, "b", state)
    try:
        out = _pdj("get", "pdj", server=url).stdout
        assert out.strip().splitlines()[1].split()[:3] == ["env-dump", "Completed", "PS"]
        st = json.loads(_pdj("get", "pdj", "env-dump", "-o", "json", server=url).stdout)["status"]
        assert st["phase"] == "Completed" and st["completionTime"]
        time.sleep(1.0)  # the controller has reconciled it by now: still finished, no pods re-created
        assert "Completed" in _pdj("get", "pdj", server=url).stdout
        assert "env-dump-worker-0" not in _pdj("get", "pods", server=url).stdout
    finally:
        _stop(p)


def test_wait_watch_and_follow(manager):
    """`pdj wait`, `pdj get -w` and `pdj logs -f` against a live manager."""
    url, _ = manager
    _pdj("apply", "-f", "examples/env_dump.yaml", server=url)
    w = subprocess.Popen(PDJ + ["get", "pdj", "-w", "--watch-interval", "0.1", "--watch-timeout", "20", "--server", url],
                         stdout=subprocess.PIPE, text=True, env=ENV, cwd=ROOT)
    r = _pdj("wait", "pdj", "env-dump", "--for", "Completed", "--timeout", "60", server=url)
    assert "condition met: Completed" in r.stdout
    bad = _pdj("wait", "pdj", "env-dump", "--for", "Failed", "--timeout", "0.5", server=url, check=False)
    assert bad.returncode == 1 and "timed out" in bad.stderr
    logs = _pdj("logs", "-f", "env-dump-worker-0", server=url)   # finished pod: prints everything and returns
    assert "PADDLE_TRAINER_ID" in logs.stdout
    time.sleep(0.5)
    w.terminate()
    out = w.communicate(timeout=10)[0]
    assert out.count("NAME") >= 2 and "Completed" in out   # re-printed as the phase moved on


def test_scale_outside_the_elastic_window_is_rejected(manager, tmp_path):
    url, _ = manager
    doc = {"apiVersion": "batch.paddlepaddle.org/v1", "kind": "PaddleJob", "metadata": {"name": "el"},
           "spec": {"elastic": 1, "worker": {"replicas": 2, "requests": 1, "limits": 3, "template": {"spec": {"containers": [
               {"name": "m", "image": "local/pdjob-b200:env-dump"}]}}}}}
    f = tmp_path / "el.json"
    f.write_text(json.dumps(doc))
    _pdj("apply", "-f", str(f), server=url)
    ok = _pdj("scale", "pdj", "el", "--replicas", "3", server=url)
    assert "scaled" in ok.stdout
    bad = _pdj("scale", "pdj", "el", "--replicas", "4", server=url, check=False)
    assert bad.returncode != 0 and "422" in (bad.stderr + bad.stdout) and "limits" in (bad.stderr + bad.stdout)
    spec = json.loads(_pdj("get", "pdj", "el", "-o", "json", server=url).stdout)["spec"]
    assert spec["worker"]["replicas"] == 3


def test_installer_renders_units_from_values(tmp_path):
    """deploy/install.py --dry-run: the Helm-values / kustomize role of the reference's deployment manifests."""
    out = tmp_path / "units"
    r = subprocess.run([sys.executable, os.path.join(ROOT, "deploy", "install.py"), "--dry-run", "--out", str(out),
                        "SCHEDULING=volcano", "GPUS=0,1", "NAMESPACE=team-a", "STATE_DIR=/srv/pdj"],
                       capture_output=True, text=True, cwd=ROOT, timeout=60)
    assert r.returncode == 0, r.stderr
    unit = (out / "pdjob-manager.service").read_text()
    for frag in ("--leader-elect", "--scheduling volcano", "--gpus 0,1", "--namespace team-a", "--state-dir /srv/pdj",
                 "MemoryMax=300M", "paddle_operator_b200.cli manager"):
        assert frag in unit, frag
    assert "cli etcd --listen 127.0.0.1:2379" in (out / "pdjob-etcd.service").read_text()
    # the manager accepts exactly the flags the unit passes
    line = [l for l in unit.splitlines() if l.startswith("ExecStart=")][0]
    flags = line.split(" manager ", 1)[1].split()
    import shlex

    from paddle_operator_b200 import cli

    ns = cli.build_parser().parse_args(["manager", *shlex.split(" ".join(flags))])
    assert ns.scheduling == "volcano" and ns.state_dir == "/srv/pdj" and ns.leader_elect
    r = subprocess.run([sys.executable, os.path.join(ROOT, "deploy", "install.py"), "--dry-run", "--uninstall", "--out",
                        str(out)], capture_output=True, text=True, cwd=ROOT, timeout=60)
    assert r.returncode == 0 and not list(out.iterdir())


def test_crd_v1beta1_variant_has_the_reference_shape(capsys):
    """deploy/v1beta1/crd.yaml of the reference: printer columns (JSONPath), subresources and `validation` at the
    top of the spec, `version: v1` next to the versions list."""
    import yaml

    from paddle_operator_b200 import cli

    assert cli.main(["crd", "--version", "v1beta1"]) == 0
    doc = yaml.safe_load(capsys.readouterr().out)
    assert doc["apiVersion"] == "apiextensions.k8s.io/v1beta1"
    spec = doc["spec"]
    assert [c["JSONPath"] for c in spec["additionalPrinterColumns"]] == [".status.phase", ".status.mode", ".metadata.creationTimestamp"]
    assert spec["version"] == "v1" and spec["versions"] == [{"name": "v1", "served": True, "storage": True}]
    assert spec["subresources"] == {"status": {}} and "replicas" in str(spec["validation"]["openAPIV3Schema"])
    assert cli.main(["crd"]) == 0
    v1 = yaml.safe_load(capsys.readouterr().out)
    assert v1["spec"]["versions"][0]["schema"]["openAPIV3Schema"] == spec["validation"]["openAPIV3Schema"]
