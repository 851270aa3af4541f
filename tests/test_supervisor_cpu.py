"""Real rank processes on CPU (tier T-proc of SURVEY.md section 4): the
supervisor runs each Pod as a subprocess; ranks dump the environment they see."""
import json
import os
import signal
import time

import pytest
import yaml

from paddle_operator_b200.runtime.cluster import LocalCluster
from paddle_operator_b200.runtime.supervisor import resolve_command

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture
def cluster(tmp_path):
    c = LocalCluster(gpus=[0, 1, 2, 3], work_root=str(tmp_path / "work"),
                     extra_env={"PDJ_ENV_DUMP_DIR": str(tmp_path / "dump")})
    c.dump = tmp_path / "dump"
    with c:
        yield c


def _dumps(c):
    return {f[:-5]: json.load(open(c.dump / f)) for f in sorted(os.listdir(c.dump))}


def test_env_dump_job_runs_to_completion_in_order(cluster):
    cluster.apply(os.path.join(ROOT, "examples", "env_dump.yaml"))
    assert cluster.wait("env-dump", timeout=60) == "Completed"
    job = cluster.get("env-dump")
    assert job.status.mode == "PS" and job.status.start_time and job.status.completion_time
    assert job.status.ps.succeeded == 2 and job.status.worker.succeeded == 2 and job.status.heter.succeeded == 1
    d = _dumps(cluster)
    assert sorted(d) == ["env-dump-heter-0", "env-dump-ps-0", "env-dump-ps-1", "env-dump-worker-0", "env-dump-worker-1"]
    w1, p0 = d["env-dump-worker-1"], d["env-dump-ps-0"]
    assert w1["PADDLE_TRAINER_ID"] == "1" and w1["TRAINING_ROLE"] == "TRAINER" == w1["PADDLE_TRAINING_ROLE"]
    assert p0["TRAINING_ROLE"] == "PSERVER" and d["env-dump-heter-0"]["TRAINING_ROLE"] == "HETER"
    assert w1["PADDLE_TRAINERS_NUM"] == "2" and w1["TRAINER_PORTS_NUM"] == "20" and w1["PADDLE_PORT"] == "2379"
    eps = w1["PADDLE_TRAINER_ENDPOINTS"].split(",")
    assert len(eps) == 2 and all(e.startswith("127.") and e.endswith(":2379") for e in eps) and len(set(eps)) == 2
    assert w1["POD_IP"] + ":2379" == eps[1] and w1["PADDLE_TRAINERS"] == ",".join(e.split(":")[0] for e in eps)
    assert p0["PADDLE_PSERVERS_IP_PORT_LIST"].split(",")[0] == p0["POD_IP"] + ":2379"
    assert w1["PADDLE_WITH_GLOO"] == "1" and w1["PADDLE_GLOO_HTTP_ENDPOINT"] == p0["POD_IP"] + ":2397"
    assert len(w1["PADDLE_HETER_ENDPOINTS"].split(",")) == 1
    # every rank of the job agrees on the global environment
    for k in ("PADDLE_TRAINER_ENDPOINTS", "PADDLE_PSERVERS_IP_PORT_LIST", "PADDLE_HETER_ENDPOINTS"):
        assert len({v[k] for v in d.values()}) == 1
    # start order: ps before worker before heter (the coord gate)
    started = {p["metadata"]["name"]: p["status"]["containerStatuses"][0] for p in cluster.pods("env-dump")}
    assert all(cs["state"]["terminated"]["exitCode"] == 0 for cs in started.values())
    logs = cluster.logs("env-dump-worker-0")
    assert "PADDLE_TRAINER_ID" in logs


def test_gpu_assignment_failure_and_restart_policy(cluster):
    tmpl = {"spec": {"containers": [{"name": "m", "image": "x", "command": ["python", "-m",
                                                                             "paddle_operator_b200.workloads.env_dump"],
                                      "resources": {"limits": {"nvidia.com/gpu": 1}}}]}}
    job = {"apiVersion": "batch.paddlepaddle.org/v1", "kind": "PaddleJob", "metadata": {"name": "col"},
           "spec": {"worker": {"replicas": 3, "template": tmpl}}}
    cluster.apply(job)
    assert cluster.wait("col", timeout=60) == "Completed"
    d = _dumps(cluster)
    assert sorted(v["CUDA_VISIBLE_DEVICES"] for v in d.values()) == ["0", "1", "2"]
    assert all(v["PDJ_LOCAL_DEVICE"] == "0" for v in d.values())
    assert cluster.get("col").status.mode == "Collective"
    assert cluster.pods("col") == []  # default cleanPodPolicy cleans a completed job
    # a failing rank with restartPolicy Never fails the job, stickily
    bad = yaml.safe_load(yaml.safe_dump(job))
    bad["metadata"]["name"] = "bad"
    bad["spec"]["worker"]["replicas"] = 1
    bad["spec"]["worker"]["template"]["spec"]["containers"][0]["env"] = [{"name": "PDJ_ENV_DUMP_EXIT", "value": "3"}]
    cluster.apply(bad)
    assert cluster.wait("bad", timeout=60) == "Failed"
    pod = cluster.pods("bad")[0]
    assert pod["status"]["containerStatuses"][0]["state"]["terminated"]["exitCode"] == 3
    # elastic jobs restart in place (OnFailure)
    el = yaml.safe_load(yaml.safe_dump(bad))
    el["metadata"]["name"] = "el"
    el["spec"]["elastic"] = 1
    cluster.apply(el)
    deadline = time.time() + 30
    while time.time() < deadline:
        pods = cluster.pods("el")
        if pods and (pods[0]["status"].get("containerStatuses") or [{}])[0].get("restartCount", 0) >= 2:
            break
        time.sleep(0.05)
    else:
        raise AssertionError("the elastic rank was not restarted in place")
    assert cluster.get("el").status.phase != "Failed"
    cluster.delete("el")
    cluster.wait_deleted("el")


def test_kill_rank_and_delete_job_terminates_processes(cluster):
    tmpl = {"spec": {"containers": [{"name": "m", "image": "x", "command": ["python", "-c", "import time; time.sleep(60)"]}]}}
    cluster.apply({"apiVersion": "batch.paddlepaddle.org/v1", "kind": "PaddleJob", "metadata": {"name": "sleepy"},
                   "spec": {"worker": {"replicas": 2, "template": tmpl}}})
    assert cluster.wait("sleepy", phases=("Running",), timeout=30) == "Running"
    pids = [p["status"]["containerStatuses"][0]["pid"] for p in cluster.pods("sleepy")]
    assert cluster.supervisor.kill_rank("default", "sleepy-worker-1", signal.SIGKILL)
    assert cluster.wait("sleepy", timeout=30) == "Failed"
    cluster.delete("sleepy")
    cluster.wait_deleted("sleepy")
    time.sleep(0.3)
    for pid in pids:
        try:
            os.kill(pid, 0)
            alive = open(f"/proc/{pid}/stat").read().split()[2] != "Z"
        except (ProcessLookupError, FileNotFoundError):
            alive = False
        assert not alive


def test_command_resolution():
    assert resolve_command({"image": "registry.baidubce.com/paddle-operator/demo-wide-and-deep:v1"})[2] == \
        "paddle_operator_b200.workloads.train_wide_deep"
    c = resolve_command({"image": "x", "command": ["python"], "args": ["-m", "paddle.distributed.launch", "train_fleet.py"]})
    assert c[1:] == ["-m", "paddle_operator_b200.workloads.train_resnet"]
    assert resolve_command({"image": "unknown"}) is None
    # args without a command are appended to the image's registered entrypoint
    c = resolve_command({"image": "local/pdjob-b200:wide-and-deep", "args": ["--emb-dim", 32]})
    assert c[1:] == ["-m", "paddle_operator_b200.workloads.train_wide_deep", "--emb-dim", "32"]
    assert resolve_command({"image": "unknown", "args": ["--x"]}) is None


def test_every_example_manifest_loads_and_resolves():
    import glob

    from paddle_operator_b200.api import loader

    files = sorted(glob.glob(os.path.join(ROOT, "examples", "*.yaml")))
    assert len(files) >= 10
    for f in files:
        job = loader.load_paddlejob(f)
        for res in (job.spec.ps, job.spec.worker, job.spec.heter):
            if res is None:
                continue
            c0 = res.template["spec"]["containers"][0]
            assert resolve_command(c0), (f, c0)


def test_volumes_map_to_host_directories_and_claims_are_shared(cluster, tmp_path):
    """hostPath -> its path, emptyDir -> a per-rank scratch directory, persistentVolumeClaim -> one directory
    shared by every pod that mounts the claim (the reference's 'Data Storage' recipe, docs/user-guide.md:260-347)."""
    host = tmp_path / "ckpt"
    tmpl = {"spec": {"volumes": [{"name": "ck", "hostPath": {"path": str(host)}},
                                 {"name": "scratch", "emptyDir": {}},
                                 {"name": "data", "persistentVolumeClaim": {"claimName": "dataset"}}],
                     "containers": [{"name": "m", "image": "x",
                                     "command": ["python", "-m", "paddle_operator_b200.workloads.env_dump"],
                                     "volumeMounts": [{"name": "ck", "mountPath": "/checkpoint"},
                                                      {"name": "scratch", "mountPath": "/scratch"},
                                                      {"name": "data", "mountPath": "/data"}]}]}}
    cluster.apply({"apiVersion": "batch.paddlepaddle.org/v1", "kind": "PaddleJob", "metadata": {"name": "vols"},
                   "spec": {"worker": {"replicas": 2, "template": tmpl}}})
    assert cluster.wait("vols", timeout=60) == "Completed"
    d = _dumps(cluster)
    w0, w1 = d["vols-worker-0"], d["vols-worker-1"]
    assert w0["PDJ_MOUNT__checkpoint"] == str(host) == w1["PDJ_MOUNT__checkpoint"] and host.is_dir()
    assert w0["PDJ_MOUNT__scratch"] != w1["PDJ_MOUNT__scratch"]                       # emptyDir: private
    assert w0["PDJ_MOUNT__data"] == w1["PDJ_MOUNT__data"]                             # claim: shared
    assert w0["PDJ_MOUNT__data"].endswith(os.path.join("pvc", "default", "dataset")) and os.path.isdir(w0["PDJ_MOUNT__data"])


def test_gpu_workloads_fail_with_one_clear_line_without_a_gpu():
    import subprocess
    import sys

    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    for mod, extra in (("train_resnet", ["--tiny"]), ("train_wide_deep", [])):
        r = subprocess.run([sys.executable, "-m", f"paddle_operator_b200.workloads.{mod}", *extra], capture_output=True,
                           text=True, cwd=ROOT, timeout=120, env=dict(os.environ, TRAINING_ROLE="TRAINER", PYTHONPATH=ROOT))
        assert r.returncode == 3, (mod, r.returncode, r.stderr[-300:])
        assert "no usable NVIDIA GPU" in r.stderr and "Traceback" not in r.stderr


def test_assigned_gpus_are_visible_in_the_pod_status(cluster):
    tmpl = {"spec": {"containers": [{"name": "m", "image": "local/pdjob-b200:env-dump",
                                     "resources": {"limits": {"nvidia.com/gpu": 1}}}]}}
    cluster.apply({"apiVersion": "batch.paddlepaddle.org/v1", "kind": "PaddleJob", "metadata": {"name": "g"},
                   "spec": {"cleanPodPolicy": "Never", "worker": {"replicas": 2, "template": tmpl}}})
    assert cluster.wait("g", timeout=60) == "Completed"
    got = sorted(tuple(p["status"].get("gpus") or []) for p in cluster.pods("g"))
    assert len(got) == 2 and all(len(g) == 1 for g in got) and got[0] != got[1]

    from paddle_operator_b200.cli import print_objects
    import io
    import contextlib

    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        print_objects("pods", cluster.pods("g"), "wide")
    assert "GPUS" in buf.getvalue().splitlines()[0]


def test_warm_standbys_hold_free_gpus_and_yield_to_real_pods(cluster):
    """Elastic jobs get their possible extra ranks (replicas .. limits - 1) parked on free GPUs; a pod of another
    job that finds no GPU evicts one; deleting the job removes the rest."""
    sleeper = {"name": "m", "image": "x", "command": ["python", "-c", "import time; time.sleep(120)"],
               "resources": {"limits": {"nvidia.com/gpu": 1}}}
    cluster.apply({"apiVersion": "batch.paddlepaddle.org/v1", "kind": "PaddleJob", "metadata": {"name": "el"},
                   "spec": {"elastic": 1, "worker": {"replicas": 1, "requests": 1, "limits": 4,
                                                     "template": {"spec": {"containers": [sleeper]}}}}})
    sup = cluster.supervisor
    deadline = time.time() + 30
    while time.time() < deadline and len(sup.standbys()) < 3:
        time.sleep(0.05)
    parked = sorted(sup.standbys())
    assert [n for _, n, _ in parked] == ["el-worker-1", "el-worker-2", "el-worker-3"]
    used = {g for _, _, gs in parked for g in gs} | {g for p in cluster.pods("el") for g in p["status"].get("gpus", [])}
    assert used == {0, 1, 2, 3}  # every GPU of the box is either a rank's or a parked rank's
    # another job needs a GPU: a standby gives way
    cluster.apply({"apiVersion": "batch.paddlepaddle.org/v1", "kind": "PaddleJob", "metadata": {"name": "other"},
                   "spec": {"worker": {"replicas": 1, "template": {"spec": {"containers": [sleeper]}}}}})
    deadline = time.time() + 30
    while time.time() < deadline and not any(p["status"].get("phase") == "Running" for p in cluster.pods("other")):
        time.sleep(0.05)
    assert any(p["status"].get("phase") == "Running" for p in cluster.pods("other"))
    assert len(sup.standbys()) == 2
    # scale-up adopts a parked rank instead of starting a process
    cluster.scale("el", 2)
    deadline = time.time() + 30
    while time.time() < deadline and not any(p["status"].get("adoptedStandby") for p in cluster.pods("el")):
        time.sleep(0.05)
    adopted = [p for p in cluster.pods("el") if p["status"].get("adoptedStandby")]
    assert len(adopted) == 1 and adopted[0]["status"]["phase"] == "Running"
    cluster.delete("el")
    cluster.wait_deleted("el", timeout=60)
    deadline = time.time() + 10
    while time.time() < deadline and sup.standbys():
        time.sleep(0.05)
    assert sup.standbys() == []
