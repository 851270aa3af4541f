"""Control-plane tests (CPU): API types, pure helpers with golden outputs, and
the reconciler under the envtest-equivalent fixture (object store with no
supervisor, then with the scriptable FakeSupervisor).

The reference's only test (controllers/paddlejob_controller_test.go:58-112)
creates a Service-mode PS job with 3 ps + 2 workers, waits for 3/2 pod refs and
mode PS, then updates to 1 ps + 4 workers and waits for 1/4;
`test_reference_scenario_scale_3_2_to_1_4` is that scenario.
BASELINE.json config 1 (wide_and_deep, 2 ps + 2 workers, plumbing on CPU) is
`test_wide_and_deep_2ps_2workers_full_lifecycle`.
"""
import glob
import os

import pytest
import yaml
from hypothesis import given, settings
from hypothesis import strategies as st

from paddle_operator_b200.api import loader
from paddle_operator_b200.api.types import (PaddleJob, PaddleJobSpec, PaddleJobStatus, ResourceSpec, ResourceStatus)
from paddle_operator_b200.controller import helpers as H
from paddle_operator_b200.controller.gang import GangScheduler
from paddle_operator_b200.controller.hostport import HostPortAllocator, parse_annotation
from paddle_operator_b200.controller.manager import ControllerManager
from paddle_operator_b200.controller.quantity import Quantity
from paddle_operator_b200.controller.reconciler import PaddleJobReconciler
from paddle_operator_b200.controller.store import Conflict, EventRecorder, NotFound, ObjectStore
from paddle_operator_b200.runtime.fake_supervisor import FakeSupervisor

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXAMPLES = os.path.join(ROOT, "examples")
REF = "/root/reference"


def _job(name="wd", ps=2, worker=2, heter=None, **spec):
    tmpl = {"spec": {"containers": [{"name": "paddle", "image": "img"}]}}
    d = {"apiVersion": "batch.paddlepaddle.org/v1", "kind": "PaddleJob", "metadata": {"name": name},
         "spec": dict(spec)}
    if ps is not None:
        d["spec"]["ps"] = {"replicas": ps, "template": yaml.safe_load(yaml.safe_dump(tmpl))}
    if worker is not None:
        d["spec"]["worker"] = {"replicas": worker, "template": yaml.safe_load(yaml.safe_dump(tmpl))}
    if heter is not None:
        d["spec"]["heter"] = {"replicas": heter, "template": yaml.safe_load(yaml.safe_dump(tmpl))}
    return d


class Env:
    """store + reconciler + manager (+ fake supervisor)."""

    def __init__(self, batch=False, scheduling="", init_image="busybox", etcd=None, gang=None, supervisor=True):
        self.store = ObjectStore()
        self.rec = EventRecorder(self.store)
        self.sup = FakeSupervisor(self.store) if supervisor else None
        self.r = PaddleJobReconciler(self.store, self.rec, scheduling=scheduling, init_image=init_image, etcd=etcd,
                                     exec_in_pod=self.sup.exec_in_pod if self.sup else None, batch=batch)
        self.gang = gang(self.store) if gang else None
        self.mgr = ControllerManager(self.store, self.r, self.gang)

    def apply(self, doc):
        return self.store.create(doc)

    def job(self, name, ns="default") -> PaddleJob:
        return PaddleJob.from_dict(self.store.get("PaddleJob", ns, name))

    def settle(self, rounds=50):
        """Reconcile and let the fake kubelet react until nothing changes."""
        for _ in range(rounds):
            n = self.mgr.run_until_idle()
            moved = 0
            if self.sup:
                moved += self.sup.schedule_all()
                moved += self.sup.start_ready()
            if n == 0 and moved == 0:
                break
            for o in self.store.list("PaddleJob"):
                self.mgr.enqueue(o["metadata"]["namespace"], o["metadata"]["name"])
            if moved == 0 and n <= len(self.store.list("PaddleJob")) + 1:
                break


# ---------------------------------------------------------------------------
# API types / loader
def test_example_manifests_load_and_roundtrip():
    files = sorted(glob.glob(os.path.join(EXAMPLES, "*.yaml")))
    assert len(files) >= 6
    for f in files:
        for job in loader.load_paddlejobs(f):
            d = job.to_dict()
            again = PaddleJob.from_dict(d)
            assert again.to_dict() == d
            raw = [x for x in yaml.safe_load_all(open(f)) if x][0]
            assert d["spec"] == raw["spec"]  # nothing lost, nothing invented


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree not mounted")
def test_reference_manifests_are_accepted_verbatim():
    files = glob.glob(f"{REF}/deploy/examples/*.yaml") + glob.glob(f"{REF}/deploy/elastic/*.yaml") + \
        [f"{REF}/config/samples/batch_v1_paddlejob.yaml"]
    seen = 0
    for f in files:
        for doc in loader.load_documents(f):
            if isinstance(doc, PaddleJob):
                seen += 1
                assert doc.spec.worker is not None
                assert H.get_paddlejob_mode(doc) in ("PS", "Collective", "Single")
    assert seen >= 8
    deepfm = loader.load_paddlejob(f"{REF}/deploy/examples/deepfm.yaml")
    assert deepfm.spec.clean_pod_policy == "OnFailed"  # not an enum value, still accepted
    assert any("behaves like Never" in w for w in loader.warnings_for(deepfm))
    etcd_docs = loader.load_documents(f"{REF}/deploy/elastic/etcd.yaml")
    assert {d["kind"] for d in etcd_docs} == {"Deployment", "Service"}


def test_validation_and_pruning():
    with pytest.raises(loader.ValidationError):
        loader.load_documents("apiVersion: batch.paddlepaddle.org/v1\nkind: PaddleJob\nmetadata: {name: x}\n"
                              "spec: {worker: {template: {}}}")
    with pytest.raises(loader.ValidationError):
        loader.load_documents("apiVersion: batch.paddlepaddle.org/v1\nkind: PaddleJob\nmetadata: {}\nspec: {}")
    d = _job(cleanPolicy="Always")
    assert loader.warnings_for(d) == ["unknown field spec.cleanPolicy is pruned"]
    assert "cleanPolicy" not in PaddleJob.from_dict(d).to_dict()["spec"]
    st_ = PaddleJobStatus(phase="Running", mode="PS", ps=ResourceStatus(running=2, refs=[{"name": "a"}]))
    assert st_.to_dict() == {"phase": "Running", "mode": "PS", "ps": {"running": 2, "refs": [{"name": "a"}]}}


def test_quantity_arithmetic():
    assert str(Quantity.parse("500m") + Quantity.parse("500m")) == "1"
    assert str(Quantity.parse("1Gi") + Quantity.parse("512Mi")) == "1536Mi"
    assert str(Quantity.parse(1) + Quantity.parse("3")) == "4"
    assert str(Quantity.parse("10m") + Quantity.parse("1")) == "1010m"
    assert int(Quantity.parse("8")) == 8 and Quantity.parse("2k") == Quantity.parse(2000)


# ---------------------------------------------------------------------------
# pure helpers: golden outputs
def test_naming_mode_and_roles():
    assert H.gen_paddle_res_name("job", "worker", 3) == "job-worker-3"
    assert H.extract_name_index("wide-ande-deep-ps-1") == ("ps", 1)
    assert H.extract_name_index("nonsense") == ("", 0)
    assert H.get_paddlejob_mode(PaddleJob.from_dict(_job())) == "PS"
    assert H.get_paddlejob_mode(PaddleJob.from_dict(_job(ps=None, worker=2))) == "Collective"
    assert H.get_paddlejob_mode(PaddleJob.from_dict(_job(ps=None, worker=1))) == "Single"


def test_construct_pod_golden():
    pdj = PaddleJob.from_dict(_job("wd", withGloo=1))
    pod = H.construct_pod(pdj, "ps", 1)
    assert pod["metadata"]["name"] == "wd-ps-1"
    assert pod["metadata"]["labels"] == {"paddle-res-name": "wd-ps-1", "paddle-res-type": "ps"}
    assert pod["metadata"]["annotations"] == {"paddle-resource": "ps"}
    assert pod["spec"]["hostname"] == "wd-ps-1" and pod["spec"]["subdomain"] == "wd-ps-1"
    c0 = pod["spec"]["containers"][0]
    assert c0["env"] == [
        {"name": "POD_IP", "valueFrom": {"fieldRef": {"fieldPath": "status.podIP"}}},
        {"name": "PADDLE_TRAINER_ID", "value": "1"},
        {"name": "TRAINING_ROLE", "value": "PSERVER"},
        {"name": "PADDLE_TRAINING_ROLE", "value": "PSERVER"},
    ]
    assert c0["envFrom"] == [{"configMapRef": {"name": "wd"}}]
    assert pod["spec"]["restartPolicy"] == "Never" and "hostNetwork" not in pod["spec"]
    # the template itself is not modified
    assert "env" not in pdj.spec.ps.template["spec"]["containers"][0]

    svc_job = PaddleJob.from_dict(_job("wd", intranet="Service"))
    w = H.construct_pod(svc_job, "worker", 0)
    assert w["spec"]["containers"][0]["env"][0] == {"name": "POD_IP", "value": "wd-worker-0"}
    assert w["spec"]["containers"][0]["ports"] == [{"containerPort": 2379}]
    assert w["spec"]["restartPolicy"] == "OnFailure"
    assert H.construct_pod(svc_job, "ps", 0)["spec"]["restartPolicy"] == "Never"
    host_job = PaddleJob.from_dict(_job("wd", intranet="Host"))
    assert H.construct_pod(host_job, "worker", 0)["spec"]["hostNetwork"] is True

    el = PaddleJob.from_dict(_job("rs", ps=None, worker=3, elastic=1))
    el.metadata.namespace = "team"
    p = H.construct_pod(el, "worker", 2)
    names = {e["name"]: e.get("value") for e in p["spec"]["containers"][0]["env"]}
    assert names["PADDLE_ELASTIC_JOB_ID"] == "team-rs" and names["PADDLE_ELASTIC_NP"] == "3"
    assert names["PADDLE_ELASTIC_TIMEOUT"] == "60" and names["TRAINING_ROLE"] == "TRAINER"
    assert "envFrom" not in p["spec"]["containers"][0] and p["spec"]["restartPolicy"] == "OnFailure"


def _pods_with_ips(pdj, ips):
    pods = []
    for res, spec in pdj.get_specs().items():
        if spec is None:
            continue
        for i in range(spec.replicas):
            p = H.construct_pod(pdj, res, i)
            p["status"] = {"podIP": ips[(res, i)]}
            pods.append(p)
    return pods


def test_construct_configmap_golden():
    pdj = PaddleJob.from_dict(_job("wd", ps=2, worker=2, heter=1, withGloo=1))
    ips = {("ps", 0): "10.0.0.1", ("ps", 1): "10.0.0.2", ("worker", 0): "10.0.0.3", ("worker", 1): "10.0.0.4",
           ("heter", 0): "10.0.0.5"}
    cm = H.construct_configmap(pdj, _pods_with_ips(pdj, ips))
    assert cm["metadata"]["name"] == "wd" and cm["metadata"]["labels"] == {"paddle-res-name": "wd"}
    assert cm["data"] == {
        "TRAINER_PORTS_NUM": "20",
        "PADDLE_PORT": "2379",
        "PADDLE_PSERVERS_IP_PORT_LIST": "10.0.0.1:2379,10.0.0.2:2379",
        "PADDLE_TRAINER_ENDPOINTS": "10.0.0.3:2379,10.0.0.4:2379",
        "PADDLE_TRAINERS": "10.0.0.3,10.0.0.4",
        "PADDLE_TRAINERS_NUM": "2",
        "PADDLE_HETER_ENDPOINTS": "10.0.0.5:2379",
        "PADDLE_WITH_GLOO": "1",
        "PADDLE_GLOO_RENDEZVOUS": "3",
        "PADDLE_GLOO_HTTP_ENDPOINT": "10.0.0.1:2397",
    }
    # an address that is not a dotted quad -> not yet (requeue)
    bad = dict(ips)
    bad[("worker", 1)] = ""
    assert H.construct_configmap(pdj, _pods_with_ips(pdj, bad)) is None
    # Service mode publishes names; Host mode publishes the annotated port, lists keep :2379
    svc = PaddleJob.from_dict(_job("wd", intranet="Service"))
    data = H.construct_configmap(svc, _pods_with_ips(svc, ips))["data"]
    assert data["PADDLE_PSERVERS_IP_PORT_LIST"] == "wd-ps-0:2379,wd-ps-1:2379"
    assert data["PADDLE_TRAINERS"] == "wd-worker-0,wd-worker-1" and "PADDLE_WITH_GLOO" not in data
    host = PaddleJob.from_dict(_job("wd", intranet="Host"))
    host.metadata.annotations["host-port"] = "35020"
    data = H.construct_configmap(host, _pods_with_ips(host, ips))["data"]
    assert data["PADDLE_PORT"] == "35020" and data["PADDLE_TRAINER_ENDPOINTS"] == "10.0.0.3:2379,10.0.0.4:2379"
    # withGloo without ps: no gloo keys (reference needs >= 1 ps)
    col = PaddleJob.from_dict(_job("c", ps=None, worker=2, withGloo=1))
    assert "PADDLE_WITH_GLOO" not in H.construct_configmap(col, _pods_with_ips(col, ips))["data"]


def test_service_and_podgroup_golden():
    pdj = loader.load_paddlejob(os.path.join(EXAMPLES, "wide_and_deep_volcano.yaml"))
    pod = H.construct_pod(pdj, "worker", 0)
    svc = H.construct_service_for_pod(pod)
    assert svc["spec"]["clusterIP"] == "None" and svc["spec"]["selector"] == {"paddle-res-name": pod["metadata"]["name"]}
    assert [p["port"] for p in svc["spec"]["ports"]] == list(range(2379, 2399))
    assert svc["spec"]["ports"][0]["name"] == "p-0"
    pg = H.construct_podgroup(pdj)
    assert pg["metadata"]["name"] == pdj.name and pg["spec"]["minMember"] == 4
    assert pg["spec"]["queue"] == "default" and pg["spec"]["priorityClassName"] == "high"
    assert pg["spec"]["minResources"] == {"cpu": "2", "nvidia.com/gpu": "2"}
    assert not H.without_volcano(pdj)
    pdj.spec.ps.template["spec"]["schedulerName"] = "default-scheduler"
    assert H.without_volcano(pdj)
    coord = H.gen_coordinate_init_container("busybox:1")
    assert coord["name"] == "coord-paddle" and "goon" in coord["command"][2]
    assert coord["resources"]["requests"] == {"cpu": "10m", "memory": "10m"}


_counts = st.builds(ResourceStatus, pending=st.integers(0, 2), starting=st.integers(0, 2), running=st.integers(0, 3),
                    failed=st.integers(0, 1), succeeded=st.integers(0, 3))


@settings(max_examples=300, deadline=None)
@given(ps=st.one_of(st.none(), _counts), worker=st.one_of(st.none(), _counts),
       prev=st.sampled_from(["", "Pending", "Starting", "Running", "Completed", "Failed"]),
       nps=st.integers(0, 3), nw=st.integers(1, 3))
def test_phase_derivation_properties(ps, worker, prev, nps, nw):
    pdj = PaddleJob(spec=PaddleJobSpec(ps=ResourceSpec(nps) if nps else None, worker=ResourceSpec(nw)),
                    status=PaddleJobStatus(phase=prev, ps=ps if nps else None, worker=worker))
    got = H.get_paddlejob_phase(pdj)
    if prev in ("Completed", "Failed"):
        assert got == prev  # sticky
        return
    sts = [s for s in (pdj.status.ps, pdj.status.worker) if s is not None]
    first = next((s for s in sts if s.failed or s.starting or s.pending), None)
    if first is not None:
        assert got == ("Failed" if first.failed else "Starting" if first.starting else "Pending")
    elif all(s is not None and sp.replicas == s.running for sp, s in
             [(x, y) for x, y in ((pdj.spec.ps, pdj.status.ps), (pdj.spec.worker, pdj.status.worker)) if x is not None]):
        assert got == "Running"
    assert got in ("Pending", "Starting", "Running", "Completed", "Failed")


# ---------------------------------------------------------------------------
# store semantics
def test_store_versions_finalizers_and_gc():
    s = ObjectStore()
    seen = []
    s.watch(lambda t, o: seen.append((t, o["kind"], o["metadata"]["name"])))
    job = s.create(_job("j"))
    assert job["metadata"]["generation"] == 1 and job["metadata"]["uid"]
    stale = dict(job)
    job["spec"]["worker"]["replicas"] = 5
    job2 = s.update(job)
    assert job2["metadata"]["generation"] == 2
    with pytest.raises(Conflict):
        s.update(stale)
    job2["status"] = {"phase": "Running"}
    job3 = s.update_status(job2)
    assert job3["metadata"]["generation"] == 2 and s.get("PaddleJob", "default", "j")["status"]["phase"] == "Running"
    pod = s.create({"kind": "Pod", "metadata": {"name": "j-worker-0", "ownerReferences": [
        {"kind": "PaddleJob", "name": "j", "controller": True, "uid": job["metadata"]["uid"]}]}})
    assert [p["metadata"]["name"] for p in s.list("Pod", "default", owner="j")] == ["j-worker-0"]
    assert s.list("Pod", "default", owner="other") == []
    job3["metadata"]["finalizers"] = ["f"]
    job4 = s.update(job3)
    s.delete("PaddleJob", "default", "j")
    assert s.get("PaddleJob", "default", "j")["metadata"]["deletionTimestamp"]  # held by the finalizer
    assert s.try_get("Pod", "default", "j-worker-0") is not None
    cur = s.get("PaddleJob", "default", "j")
    cur["metadata"]["finalizers"] = []
    s.update(cur)
    with pytest.raises(NotFound):
        s.get("PaddleJob", "default", "j")
    assert s.try_get("Pod", "default", "j-worker-0") is None  # garbage collected with its owner
    assert ("DELETED", "Pod", "j-worker-0") in seen and pod and job4


def test_host_port_allocator():
    a = HostPortAllocator(35000, 35100)
    got = [a.allocate() for _ in range(5)]
    assert got == [35020, 35040, 35060, 35080, 35000]  # cursor moves first, wraps at the end of the range
    assert a.allocate() == 40000  # exhausted -> the reference's fixed fallback
    assert a.release(35040) and a.allocate() == 35040
    b = HostPortAllocator.from_flag("35000,65000")
    b.mark("35020")
    assert b.allocate() == 35040 and b.is_allocated(35020)
    assert parse_annotation("30001, 30002") == [30001, 30002]
    with pytest.raises(ValueError):
        parse_annotation("70000")


# ---------------------------------------------------------------------------
# reconciler: envtest-equivalent (no supervisor) -------------------------------
@pytest.mark.parametrize("batch", [False, True])
def test_reference_scenario_scale_3_2_to_1_4(batch):
    env = Env(batch=batch, init_image="", supervisor=False)
    env.apply(_job("wide-and-deep-service", ps=3, worker=2, cleanPodPolicy="Never", intranet="Service"))
    env.mgr.run_until_idle()
    job = env.job("wide-and-deep-service")
    assert len(job.status.ps.refs) == 3 and len(job.status.worker.refs) == 2
    assert job.status.mode == "PS" and job.status.phase == "Pending"
    assert H.FINALIZER_NAME in job.metadata.finalizers
    names = sorted(p["metadata"]["name"] for p in env.store.list("Pod"))
    assert names == ["wide-and-deep-service-ps-0", "wide-and-deep-service-ps-1", "wide-and-deep-service-ps-2",
                     "wide-and-deep-service-worker-0", "wide-and-deep-service-worker-1"]
    assert sorted(s["metadata"]["name"] for s in env.store.list("Service")) == names
    for p in env.store.list("Pod"):
        assert "initContainers" not in p["spec"]  # --initImage "" disables the gate
        assert p["metadata"]["ownerReferences"][0]["name"] == "wide-and-deep-service"
    obj = env.store.get("PaddleJob", "default", "wide-and-deep-service")
    obj["spec"]["ps"]["replicas"] = 1
    obj["spec"]["worker"]["replicas"] = 4
    env.store.update(obj)
    env.mgr.run_until_idle()
    job = env.job("wide-and-deep-service")
    assert len(job.status.ps.refs) == 1 and len(job.status.worker.refs) == 4
    assert job.status.observed_generation == 2
    reasons = [e["reason"] for e in env.rec.events_for("PaddleJob", "default", "wide-and-deep-service")]
    assert reasons.count("Deleted") == 2 and reasons.count("Created") >= 5 + 5 + 2
    assert env.store.try_get("ConfigMap", "default", "wide-and-deep-service") is None  # no pod has an IP


def test_one_mutation_per_pass_when_not_batched():
    env = Env(batch=False, init_image="", supervisor=False)
    env.apply(_job("j", ps=1, worker=2))
    counts = []
    for _ in range(6):
        env.r.reconcile("default", "j")
        counts.append(len(env.store.list("Pod")))
    assert counts == [1, 2, 3, 3, 3, 3]
    env2 = Env(batch=True, init_image="", supervisor=False)
    env2.apply(_job("j", ps=1, worker=2))
    env2.r.reconcile("default", "j")
    assert len(env2.store.list("Pod")) == 3


# reconciler with the fake kubelet ----------------------------------------------
@pytest.mark.parametrize("batch", [False, True])
def test_wide_and_deep_2ps_2workers_full_lifecycle(batch):
    env = Env(batch=batch)
    job = loader.load_paddlejob(os.path.join(EXAMPLES, "wide_and_deep.yaml"))
    env.apply(job.to_dict())
    name = job.name
    env.mgr.run_until_idle()
    assert len(env.store.list("Pod")) == 4
    assert all(p["spec"]["initContainers"][0]["name"] == "coord-paddle" for p in env.store.list("Pod"))
    env.sup.schedule_all()
    env.mgr.run_until_idle()
    j = env.job(name)
    # both ps gates were opened in this round (their init containers exited, main not yet started),
    # the workers are still held by theirs
    assert j.status.ps.pending == 2 and j.status.worker.starting == 2 and j.status.phase == "Pending"
    cm = env.store.get("ConfigMap", "default", name)
    assert cm["data"]["PADDLE_TRAINERS_NUM"] == "2" and cm["data"]["PADDLE_WITH_GLOO"] == "1"
    assert cm["data"]["PADDLE_PSERVERS_IP_PORT_LIST"].count(":2379") == 2
    # gate: ps released first, workers only once both ps are running
    released = [e[1] for e in env.sup.exec_log]
    assert sorted(released) == [f"{name}-ps-0", f"{name}-ps-1"]
    assert all(e[2] == "coord-paddle" and e[3] == ("touch", "goon") for e in env.sup.exec_log)
    env.sup.start_ready()
    env.mgr.run_until_idle()
    released = [e[1] for e in env.sup.exec_log]
    assert f"{name}-worker-0" in released and f"{name}-worker-1" in released
    assert released.index(f"{name}-worker-0") > released.index(f"{name}-ps-1")
    env.sup.start_ready()
    env.mgr.run_until_idle()
    j = env.job(name)
    assert j.status.phase == "Running" and j.status.start_time and j.status.ps.running == 2
    env.sup.finish_all(succeeded=True)
    env.mgr.run_until_idle()
    j = env.job(name)
    assert j.status.phase == "Completed" and j.status.completion_time and j.status.worker.succeeded == 2
    assert len(env.store.list("Pod")) == 4  # cleanPodPolicy: Never
    # delete: finalizer removed, children garbage collected
    env.store.delete("PaddleJob", "default", name)
    env.mgr.run_until_idle()
    assert env.store.try_get("PaddleJob", "default", name) is None
    assert env.store.list("Pod") == [] and env.store.try_get("ConfigMap", "default", name) is None


@pytest.mark.parametrize("policy,fail,left", [("", False, 0), ("", True, 2), ("Always", True, 0), ("OnFailure", True, 0),
                                              ("OnFailure", False, 2), ("OnCompletion", False, 0),
                                              ("OnCompletion", True, 2), ("Never", False, 2), ("OnFailed", True, 2)])
def test_clean_pod_policy(policy, fail, left):
    env = Env(batch=False)
    spec = {"cleanPodPolicy": policy} if policy else {}
    env.apply(_job("c", ps=None, worker=2, **spec))
    env.settle()
    assert env.job("c").status.phase == "Running"
    if fail:
        env.sup.finish("default", "c-worker-1", succeeded=False)
    else:
        env.sup.finish_all()
    env.mgr.run_until_idle()
    j = env.job("c")
    assert j.status.phase == ("Failed" if fail else "Completed")
    assert len(env.store.list("Pod")) == left
    # a failed job stays failed even if the other rank later succeeds
    if fail and left:
        env.sup.finish_all()
        env.mgr.run_until_idle()
        assert env.job("c").status.phase == "Failed"


def test_service_mode_cleanup_deletes_services_too():
    env = Env()
    env.apply(loader.load_paddlejob(os.path.join(EXAMPLES, "wide_and_deep_service.yaml")).to_dict())
    env.settle()
    assert len(env.store.list("Service")) == 4 and env.job("wide-ande-deep-service").status.phase == "Running"
    env.sup.finish_all()
    env.mgr.run_until_idle()
    assert env.store.list("Pod") == [] and env.store.list("Service") == []


def test_host_mode_allocates_and_releases_port_block():
    env = Env()
    env.apply(loader.load_paddlejob(os.path.join(EXAMPLES, "wide_and_deep_host.yaml")).to_dict())
    env.settle()
    j = env.job("wide-ande-deep-host")
    port = j.metadata.annotations["host-port"]
    assert port == "35020" and env.r.ports.is_allocated(port)
    assert env.store.get("ConfigMap", "default", j.name)["data"]["PADDLE_PORT"] == port
    assert all(p["spec"]["hostNetwork"] for p in env.store.list("Pod"))
    # a restarted manager rebuilds its allocator from the annotation
    r2 = PaddleJobReconciler(env.store, env.rec, init_image="busybox", exec_in_pod=env.sup.exec_in_pod)
    assert not r2.ports.is_allocated(port)
    r2.reconcile("default", j.name)
    assert r2.ports.is_allocated(port)
    env.store.delete("PaddleJob", "default", j.name)
    env.mgr.run_until_idle()
    assert not env.r.ports.is_allocated(port) and env.store.try_get("PaddleJob", "default", j.name) is None


def test_volcano_gang_admission_and_queueing():
    env = Env(scheduling="volcano", gang=lambda s: GangScheduler(s, {"nvidia.com/gpu": 2}, {"high": 10}))
    a = loader.load_paddlejob(os.path.join(EXAMPLES, "wide_and_deep_volcano.yaml")).to_dict()
    env.apply(a)
    b = loader.load_paddlejob(os.path.join(EXAMPLES, "wide_and_deep_volcano.yaml")).to_dict()
    b["metadata"]["name"] = "second"
    env.apply(b)
    env.settle()
    first = "wide-ande-deep-gang"
    pg = env.store.get("PodGroup", "default", first)
    assert pg["spec"]["minMember"] == 4 and pg["status"]["phase"] == "Running"
    pods = env.store.list("Pod", "default", owner=first)
    assert len(pods) == 4 and all(p["spec"]["schedulerName"] == "volcano" for p in pods)
    ann = pods[0]["metadata"]["annotations"]
    assert ann["scheduling.k8s.io/group-name"] == first and ann["volcano.sh/queue-name"] == "default"
    assert ann["volcano.sh/job-name"] == first and ann["volcano.sh/task-spec"] in ("ps", "worker")
    # the second job needs 2 more GPUs than the pool has: its group stays Pending, no pods
    assert env.store.get("PodGroup", "default", "second")["status"]["phase"] == "Pending"
    assert env.store.list("Pod", "default", owner="second") == []
    # when the first finishes, its PodGroup is deleted and the second is admitted
    env.sup.finish_all()
    env.settle()
    assert env.store.try_get("PodGroup", "default", first) is None
    env.settle()
    assert len(env.store.list("Pod", "default", owner="second")) == 4
    # a template that names another scheduler opts out of the gang
    c = _job("optout", ps=None, worker=1)
    c["spec"]["worker"]["template"]["spec"]["schedulerName"] = "default-scheduler"
    env.apply(c)
    env.settle()
    assert env.store.try_get("PodGroup", "default", "optout") is None
    assert len(env.store.list("Pod", "default", owner="optout")) == 1


class _Etcd:
    def __init__(self):
        self.kv = {}
        self.puts = []

    def endpoints(self):
        return ["127.0.0.1:2379"]

    def get(self, key):
        return [(key, self.kv[key])] if key in self.kv else []

    def put(self, key, value):
        self.kv[key] = value
        self.puts.append((key, value))


def test_elastic_np_sync_and_env():
    etcd = _Etcd()
    env = Env(etcd=etcd)
    env.apply(_job("resnet-elastic", ps=None, worker=3, elastic=1))
    env.settle()
    pods = env.store.list("Pod")
    assert len(pods) == 3 and env.store.try_get("ConfigMap", "default", "resnet-elastic") is None
    e = {x["name"]: x.get("value") for x in pods[0]["spec"]["containers"][0]["env"]}
    assert e["PADDLE_ELASTIC_SERVER"] == "127.0.0.1:2379" and e["PADDLE_ELASTIC_JOB_ID"] == "default-resnet-elastic"
    assert pods[0]["spec"]["restartPolicy"] == "OnFailure"
    assert etcd.puts == []  # the controller never creates the key
    etcd.kv["/paddle/default-resnet-elastic/np"] = "3"  # the workers did
    obj = env.store.get("PaddleJob", "default", "resnet-elastic")
    obj["spec"]["worker"]["replicas"] = 2
    env.store.update(obj)
    env.settle()
    assert etcd.puts == [("/paddle/default-resnet-elastic/np", "2")]
    assert len(env.store.list("Pod")) == 2
    assert "Scaled" in [ev["reason"] for ev in env.rec.events_for("PaddleJob", "default", "resnet-elastic")]
    obj = env.store.get("PaddleJob", "default", "resnet-elastic")
    obj["spec"]["worker"]["replicas"] = 4
    env.store.update(obj)
    env.settle()
    assert etcd.kv["/paddle/default-resnet-elastic/np"] == "4" and len(env.store.list("Pod")) == 4
    new = [p for p in env.store.list("Pod") if p["metadata"]["name"].endswith("-3")][0]
    assert {x["name"]: x.get("value") for x in new["spec"]["containers"][0]["env"]}["PADDLE_ELASTIC_NP"] == "4"


def test_status_written_only_on_change():
    env = Env(init_image="", supervisor=False)
    env.apply(_job("j", ps=None, worker=1))
    env.mgr.run_until_idle()
    rv = env.store.get("PaddleJob", "default", "j")["metadata"]["resourceVersion"]
    for _ in range(3):
        env.r.reconcile("default", "j")
    assert env.store.get("PaddleJob", "default", "j")["metadata"]["resourceVersion"] == rv


def test_legacy_hostport_manager_semantics_and_crd():
    """third_party/hostport-allocator of the reference: N ports per job, comma annotation, release."""
    from paddle_operator_b200.controller.hostport import (LEGACY_HOSTPORT_ANNOTATION, LEGACY_PORTNUM_ANNOTATION,
                                                          format_annotation)

    a = HostPortAllocator(30000, 30000 + 200, block=1)
    ports = a.allocate_many(3)
    assert len(set(ports)) == 3 and all(30000 <= p <= 30200 for p in ports)
    assert parse_annotation(format_annotation(ports)) == ports
    assert LEGACY_PORTNUM_ANNOTATION == "hostport-manager/portnum" and LEGACY_HOSTPORT_ANNOTATION == "hostport-manager/hostport"
    for p in ports:
        assert a.release(p)
    small = HostPortAllocator(30000, 30002, block=1)
    with pytest.raises(RuntimeError):
        small.allocate_many(5)
    schema = loader.openapi_schema()
    assert schema["properties"]["spec"]["properties"]["worker"]["required"] == ["replicas"]
    assert set(schema["properties"]["status"]["properties"]) >= {"phase", "mode", "ps", "worker", "heter", "elastic",
                                                                  "startTime", "completionTime", "observedGeneration"}


def test_trace_spans_cover_reconcile_passes(tmp_path):
    from paddle_operator_b200.utils import trace

    trace.enable()
    n0 = len(trace.events())
    env = Env(init_image="", supervisor=False)
    env.apply(_job("t", ps=None, worker=1))
    env.mgr.run_until_idle()
    evs = [e for e in trace.events()[n0:] if e["name"] == "reconcile"]
    assert evs and evs[0]["args"]["job"] == "t" and evs[0]["dur"] >= 0
    out = trace.flush(str(tmp_path / "trace.json"))
    import json as _json
    assert _json.load(open(out))["traceEvents"]


def test_jobs_and_status_survive_a_manager_restart(tmp_path):
    """--state-dir: a PaddleJob and its status are reloaded by the next manager (the reference gets this from
    the API server's storage); finished jobs stay finished; deleting a job removes its file."""
    import json as _json

    from paddle_operator_b200.runtime.cluster import LocalCluster

    state = str(tmp_path / "state")
    c1 = LocalCluster(supervisor=False, init_image="", state_dir=state)
    c1.apply(os.path.join(EXAMPLES, "wide_and_deep.yaml"))
    c1.manager.run_until_idle()
    job = c1.get("wide-ande-deep")
    assert len(job.status.ps.refs) == 2 and len(job.status.worker.refs) == 2 and job.status.mode == "PS"
    # mark it Completed the way the supervisor would: every pod Succeeded
    for pod in c1.store.list("Pod", "default"):
        pod["status"] = {"phase": "Succeeded"}
        c1.store.update_status(pod)
    c1.manager.run_until_idle()
    assert c1.get("wide-ande-deep").status.phase == "Completed"
    uid = c1.store.get("PaddleJob", "default", "wide-ande-deep")["metadata"]["uid"]
    path = c1.persister.path_of("PaddleJob", "default", "wide-ande-deep")
    on_disk = _json.load(open(path))
    assert on_disk["status"]["phase"] == "Completed" and on_disk["status"]["mode"] == "PS"
    assert set(on_disk["status"]) >= {"phase", "mode", "ps", "worker"}
    c1.stop()

    c2 = LocalCluster(supervisor=False, init_image="", state_dir=state)
    assert c2.restored == 1
    job2 = c2.get("wide-ande-deep")
    assert job2.status.phase == "Completed" and job2.status.mode == "PS"
    assert c2.store.get("PaddleJob", "default", "wide-ande-deep")["metadata"]["uid"] == uid
    c2.manager.enqueue("default", "wide-ande-deep")
    c2.manager.run_until_idle()
    assert c2.get("wide-ande-deep").status.phase == "Completed"  # sticky: nothing is re-run
    c2.delete("wide-ande-deep")
    c2.manager.run_until_idle()
    assert c2.store.try_get("PaddleJob", "default", "wide-ande-deep") is None
    assert not os.path.exists(path)
    c2.stop()

    c3 = LocalCluster(supervisor=False, init_image="", state_dir=state)
    assert c3.restored == 0 and c3.list_jobs() == []
    c3.stop()


def test_running_job_is_recreated_from_its_persisted_spec(tmp_path):
    from paddle_operator_b200.runtime.cluster import LocalCluster

    state = str(tmp_path / "state")
    c1 = LocalCluster(supervisor=False, init_image="", state_dir=state)
    c1.apply(os.path.join(EXAMPLES, "resnet.yaml"))
    c1.manager.run_until_idle()
    name = c1.list_jobs()[0].metadata.name
    n_pods = len(c1.store.list("Pod", "default"))
    assert n_pods >= 2
    c1.stop()
    c2 = LocalCluster(supervisor=False, init_image="", state_dir=state)
    assert c2.restored == 1 and c2.store.list("Pod", "default") == []  # children are not persisted ...
    c2.manager.enqueue("default", name)
    c2.manager.run_until_idle()
    assert len(c2.store.list("Pod", "default")) == n_pods             # ... the reconciler brings them back
    c2.stop()


@pytest.mark.parametrize("seed", range(12))
def test_random_operation_sequences_keep_the_invariants(seed):
    """Chaos test: random apply / scale / pod-exit / delete sequences against the fake kubelet.  After
    every settle: refs == child pods, the per-kind counters add up to the pods present, pod names are
    exactly <job>-<kind>-<idx> for idx < replicas (plus pods still being torn down), no child outlives its
    job, and the derived phase is one the reference can produce."""
    import random

    rnd = random.Random(seed)
    env = Env(batch=rnd.random() < 0.5, init_image=rnd.choice(["busybox", ""]))
    jobs = {}
    produced = {"", "Pending", "Starting", "Running", "Failed", "Completed"}

    def check():
        live = {o["metadata"]["name"] for o in env.store.list("PaddleJob")}
        for kind in ("Pod", "Service", "ConfigMap"):
            for o in env.store.list(kind):
                owners = [r["name"] for r in o["metadata"].get("ownerReferences") or []]
                assert owners and owners[0] in live, (kind, o["metadata"]["name"], owners)
        for name in live:
            job = env.job(name)
            pods = [p for p in env.store.list("Pod") if p["metadata"]["ownerReferences"][0]["name"] == name]
            assert job.status.phase in produced, job.status.phase
            total_refs = 0
            for kind in ("ps", "worker", "heter"):
                st, sp = getattr(job.status, kind), getattr(job.spec, kind)
                mine = [p for p in pods if p["metadata"]["labels"]["paddle-res-type"] == kind]
                if sp is None:
                    assert not mine
                    continue
                if st is None:
                    continue
                total_refs += len(st.refs)
                assert st.pending + st.starting + st.running + st.failed + st.succeeded <= len(st.refs)
                for p in mine:
                    n = p["metadata"]["name"]
                    assert n.startswith(f"{name}-{kind}-") and n.rsplit("-", 1)[1].isdigit()
            assert total_refs == len(pods), (name, total_refs, [p["metadata"]["name"] for p in pods])

    for step in range(40):
        op = rnd.choice(["apply", "scale", "scale", "exit", "finish", "delete", "settle"])
        names = list(jobs)
        if op == "apply" and len(jobs) < 3:
            name = f"j{seed}-{step}"
            ps = rnd.choice([None, 1, 2])
            env.apply(_job(name, ps=ps, worker=rnd.randint(1, 3), heter=rnd.choice([None, None, 1]),
                           cleanPodPolicy=rnd.choice(["", "Never", "Always", "OnFailure", "OnCompletion"]),
                           intranet=rnd.choice(["", "PodIP", "Service"])))
            jobs[name] = True
        elif op == "scale" and names:
            name = rnd.choice(names)
            obj = env.store.get("PaddleJob", "default", name)
            kind = rnd.choice([k for k in ("ps", "worker", "heter") if obj["spec"].get(k)])
            obj["spec"][kind]["replicas"] = rnd.randint(1, 4)
            env.store.update(obj)
        elif op == "exit" and names:
            pods = env.store.list("Pod")
            if pods:
                p = rnd.choice(pods)
                if (p.get("status") or {}).get("phase") == "Running":
                    env.sup.finish("default", p["metadata"]["name"], succeeded=rnd.random() < 0.7)
        elif op == "finish" and names:
            env.sup.finish_all(succeeded=True)
        elif op == "delete" and names:
            name = rnd.choice(names)
            env.store.delete("PaddleJob", "default", name)
            del jobs[name]
        env.settle()
        check()
    for name in list(jobs):
        env.store.delete("PaddleJob", "default", name)
    env.settle()
    assert env.store.list("PaddleJob") == [] and env.store.list("Pod") == [] and env.store.list("ConfigMap") == []


def test_elastic_jobs_scale_only_inside_requests_limits():
    """ResourceSpec.requests / limits (min / max replicas): stored and ignored for ordinary jobs like the reference,
    enforced as the scaling window of elastic jobs."""
    ok = _job("e", ps=None, worker=4, elastic=1)
    ok["spec"]["worker"].update(requests=2, limits=8)
    assert loader.paddlejob_from_doc(ok).spec.worker.limits == 8
    for bad in (1, 9):
        doc = _job("e", ps=None, worker=bad, elastic=1)
        doc["spec"]["worker"].update(requests=2, limits=8)
        with pytest.raises(loader.ValidationError):
            loader.paddlejob_from_doc(doc)
    plain = _job("p", ps=None, worker=9)
    plain["spec"]["worker"].update(requests=2, limits=8)
    assert loader.paddlejob_from_doc(plain).spec.worker.replicas == 9      # not elastic: as in the reference
    inverted = _job("i", ps=None, worker=3)
    inverted["spec"]["worker"].update(requests=5, limits=2)
    with pytest.raises(loader.ValidationError):
        loader.paddlejob_from_doc(inverted)


# ---- names are validated the way the API server validates them --------------------------------------------
@pytest.mark.parametrize("ns,name", [("..", "job"), ("default", "../evil"), ("default", "a/b"), ("default", "Job"),
                                     ("default", "a..b"), ("default", "-x"), ("Default", "job"), ("default", "x" * 64)])
def test_invalid_names_are_rejected_everywhere(ns, name):
    from paddle_operator_b200.api import loader
    from paddle_operator_b200.controller.store import Invalid, ObjectStore

    doc = {"apiVersion": "batch.paddlepaddle.org/v1", "kind": "PaddleJob", "metadata": {"name": name, "namespace": ns},
           "spec": {"worker": {"replicas": 1, "template": {"spec": {"containers": [{"name": "c", "image": "x"}]}}}}}
    with pytest.raises(loader.ValidationError):
        loader.paddlejob_from_doc(doc)
    if len(name) <= 63 or ns != "default":
        with pytest.raises(Invalid):
            ObjectStore().create({"kind": "Pod", "metadata": {"name": name, "namespace": ns}})


def test_supervisor_never_leaves_its_work_root(tmp_path):
    from paddle_operator_b200.controller.store import ObjectStore
    from paddle_operator_b200.runtime.supervisor import Supervisor

    sup = Supervisor(ObjectStore(), work_root=str(tmp_path / "root"), gpus=[])
    assert sup._contained("default", "job-worker-0") == str(tmp_path / "root" / "default" / "job-worker-0")
    assert sup._contained("..", "x") is None and sup._contained("default", "../../etc") is None
    outside = tmp_path / "secret"
    outside.mkdir()
    (outside / "log.txt").write_text("do not read me")
    assert sup.logs("..", "secret") == ""


def test_owner_index_and_event_policy():
    from paddle_operator_b200.controller.store import EventRecorder, ObjectStore

    st = ObjectStore()
    job = st.create({"kind": "PaddleJob", "metadata": {"name": "j", "namespace": "default"}, "spec": {}})
    ref = [{"kind": "PaddleJob", "name": "j", "uid": job["metadata"]["uid"], "controller": True}]
    for i in range(5):
        st.create({"kind": "Pod", "metadata": {"name": f"j-worker-{i}", "namespace": "default", "ownerReferences": ref}})
    st.create({"kind": "Pod", "metadata": {"name": "stranger", "namespace": "default"}})
    assert [p["metadata"]["name"] for p in st.list("Pod", "default", owner="j")] == [f"j-worker-{i}" for i in range(5)]
    assert ("PaddleJob", "default", "j") in st._by_owner  # served from the index, not from a scan
    rec = EventRecorder(st, max_per_object=4)
    for _ in range(10):
        rec.event(job, "Normal", "Created", "pod created")  # a repeat: one object, count 10
    evs = rec.events_for("PaddleJob", "default", "j")
    assert len(evs) == 1 and evs[0]["count"] == 10
    for i in range(10):
        rec.event(job, "Normal", "Scaled", f"to {i}")
    assert len(rec.events_for("PaddleJob", "default", "j")) <= 4  # capped per object
    st.delete("PaddleJob", "default", "j")
    assert st.list("Event", "default") == [] and st.list("Pod", "default", owner="j") == []  # collected with the job
    assert ("PaddleJob", "default", "j") not in st._by_owner


def test_legacy_hostport_manager_runs_as_a_controller():
    """third_party/hostport-allocator: watch -> queue -> allocate N ports -> annotate; re-mark on restart; release on delete."""
    from paddle_operator_b200.controller.hostport import (LEGACY_HOSTPORT_ANNOTATION, LEGACY_PORTNUM_ANNOTATION,
                                                          LegacyHostPortManager, parse_annotation)
    from paddle_operator_b200.controller.store import ObjectStore

    st = ObjectStore()
    st.create({"kind": "TrainingJob", "metadata": {"name": "old", "annotations": {LEGACY_PORTNUM_ANNOTATION: "2",
                                                                                   LEGACY_HOSTPORT_ANNOTATION: "35001,35002"}}})
    mgr = LegacyHostPortManager(st, start=35000, end=35010)
    mgr.start()
    try:
        assert mgr.alloc.is_allocated(35001) and mgr.alloc.is_allocated(35002)  # re-marked from the annotation
        st.create({"kind": "TrainingJob", "metadata": {"name": "a", "annotations": {LEGACY_PORTNUM_ANNOTATION: "3"}}})
        st.create({"kind": "TrainingJob", "metadata": {"name": "plain"}})  # no portnum: ignored
        mgr.drain()
        ports = parse_annotation(st.get("TrainingJob", "default", "a")["metadata"]["annotations"][LEGACY_HOSTPORT_ANNOTATION])
        assert len(ports) == 3 and len(set(ports)) == 3 and not {35001, 35002} & set(ports)
        assert LEGACY_HOSTPORT_ANNOTATION not in (st.get("TrainingJob", "default", "plain")["metadata"].get("annotations") or {})
        st.delete("TrainingJob", "default", "a")
        assert not any(mgr.alloc.is_allocated(p) for p in ports)  # released by the delete handler
        st.create({"kind": "TrainingJob", "metadata": {"name": "big", "annotations": {LEGACY_PORTNUM_ANNOTATION: "50"}}})
        mgr.drain()
        assert LEGACY_HOSTPORT_ANNOTATION not in st.get("TrainingJob", "default", "big")["metadata"]["annotations"]
        assert not mgr.alloc.is_allocated(35005)  # all-or-nothing: nothing leaked by the failed request
    finally:
        mgr.stop()
