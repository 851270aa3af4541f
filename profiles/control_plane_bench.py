"""Control-plane timings on CPU (no GPU, no rank processes): the store + reconciler + manager against the scriptable
fake kubelet.  What the reference would spend between `kubectl apply` and "all pods created / released", where
its own pace is set by one mutation per reconcile and 1 s requeues (controllers/paddlejob_controller.go:113-327).

    python profiles/control_plane_bench.py            # prints a table and one JSON line
"""
import json
import statistics
import sys
import time

sys.path.insert(0, ".")
from paddle_operator_b200.controller.manager import ControllerManager  # noqa: E402
from paddle_operator_b200.controller.reconciler import PaddleJobReconciler  # noqa: E402
from paddle_operator_b200.controller.store import EventRecorder, ObjectStore  # noqa: E402
from paddle_operator_b200.runtime.fake_supervisor import FakeSupervisor  # noqa: E402


def job(name, ps, worker):
    tmpl = {"spec": {"containers": [{"name": "paddle", "image": "img"}]}}
    d = {"apiVersion": "batch.paddlepaddle.org/v1", "kind": "PaddleJob", "metadata": {"name": name}, "spec": {}}
    if ps:
        d["spec"]["ps"] = {"replicas": ps, "template": json.loads(json.dumps(tmpl))}
    d["spec"]["worker"] = {"replicas": worker, "template": json.loads(json.dumps(tmpl))}
    return d


class Env:
    def __init__(self, batch):
        self.store = ObjectStore()
        self.sup = FakeSupervisor(self.store)
        self.r = PaddleJobReconciler(self.store, EventRecorder(self.store), init_image="busybox", batch=batch,
                                     exec_in_pod=self.sup.exec_in_pod)
        self.mgr = ControllerManager(self.store, self.r, None)

    def settle(self):
        passes = 0
        for _ in range(200):
            n = self.mgr.run_until_idle()
            passes += n
            moved = self.sup.schedule_all() + self.sup.start_ready()
            for o in self.store.list("PaddleJob"):
                self.mgr.enqueue(o["metadata"]["namespace"], o["metadata"]["name"])
            if moved == 0 and n <= len(self.store.list("PaddleJob")) + 1:
                break
        return passes

    def phase(self, name):
        return (self.store.get("PaddleJob", "default", name).get("status") or {}).get("phase")


def timed(fn, reps=5):
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        out = fn()
        ts.append((time.perf_counter() - t0) * 1e3)
    return statistics.median(ts), out


def bring_up(batch, ps, worker):
    def run():
        e = Env(batch)
        e.store.create(job("j", ps, worker))
        passes = e.settle()
        assert e.phase("j") == "Running", e.phase("j")
        return passes
    return timed(run)


def rescale(batch, frm, to):
    def run():
        e = Env(batch)
        e.store.create(job("j", 0, frm))
        e.settle()
        obj = e.store.get("PaddleJob", "default", "j")
        obj["spec"]["worker"]["replicas"] = to
        t0 = time.perf_counter()
        e.store.update(obj)
        e.settle()
        dt = (time.perf_counter() - t0) * 1e3
        assert len(e.store.list("Pod")) == to and e.phase("j") == "Running"
        return dt
    return statistics.median([run() for _ in range(5)])


def many(n):
    def run():
        e = Env(True)
        for i in range(n):
            e.store.create(job(f"j{i}", 2, 2))
        e.settle()
        assert all(e.phase(f"j{i}") == "Running" for i in range(n))
        return len(e.store.list("Pod"))
    return timed(run, reps=3)


def main():
    out = {}
    for batch in (True, False):
        ms, passes = bring_up(batch, 2, 2)
        out[f"bring_up_2ps_2w_batch{int(batch)}"] = {"ms": round(ms, 2), "reconcile_passes": passes}
        ms, passes = bring_up(batch, 0, 8)
        out[f"bring_up_8w_batch{int(batch)}"] = {"ms": round(ms, 2), "reconcile_passes": passes}
    out["rescale_8_to_4_ms"] = round(rescale(True, 8, 4), 2)
    out["rescale_4_to_8_ms"] = round(rescale(True, 4, 8), 2)
    ms, pods = many(100)
    out["100_jobs_2ps_2w"] = {"ms": round(ms, 1), "pods": pods}
    for k, v in out.items():
        print(f"{k:32s} {v}")
    print(json.dumps(out))


if __name__ == "__main__":
    main()
