"""Elastic rescale harness (BASELINE.json config 4): runs the elastic ResNet
PaddleJob on this box through the whole control plane - etcd-compatible store,
reconciler, supervisor - scales it A -> B -> A with `scale`, and reports the
recovery time of every rescale and the throughput of every phase.

    python -m paddle_operator_b200.workloads.elastic_demo --sizes 8,4,8 --steps-per-phase 40

"Recovery seconds" of a rescale = wall time from the scale request to the first
progress report of the new generation (it includes controller latency, pod
start-up for joiners, the stop-step agreement, re-forming the symmetric-memory
group, the GPU-to-GPU state broadcast and re-capturing the step graph).
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import tempfile
import time


from ..runtime.cluster import LocalCluster
from ..runtime.etcd_client import EtcdClient
from ..runtime.etcd_server import EtcdServer

JOB = "resnet-elastic"


def job_doc(replicas: int, args_list: list, gpu: bool) -> dict:
    c = {"name": "resnet", "image": "local/pdjob-b200:resnet-elastic", "command": ["python"],
         "args": ["-m", "paddle_operator_b200.workloads.train_elastic_resnet"] + args_list,
         "volumeMounts": [{"mountPath": "/checkpoint", "name": "chkpt"}]}
    if gpu:
        c["resources"] = {"limits": {"nvidia.com/gpu": 1}}
    return {"apiVersion": "batch.paddlepaddle.org/v1", "kind": "PaddleJob", "metadata": {"name": JOB},
            "spec": {"cleanPodPolicy": "Never", "elastic": 1,
                     "worker": {"replicas": replicas, "requests": 1, "limits": 8,
                                "template": {"spec": {"terminationGracePeriodSeconds": 30, "containers": [c],
                                                      "volumes": [{"name": "chkpt", "hostPath": {"path": "CKPT"}}]}}}}}


def _reports(etcd, prefix: str) -> dict:
    out = {}
    for k, v in etcd.get_prefix(prefix + "/report/"):
        _, _, rest = k.partition("/report/")
        host, _, key = rest.partition("/")
        out.setdefault(host, {})[key] = json.loads(v)
    return out


def _wait_progress(etcd, prefix: str, world: int, min_step: int, timeout: float, after_gen: int = -1):
    """Waits until every one of `world` ranks reports progress at world size `world`, generation >
    after_gen and step >= min_step.  Returns (time of the earliest such report of the slowest rank, reports)."""
    deadline = time.time() + timeout
    while time.time() < deadline:
        rep = _reports(etcd, prefix)
        ok = [r["progress"] for r in rep.values() if "progress" in r and r["progress"]["world"] == world
              and r["progress"]["gen"] > after_gen and r["progress"]["step"] >= min_step]
        if len(ok) >= world:
            return rep
        time.sleep(0.02)
    raise TimeoutError(f"no progress at world {world} within {timeout}s: {json.dumps(_reports(etcd, prefix))[:600]}")


def _wait_standbys(cluster, lo: int, hi: int, timeout: float) -> float:
    """Waits until the parked ranks of worker indices lo .. hi-1 have finished booting (their log says so);
    returns the seconds waited."""
    t0 = time.time()
    want = {f"{JOB}-worker-{i}" for i in range(lo, hi)}
    while time.time() - t0 < timeout:
        ready = {name for ns, name, _ in cluster.supervisor.standbys()
                 if name in want and '"event": "standby"' in cluster.supervisor.logs(ns, name)}
        if ready == want:
            break
        time.sleep(0.05)
    return round(time.time() - t0, 3)


def run(sizes: list, steps_per_phase: int, dry: bool, batch: int, tiny: bool, gpus, work_root: str, timeout: float,
        fail_rank: int = -1, standby: bool = True) -> dict:
    ckpt_dir = os.path.join(work_root, "checkpoint")
    os.makedirs(ckpt_dir, exist_ok=True)
    total_steps = steps_per_phase * len(sizes) * 50  # the harness ends the job itself
    rank_args = ["--steps", str(total_steps), "--batch", str(batch), "--report-every", "2", "--ckpt-every", "20"]
    if tiny:
        rank_args += ["--tiny", "--image", "64"]
    if dry:
        rank_args += ["--cpu-dry-run"]
    if fail_rank >= 0:
        rank_args += ["--fail-rank", str(fail_rank), "--fail-at-step", str(steps_per_phase // 2)]
    doc = job_doc(sizes[0], rank_args, gpu=not dry)
    doc["spec"]["worker"]["template"]["spec"]["volumes"][0]["hostPath"]["path"] = ckpt_dir
    result = {"sizes": sizes, "phases": [], "rescales": []}
    with EtcdServer("127.0.0.1", 0) as srv:
        etcd = EtcdClient(srv.endpoint)
        cluster = LocalCluster(etcd=etcd, gpus=gpus, work_root=os.path.join(work_root, "pods"),
                               extra_env={"PDJ_ELASTIC_TTL": "3"})
        cluster.supervisor.warm_standby = standby
        result["warm_standby"] = standby
        prefix = f"/paddle/default-{JOB}"
        with cluster:
            t_apply = time.time()
            cluster.apply(doc)
            gen_seen = -1
            cur_step = 0
            for i, n in enumerate(sizes):
                if i > 0 and standby and n > sizes[i - 1]:
                    # a scale-up request that arrives while the parked ranks are still booting would measure their
                    # boot, not the rescale: give them time (an operator scales up minutes after a scale-down)
                    result.setdefault("standby_wait_seconds", []).append(
                        _wait_standbys(cluster, sizes[i - 1], min(n, 8), min(timeout, 60.0)))
                t_req = time.time()
                if i > 0:
                    cluster.scale(JOB, n)
                rep = _wait_progress(etcd, prefix, n, cur_step + 2, timeout, after_gen=gen_seen)
                t_first = min(max(r["progress"]["t"] for r in rep.values() if "progress" in r
                                  and r["progress"]["world"] == n and r["progress"]["gen"] > gen_seen), time.time())
                new_gen = max(r["progress"]["gen"] for r in rep.values() if "progress" in r and r["progress"]["world"] == n)
                forms = [r[f"formed/{new_gen}"] for r in rep.values() if f"formed/{new_gen}" in r]
                if i > 0:
                    stops = [r[f"stopped/{gen_seen}"] for r in rep.values() if f"stopped/{gen_seen}" in r]
                    result["rescales"].append({
                        "from": sizes[i - 1], "to": n, "recovery_seconds": round(t_first - t_req, 3),
                        "stop_step": stops[0]["step"] if stops else None,
                        "resume_step": forms[0]["step"] if forms else None,
                        "request_to_stop_seconds": round(min(s["t"] for s in stops) - t_req, 3) if stops else None,
                        "form_seconds_max": max(f["form_seconds"] for f in forms) if forms else None,
                        # joiners that were parked warm standbys (supervisor) rather than fresh processes
                        "adopted_standbys": sum(1 for p in cluster.pods(JOB) if (p.get("status") or {}).get("adoptedStandby")),
                    })
                else:
                    result["startup_seconds"] = round(t_first - t_apply, 3)
                gen_seen = new_gen
                # let the phase run, then take its throughput from the ranks' own reports
                start_step = max(r["progress"]["step"] for r in rep.values() if "progress" in r and r["progress"]["world"] == n)
                rep = _wait_progress(etcd, prefix, n, start_step + steps_per_phase, timeout, after_gen=gen_seen - 1)
                pr = [r["progress"] for r in rep.values() if "progress" in r and r["progress"]["world"] == n
                      and r["progress"]["gen"] >= gen_seen]
                if max(p["gen"] for p in pr) != gen_seen:  # the job re-formed on its own (a rank failed and came back)
                    result.setdefault("unplanned_reforms", []).append({"world": n, "from_gen": gen_seen,
                                                                       "to_gen": max(p["gen"] for p in pr)})
                    gen_seen = max(p["gen"] for p in pr)
                    pr = [p for p in pr if p["gen"] == gen_seen]
                cur_step = max(p["step"] for p in pr)
                result["phases"].append({"world": n, "gen": gen_seen, "step": cur_step,
                                         "img_per_s": round(min(p["img_per_s"] for p in pr), 1),
                                         "loss": pr[0]["loss"]})
            job = cluster.get(JOB)
            result["job_phase"] = job.status.phase
            result["final_step"] = cur_step
            result["events"] = [e["reason"] + ": " + e["message"] for e in cluster.events(JOB)
                                if e["reason"] == "Scaled"]
            cluster.delete(JOB)
            cluster.wait_deleted(JOB, timeout=60)
    return result


def run_failure(n: int, fail_rank: int, fail_step: int, batch: int, tiny: bool, gpus, work_root: str, timeout: float,
                dry: bool = False, spin_timeout_ms: int = 2000) -> dict:
    """Fault injection: rank `fail_rank` SIGKILLs itself at `fail_step`.  On GPUs its peers are
    inside a collective, trap after the spin time-out and exit; the supervisor restarts every rank
    (restartPolicy OnFailure), the job re-forms as a fresh generation and resumes from /checkpoint.
    Reports the seconds from the last progress of the failed generation to the first progress of the
    new one, and the steps that had to be redone."""
    ckpt_dir = os.path.join(work_root, "checkpoint")
    os.makedirs(ckpt_dir, exist_ok=True)
    rank_args = ["--steps", "100000", "--batch", str(batch), "--report-every", "1", "--ckpt-every", "10",
                 "--fail-rank", str(fail_rank), "--fail-at-step", str(fail_step)]
    if tiny:
        rank_args += ["--tiny", "--image", "64"]
    if dry:
        rank_args += ["--cpu-dry-run"]
    doc = job_doc(n, rank_args, gpu=not dry)
    doc["spec"]["worker"]["template"]["spec"]["volumes"][0]["hostPath"]["path"] = ckpt_dir
    with EtcdServer("127.0.0.1", 0) as srv:
        etcd = EtcdClient(srv.endpoint)
        cluster = LocalCluster(etcd=etcd, gpus=gpus, work_root=os.path.join(work_root, "pods"),
                               extra_env={"PDJ_ELASTIC_TTL": "3", "PDJ_SPIN_TIMEOUT_MS": str(spin_timeout_ms)})
        prefix = f"/paddle/default-{JOB}"
        with cluster:
            cluster.apply(doc)
            last_old, first_new = None, None
            deadline = time.time() + timeout
            while time.time() < deadline and first_new is None:
                for r in _reports(etcd, prefix).values():
                    p = r.get("progress")
                    if not p or p["world"] != n:
                        continue
                    if p["gen"] == 0 and (last_old is None or p["t"] > last_old["t"]):
                        last_old = p
                    if p["gen"] > 0 and p["step"] > 0:
                        first_new = p
                time.sleep(0.02)
            restarts = sum((p["status"].get("containerStatuses") or [{}])[0].get("restartCount", 0)
                           for p in cluster.pods(JOB))
            phase = cluster.get(JOB).status.phase
            cluster.delete(JOB)
            cluster.wait_deleted(JOB, timeout=60)
    if first_new is None or last_old is None:
        raise TimeoutError("the job did not recover from the injected failure")
    return {"world": n, "failed_rank": fail_rank, "failed_at_step": fail_step, "last_step_before": last_old["step"],
            "resumed_generation": first_new["gen"], "resumed_at_step": first_new["step"],
            "steps_redone": max(0, last_old["step"] - first_new["step"] + 1),
            "recovery_seconds": round(first_new["t"] - last_old["t"], 3), "spin_timeout_ms": spin_timeout_ms,
            "rank_restarts": restarts, "job_phase_after": phase}


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", default="8,4,8")
    ap.add_argument("--steps-per-phase", type=int, default=40)
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--tiny", action="store_true")
    ap.add_argument("--cpu-dry-run", action="store_true")
    ap.add_argument("--timeout", type=float, default=240.0)
    ap.add_argument("--out", default="")
    ap.add_argument("--fail-rank", type=int, default=-1, help="fault injection instead of a rescale: rank to kill")
    ap.add_argument("--fail-step", type=int, default=25)
    ap.add_argument("--no-standby", action="store_true", help="scale-ups boot fresh processes (the round-1 behaviour)")
    args = ap.parse_args()
    sizes = [int(x) for x in args.sizes.split(",")]
    gpus = None if not args.cpu_dry_run else list(range(max(sizes)))
    work = tempfile.mkdtemp(prefix="pdj-elastic-")
    try:
        if args.fail_rank >= 0:
            res = run_failure(sizes[0], args.fail_rank, args.fail_step, args.batch, args.tiny, gpus, work, args.timeout,
                              dry=args.cpu_dry_run)
        else:
            res = run(sizes, args.steps_per_phase, args.cpu_dry_run, args.batch, args.tiny, gpus, work, args.timeout,
                      standby=not args.no_standby)
    except Exception:
        # post-mortem: every rank's log
        import glob

        for f in sorted(glob.glob(os.path.join(work, "pods", "*", "*", "log.txt"))):
            print(f"===== {f}", flush=True)
            with open(f, errors="replace") as fh:
                print(fh.read()[-3000:], flush=True)
        raise
    print(json.dumps(res), flush=True)
    if args.out:
        with open(args.out, "w") as f:
            json.dump(res, f, indent=1)
    return 0


if __name__ == "__main__":
    sys.exit(main())
