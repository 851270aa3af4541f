"""Rank supervisor: the kubelet + scheduler + device plugin of a single box.

The reference creates Pod objects and leaves running them to Kubernetes.  Here
every Pod in the object store becomes one local process:

  * scheduling   - assigns a loopback address per rank (`127.x.y.z`, so the
                   reference's "<ip>:2379" endpoints stay collision-free and
                   pass its dotted-quad check, controllers/paddlejob_helper.go:
                   226-228) and GPUs from the local pool according to
                   `resources.limits["nvidia.com/gpu"]`
                   (deploy/examples/resnet.yaml:22-24) -> CUDA_VISIBLE_DEVICES;
  * start gate   - a pod with the `coord-paddle` init container stays Pending
                   with that container "running" until a file named `goon`
                   appears in its work directory - the reference's coord
                   container spins on exactly that file
                   (controllers/paddlejob_helper.go:39-41) and the reconciler
                   creates it with exec `touch goon` (controller:313);
  * environment  - containers[0].env (+ downward-API fieldRefs) and envFrom the
                   job ConfigMap; a missing ConfigMap keeps the container
                   waiting, as CreateContainerConfigError does (faq.md:18-26);
  * lifecycle    - exit 0 -> Succeeded, else restartPolicy Never -> Failed,
                   OnFailure/Always -> restarted in place with back-off
                   (controllers/paddlejob_helper.go:366-374 chooses the policy);
  * volumes      - hostPath volumes map to their path, emptyDir to a scratch
                   directory (medium Memory -> /dev/shm), a persistentVolumeClaim
                   to <work_root>/pvc/<namespace>/<claimName> shared by all pods
                   that mount the claim; each mountPath is
                   exported as PDJ_MOUNT_<path> (deploy/elastic/resnet.yaml:
                   25-36: /checkpoint <- hostPath /tmp/checkpoint);
  * deletion     - a deleted Pod's process group is terminated.

There are no container images here: `image` is only a key into IMAGE_COMMANDS
for templates that give no command (deploy/examples/wide_and_deep.yaml has
none), and `python` in a command is this interpreter.
"""
from __future__ import annotations

import fnmatch
import logging
import os
import shlex
import shutil
import signal
import subprocess
import sys
import threading
import time
from dataclasses import dataclass, field
from typing import Optional

from ..controller import helpers as H
from ..controller.store import NotFound, ObjectStore

log = logging.getLogger("pdjob-supervisor")

# image glob -> argv, for templates without a command.  The reference's demo
# images map onto this repository's workloads.
IMAGE_COMMANDS: dict[str, list[str]] = {
    "*demo-wide-and-deep*": ["python", "-m", "paddle_operator_b200.workloads.train_wide_deep"],
    "*wide-and-deep*": ["python", "-m", "paddle_operator_b200.workloads.train_wide_deep"],
    "*deepfm*": ["python", "-m", "paddle_operator_b200.workloads.train_wide_deep"],
    "*demo-resnet:elastic*": ["python", "-m", "paddle_operator_b200.workloads.train_elastic_resnet"],
    "*resnet-elastic*": ["python", "-m", "paddle_operator_b200.workloads.train_elastic_resnet"],
    "*resnet*": ["python", "-m", "paddle_operator_b200.workloads.train_resnet"],
    "*env-dump*": ["python", "-m", "paddle_operator_b200.workloads.env_dump"],
}
# scripts of the reference's examples -> local modules (python -m paddle.distributed.launch <script>)
SCRIPT_MODULES = {
    "train_fleet.py": "paddle_operator_b200.workloads.train_resnet",
    "train_fleet_dygraph_ckpt.py": "paddle_operator_b200.workloads.train_elastic_resnet",
}


def resolve_command(container: dict) -> Optional[list[str]]:
    # Kubernetes semantics: `command` replaces the image's entrypoint, `args` are appended to
    # whichever of the two is in effect.
    cmd = list(container.get("command") or [])
    if not cmd:
        image = container.get("image", "")
        for pat, argv in IMAGE_COMMANDS.items():
            if fnmatch.fnmatch(image, pat):
                cmd = list(argv)
                break
    if not cmd:
        return None
    cmd += [str(a) for a in container.get("args") or []]
    # `python -m paddle.distributed.launch script.py ...` -> the local workload for that script
    if len(cmd) >= 4 and cmd[1] == "-m" and cmd[2] == "paddle.distributed.launch":
        script = os.path.basename(cmd[3])
        if script in SCRIPT_MODULES:
            cmd = [cmd[0], "-m", SCRIPT_MODULES[script]] + cmd[4:]
    if os.path.basename(cmd[0]) in ("python", "python3"):
        cmd[0] = sys.executable
    return cmd


def mount_env_name(mount_path: str) -> str:
    return "PDJ_MOUNT_" + "".join(c if c.isalnum() else "_" for c in mount_path)


def resolve_mount(mount_path: str, default: Optional[str] = None) -> str:
    """For rank programs: where a container mountPath lives on this host."""
    return os.environ.get(mount_env_name(mount_path), default if default is not None else mount_path)


@dataclass
class _Rank:
    namespace: str
    name: str
    uid: str
    workdir: str
    ip: str
    gpus: list = field(default_factory=list)
    proc: Optional[subprocess.Popen] = None
    restarts: int = 0
    next_start: float = 0.0
    gate_open: bool = False
    log_path: str = ""
    started: bool = False
    done: bool = False
    kill_at: float = 0.0   # monotonic deadline after which a terminating rank is SIGKILLed
    grace: float = 30.0    # terminationGracePeriodSeconds


@dataclass
class _Standby:
    """A rank of an elastic job started ahead of its pod (see Supervisor._sync_standbys)."""
    namespace: str
    name: str          # the name the pod will have: <job>-worker-<idx>
    job_uid: str
    workdir: str
    ip: str
    gpus: list
    proc: subprocess.Popen
    gate: str          # file whose appearance lets the parked process register and join
    log_path: str
    host_network: bool = False


class Supervisor:
    def __init__(self, store: ObjectStore, work_root: str = "/tmp/pdjob", gpus: Optional[list] = None,
                 isolate_gpus: bool = True, poll: float = 0.05, extra_env: Optional[dict] = None,
                 backoff: float = 0.5, backoff_max: float = 10.0, warm_standby: bool = True):
        self.store = store
        self.work_root = work_root
        os.makedirs(work_root, exist_ok=True)
        self.gpu_pool = list(gpus) if gpus is not None else self._detect_gpus()
        self.isolate_gpus = isolate_gpus
        self.poll = poll
        self.extra_env = dict(extra_env or {})
        self.backoff, self.backoff_max = backoff, backoff_max
        self._ranks: dict[tuple, _Rank] = {}
        # warm standbys of elastic jobs, keyed like ranks; PDJ_WARM_STANDBY=0 turns the feature off
        self.warm_standby = warm_standby and os.environ.get("PDJ_WARM_STANDBY", "1") != "0"
        self._standbys: dict[tuple, _Standby] = {}
        # builds the pod the reconciler would create for (job, kind, index) - the reconciler's own constructor, wired
        # in by LocalCluster; without it there are no standbys (the env of a parked rank must match its pod's)
        self.pod_factory = None
        self._gpu_owner: dict[int, tuple] = {}
        self._gpu_starved = False
        self._ip_next = 1
        self._lock = threading.RLock()
        self._stop = threading.Event()
        self._wake = threading.Event()
        self._thread: Optional[threading.Thread] = None
        self._cancel = store.watch(self._on_event)

    @staticmethod
    def _detect_gpus() -> list:
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        if vis is not None:
            return [int(x) for x in vis.split(",") if x.strip().isdigit()]
        try:
            out = subprocess.run(["nvidia-smi", "-L"], capture_output=True, text=True, timeout=10).stdout
            return list(range(sum(1 for l in out.splitlines() if l.startswith("GPU "))))
        except (OSError, subprocess.SubprocessError):
            return []

    # -- plumbing -------------------------------------------------------------
    def _on_event(self, typ: str, obj: dict) -> None:
        if obj.get("kind") in ("Pod", "ConfigMap", "PaddleJob"):
            self._wake.set()

    def _alloc_ip(self, host_network: bool) -> str:
        if host_network:
            return "127.0.0.1"
        n = self._ip_next
        self._ip_next += 1
        return f"127.{1 + (n >> 16) % 254}.{(n >> 8) & 255}.{n & 255}"

    def _alloc_gpus(self, key: tuple, n: int) -> Optional[list]:
        free = [g for g in self.gpu_pool if g not in self._gpu_owner]
        if len(free) < n:
            return None
        got = free[:n]
        for g in got:
            self._gpu_owner[g] = key
        return got

    def _free_gpus(self, key: tuple) -> None:
        for g in [g for g, k in self._gpu_owner.items() if k == key]:
            del self._gpu_owner[g]

    def hosts(self) -> dict:
        """name -> address table (what cluster DNS provides in Service mode)."""
        with self._lock:
            return {r.name: r.ip for r in self._ranks.values()}

    def _write_hosts(self) -> None:
        path = os.path.join(self.work_root, "hosts")
        with open(path + ".tmp", "w") as f:
            for name, ip in sorted(self.hosts().items()):
                f.write(f"{ip} {name}\n")
        os.replace(path + ".tmp", path)

    # -- the pods/exec endpoint ---------------------------------------------------
    def exec_in_pod(self, namespace: str, pod_name: str, container: str, cmd: list) -> int:
        with self._lock:
            r = self._ranks.get((namespace, pod_name))
        if r is None:
            return 1
        if len(cmd) == 2 and cmd[0] == "touch":
            open(os.path.join(r.workdir, cmd[1]), "a").close()
            self._wake.set()
            return 0
        return subprocess.run(cmd, cwd=r.workdir, timeout=3).returncode

    # -- reconcile local processes with the store ------------------------------------
    def sync(self) -> None:
        with self._lock:
            self._gpu_starved = False
            pods = {(p["metadata"]["namespace"], p["metadata"]["name"]): p for p in self.store.list("Pod")}
            for key in list(self._ranks):
                r = self._ranks[key]
                p = pods.get(key)
                if p is None or p["metadata"].get("uid") != r.uid or p["metadata"].get("deletionTimestamp"):
                    # Graceful termination without blocking this loop: SIGTERM now, SIGKILL once the
                    # pod's terminationGracePeriodSeconds (default 30, as in Kubernetes) has passed.
                    alive = r.proc is not None and r.proc.poll() is None
                    if alive and r.kill_at == 0.0:
                        try:
                            os.killpg(r.proc.pid, signal.SIGTERM)
                        except (ProcessLookupError, PermissionError):
                            pass
                        r.kill_at = time.monotonic() + r.grace
                        continue
                    if alive and time.monotonic() < r.kill_at:
                        continue
                    if alive:
                        self._kill(r, grace=0.0)
                    self._free_gpus(key)
                    del self._ranks[key]  # the work directory (log.txt) is kept for `pdj logs`
            for key, pod in pods.items():
                if pod["metadata"].get("deletionTimestamp"):
                    continue
                old = self._ranks.get(key)
                if old is not None and old.uid != pod["metadata"].get("uid"):
                    continue  # a previous incarnation of this name is still terminating
                try:
                    self._sync_pod(key, pod)
                except NotFound:
                    continue
            if self.warm_standby or self._standbys:
                self._sync_standbys(pods)

    # -- warm standby ranks of elastic jobs --------------------------------------------------------------
    def _sync_standbys(self, pods: dict) -> None:
        """For every running elastic job, keeps the ranks it may still grow into (worker indices replicas ..
        limits - 1) started and parked while GPUs are free.  A parked process has the interpreter, the CUDA
        context and the model up and waits on its gate file without registering as a member; when `scale` makes
        the reconciler create that pod, _sync_pod adopts the process instead of booting one.  Measured on the
        8 -> 4 -> 8 rescale of ResNet-50: 8.8 of the 10.4 s of a scale-up were four fresh processes booting.
        Standbys never compete with real pods: one is evicted whenever a pod finds no free GPU."""
        from ..api.types import PaddleJob, PaddleJobPhase

        want: dict[tuple, tuple] = {}   # ranks to keep parked: indices replicas .. limits - 1
        keep: dict[tuple, str] = {}     # ranks that may stay parked: every index below limits (job uid)
        if self.warm_standby and self.pod_factory is not None:
            for obj in self.store.list("PaddleJob"):
                try:
                    pdj = PaddleJob.from_dict(obj)
                except (ValueError, KeyError, TypeError):
                    continue
                w = pdj.spec.worker
                if (pdj.spec.elastic is None or w is None or not w.limits or obj["metadata"].get("deletionTimestamp")
                        or pdj.status.phase != PaddleJobPhase.RUNNING.value):
                    continue
                for idx in range(int(w.limits)):
                    key = (pdj.namespace, H.gen_paddle_res_name(pdj.name, "worker", idx))
                    # between `scale` raising replicas and the reconciler creating the pod, the rank is neither
                    # wanted as a standby nor a pod yet: it must survive that window to be adopted
                    keep[key] = pdj.metadata.uid
                    if idx >= int(w.replicas):
                        want[key] = (pdj, idx)
        for key in list(self._standbys):
            sb = self._standbys[key]
            dead = sb.proc.poll() is not None
            if key in pods:
                continue  # its pod exists: _sync_pod adopts it (or has done so)
            if dead or keep.get(key) != sb.job_uid:
                self._drop_standby(key)
        for key, (pdj, idx) in want.items():
            if key in pods or key in self._standbys or key in self._ranks or self._gpu_starved:
                continue
            try:
                pod = self.pod_factory(pdj, "worker", idx)  # exactly the pod the reconciler will create
            except ValueError:
                continue
            need = H.gpus_requested(pod["spec"])
            if need > len([g for g in self.gpu_pool if g not in self._gpu_owner]):
                continue
            self._spawn_standby(key, pdj, pod, need)

    def _spawn_standby(self, key: tuple, pdj, pod: dict, need: int) -> None:
        ns, name = key
        c0 = pod["spec"]["containers"][0]
        cmd = resolve_command(c0)
        workdir = self._contained(ns, name)
        if cmd is None or workdir is None:
            return
        skey = ("standby",) + key
        gpus = self._alloc_gpus(skey, need) if need else []
        if gpus is None:
            return
        shutil.rmtree(workdir, ignore_errors=True)
        os.makedirs(workdir, exist_ok=True)
        host_net = bool(pod["spec"].get("hostNetwork"))
        r = _Rank(ns, name, "", workdir, self._alloc_ip(host_net), gpus or [])
        r.log_path = os.path.join(workdir, "log.txt")
        env = self._build_env(r, pod)
        if env is None:
            self._free_gpus(skey)
            return
        gate = os.path.join(workdir, "standby-go")
        env["PDJ_STANDBY_GATE"] = gate
        with open(r.log_path, "ab") as logf:
            logf.write(f"--- standby start: {shlex.join(cmd)}\n".encode())
            logf.flush()
            try:
                proc = subprocess.Popen(cmd, cwd=c0.get("workingDir") or workdir, env=env, stdout=logf,
                                        stderr=subprocess.STDOUT, start_new_session=True)
            except OSError:
                self._free_gpus(skey)
                return
        self._standbys[key] = _Standby(ns, name, pdj.metadata.uid, workdir, r.ip, list(gpus or []), proc, gate, r.log_path, host_net)

    def _drop_standby(self, key: tuple) -> None:
        sb = self._standbys.pop(key, None)
        if sb is None:
            return
        if sb.proc.poll() is None:
            try:
                os.killpg(sb.proc.pid, signal.SIGKILL)  # it holds no state worth a graceful exit
            except (ProcessLookupError, PermissionError):
                pass
            try:
                sb.proc.wait(2)
            except subprocess.TimeoutExpired:
                pass
        self._free_gpus(("standby",) + key)

    def _adopt_standby(self, key: tuple, pod: dict) -> bool:
        """The pod of a parked rank has been created: hand the process over and open its gate."""
        sb = self._standbys.get(key)
        if sb is None:
            return False
        owner = H.controller_of(pod) or {}
        if sb.proc.poll() is not None or (owner.get("uid") and owner.get("uid") != sb.job_uid) or \
                H.gpus_requested(pod["spec"]) != len(sb.gpus) or bool(pod["spec"].get("hostNetwork")) != sb.host_network:
            self._drop_standby(key)
            return False
        del self._standbys[key]
        ns, name = key
        for g in sb.gpus:
            self._gpu_owner[g] = key
        r = _Rank(ns, name, pod["metadata"].get("uid", ""), sb.workdir, sb.ip, list(sb.gpus))
        r.proc, r.log_path, r.started, r.gate_open = sb.proc, sb.log_path, True, True
        r.grace = float(pod["spec"].get("terminationGracePeriodSeconds", 30))
        self._ranks[key] = r
        self._write_hosts()
        open(sb.gate, "a").close()
        c0 = pod["spec"]["containers"][0]
        inits = [{"name": c["name"], "ready": True, "restartCount": 0,
                  "state": {"terminated": {"exitCode": 0, "reason": "Completed", "finishedAt": H.now_rfc3339()}}}
                 for c in pod["spec"].get("initContainers") or []]
        cs = [{"name": c0.get("name", "main"), "ready": True, "restartCount": 0,
               "state": {"running": {"startedAt": H.now_rfc3339()}}, "pid": sb.proc.pid}]

        def fn(st, ip=sb.ip, gpus=list(sb.gpus)):
            st.update({"phase": "Running", "podIP": ip, "hostIP": "127.0.0.1", "containerStatuses": cs,
                       "conditions": [{"type": "PodScheduled", "status": "True"}],
                       "startTime": st.get("startTime") or H.now_rfc3339(), "adoptedStandby": True})
            if gpus:
                st["gpus"] = gpus
            if inits:
                st["initContainerStatuses"] = inits

        self.store.patch_status("Pod", ns, name, fn)
        return True

    def standbys(self) -> list:
        """(namespace, pod name, GPUs) of the parked ranks (`pdj` debugging, tests)."""
        with self._lock:
            return [(sb.namespace, sb.name, list(sb.gpus)) for sb in self._standbys.values()]

    def _sync_pod(self, key: tuple, pod: dict) -> None:
        ns, name = key
        r = self._ranks.get(key)
        phase = H.pod_phase(pod)
        if r is None:
            if phase in ("Succeeded", "Failed"):
                return
            if key in self._standbys and self._adopt_standby(key, pod):
                return
            need = H.gpus_requested(pod["spec"])
            gpus = self._alloc_gpus(key, need) if need else []
            if gpus is None:
                self._gpu_starved = True  # no standby is (re)started in a pass in which a pod went without a GPU
            if gpus is None and self._standbys:
                # a real pod always beats a parked one: free a standby's GPUs and try again on the next pass
                # (the highest worker index first: a scale-up takes the lowest missing indices)
                self._drop_standby(max(self._standbys, key=lambda k: (H.extract_name_index(k[1])[1], k)))
                self._wake.set()
            if gpus is None:
                if phase != "Pending" or not (pod.get("status") or {}).get("conditions"):
                    self.store.patch_status("Pod", ns, name, lambda st: st.update({
                        "phase": "Pending",
                        "conditions": [{"type": "PodScheduled", "status": "False", "reason": "Unschedulable",
                                        "message": f"0/{len(self.gpu_pool)} GPUs free, need {need}"}]}))
                return
            workdir = self._contained(ns, name)
            if workdir is None:  # never reached through the API (names are validated there); belt and braces
                self.store.patch_status("Pod", ns, name, lambda st: st.update({
                    "phase": "Failed", "reason": "InvalidName",
                    "message": "namespace / name would place the work directory outside the work root"}))
                return
            shutil.rmtree(workdir, ignore_errors=True)
            os.makedirs(workdir, exist_ok=True)
            r = _Rank(ns, name, pod["metadata"].get("uid", ""), workdir,
                      self._alloc_ip(bool(pod["spec"].get("hostNetwork"))), gpus or [])
            r.log_path = os.path.join(workdir, "log.txt")
            r.grace = float(pod["spec"].get("terminationGracePeriodSeconds", 30))
            self._ranks[key] = r
            self._write_hosts()
            inits = [{"name": c["name"], "ready": False, "restartCount": 0,
                      "state": {"running": {"startedAt": H.now_rfc3339()}}}
                     for c in pod["spec"].get("initContainers") or []]
            r.gate_open = not any(c["name"] == H.COORD_CONTAINER_NAME for c in inits)

            def fn(st, ip=r.ip, inits=inits, gpus=list(r.gpus)):
                st.update({"phase": "Pending", "podIP": ip, "hostIP": "127.0.0.1",
                           "conditions": [{"type": "PodScheduled", "status": "True"}]})
                if gpus:
                    st["gpus"] = gpus  # which GPUs of the box the rank was given (`pdj get pods -o wide`)
                if inits:
                    st["initContainerStatuses"] = inits

            self.store.patch_status("Pod", ns, name, fn)
            return
        if r.done:
            return
        if not r.gate_open:
            if os.path.exists(os.path.join(r.workdir, "goon")):
                r.gate_open = True

                def fn(st):
                    for c in st.get("initContainerStatuses") or []:
                        c["ready"] = True
                        c["state"] = {"terminated": {"exitCode": 0, "reason": "Completed",
                                                     "finishedAt": H.now_rfc3339()}}

                self.store.patch_status("Pod", ns, name, fn)
            else:
                return
        if r.proc is None:
            if time.monotonic() < r.next_start:
                return
            self._start(r, pod)
            return
        rc = r.proc.poll()
        if rc is None:
            return
        self._on_exit(r, pod, rc)

    def _build_env(self, r: _Rank, pod: dict) -> Optional[dict]:
        c0 = pod["spec"]["containers"][0]
        env = dict(os.environ)
        env.update(self.extra_env)
        for ef in c0.get("envFrom") or []:
            ref = (ef.get("configMapRef") or {}).get("name")
            if not ref:
                continue
            cm = self.store.try_get("ConfigMap", r.namespace, ref)
            if cm is None:
                if (ef.get("configMapRef") or {}).get("optional"):
                    continue
                return None  # CreateContainerConfigError: wait for the ConfigMap
            env.update({k: str(v) for k, v in (cm.get("data") or {}).items()})
        fields = {"status.podIP": r.ip, "status.hostIP": "127.0.0.1", "metadata.name": r.name,
                  "metadata.namespace": r.namespace, "spec.nodeName": "localhost"}
        for e in c0.get("env") or []:
            if "value" in e:
                env[e["name"]] = str(e["value"])
            elif "valueFrom" in e:
                fp = ((e["valueFrom"] or {}).get("fieldRef") or {}).get("fieldPath")
                if fp in fields:
                    env[e["name"]] = fields[fp]
        owner = H.controller_of(pod) or {}
        env["PDJ_POD_NAME"] = r.name
        env["PDJ_NAMESPACE"] = r.namespace
        env["PDJ_JOB_NAME"] = owner.get("name", "")
        env["PDJ_JOB_KEY"] = f"{r.namespace}-{owner.get('name', r.name)}-{str(owner.get('uid', ''))[:8]}"
        env["PDJ_HOSTS_FILE"] = os.path.join(self.work_root, "hosts")
        env["PDJ_RESTART_COUNT"] = str(r.restarts)
        env["PYTHONUNBUFFERED"] = "1"
        repo = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        env["PYTHONPATH"] = repo + (os.pathsep + env["PYTHONPATH"] if env.get("PYTHONPATH") else "")
        if r.gpus:
            if self.isolate_gpus:
                env["CUDA_VISIBLE_DEVICES"] = ",".join(str(g) for g in r.gpus)
                env["PDJ_LOCAL_DEVICE"] = "0"
            else:
                env["PDJ_LOCAL_DEVICE"] = str(r.gpus[0])
            env["PDJ_GPU_IDS"] = ",".join(str(g) for g in r.gpus)
        # volumes
        vols = {v.get("name"): v for v in pod["spec"].get("volumes") or []}
        for vm in c0.get("volumeMounts") or []:
            v = vols.get(vm.get("name"))
            if v is None:
                continue
            if "hostPath" in v:
                path = v["hostPath"].get("path", "")
            elif "emptyDir" in v:
                mem = (v["emptyDir"] or {}).get("medium") == "Memory"
                base = "/dev/shm/pdjob" if mem and os.path.isdir("/dev/shm") else os.path.join(r.workdir, "volumes")
                path = os.path.join(base, r.namespace, r.name, v["name"]) if mem else os.path.join(base, v["name"])
            elif "persistentVolumeClaim" in v:
                # a claim is one directory shared by every pod (and job) of the namespace that mounts it - what a
                # PV/PVC pair gives the reference's users for data sets and models (docs/user-guide.md:260-347)
                claim = (v["persistentVolumeClaim"] or {}).get("claimName") or v.get("name", "claim")
                path = os.path.join(self.work_root, "pvc", r.namespace, claim)
            else:
                path = os.path.join(r.workdir, "volumes", v.get("name", "vol"))
            try:
                os.makedirs(path, exist_ok=True)
            except OSError:
                pass
            env[mount_env_name(vm.get("mountPath", ""))] = path
        return env

    def _start(self, r: _Rank, pod: dict) -> None:
        c0 = pod["spec"]["containers"][0]
        cmd = resolve_command(c0)
        if cmd is None:
            r.done = True
            self._set_terminal(r, pod, "Failed", 127, "ErrImageNeverPull",
                               f"no command in the template and image {c0.get('image')!r} is not registered")
            return
        env = self._build_env(r, pod)
        if env is None:
            cs = [{"name": c0.get("name", "main"), "ready": False, "restartCount": r.restarts,
                   "state": {"waiting": {"reason": "CreateContainerConfigError"}}}]
            cur = (pod.get("status") or {}).get("containerStatuses")
            if cur != cs:
                self.store.patch_status("Pod", r.namespace, r.name, lambda st: st.update({"containerStatuses": cs}))
            return
        logf = open(r.log_path, "ab")
        logf.write(f"--- start #{r.restarts}: {shlex.join(cmd)}\n".encode())
        logf.flush()
        try:
            r.proc = subprocess.Popen(cmd, cwd=c0.get("workingDir") or r.workdir, env=env, stdout=logf,
                                      stderr=subprocess.STDOUT, start_new_session=True)
        except OSError as e:
            logf.write(f"exec failed: {e}\n".encode())
            logf.close()
            r.done = True
            self._set_terminal(r, pod, "Failed", 127, "StartError", str(e))
            return
        logf.close()
        r.started = True
        cs = [{"name": c0.get("name", "main"), "ready": True, "restartCount": r.restarts,
               "state": {"running": {"startedAt": H.now_rfc3339()}}, "pid": r.proc.pid}]
        self.store.patch_status("Pod", r.namespace, r.name,
                                lambda st: st.update({"phase": "Running", "containerStatuses": cs,
                                                      "startTime": st.get("startTime") or H.now_rfc3339()}))

    def _set_terminal(self, r: _Rank, pod: dict, phase: str, rc: int, reason: str, message: str = "") -> None:
        c0 = pod["spec"]["containers"][0]
        cs = [{"name": c0.get("name", "main"), "ready": False, "restartCount": r.restarts,
               "state": {"terminated": {"exitCode": rc, "reason": reason, "message": message,
                                        "finishedAt": H.now_rfc3339()}}}]
        self._free_gpus((r.namespace, r.name))
        self.store.patch_status("Pod", r.namespace, r.name,
                                lambda st: st.update({"phase": phase, "containerStatuses": cs}))

    def _on_exit(self, r: _Rank, pod: dict, rc: int) -> None:
        policy = pod["spec"].get("restartPolicy") or "Always"
        r.proc = None
        if rc == 0 and policy != "Always":
            r.done = True
            self._set_terminal(r, pod, "Succeeded", 0, "Completed")
            return
        if policy == "Never":
            r.done = True
            self._set_terminal(r, pod, "Failed", rc, "Error")
            return
        # OnFailure / Always: restart in place with exponential back-off
        r.restarts += 1
        delay = min(self.backoff * (2 ** (r.restarts - 1)), self.backoff_max)
        r.next_start = time.monotonic() + delay
        c0 = pod["spec"]["containers"][0]
        cs = [{"name": c0.get("name", "main"), "ready": False, "restartCount": r.restarts,
               "state": {"waiting": {"reason": "CrashLoopBackOff"}},
               "lastState": {"terminated": {"exitCode": rc, "reason": "Error" if rc else "Completed"}}}]
        self.store.patch_status("Pod", r.namespace, r.name, lambda st: st.update({"containerStatuses": cs}))

    def _kill(self, r: _Rank, grace: float = 2.0) -> None:
        p = r.proc
        if p is None or p.poll() is not None:
            return
        if grace > 0:
            try:
                os.killpg(p.pid, signal.SIGTERM)
            except (ProcessLookupError, PermissionError):
                pass
        try:
            p.wait(grace if grace > 0 else 0.001)
        except subprocess.TimeoutExpired:
            try:
                os.killpg(p.pid, signal.SIGKILL)
            except (ProcessLookupError, PermissionError):
                pass
            try:
                p.wait(2)
            except subprocess.TimeoutExpired:
                pass

    def kill_rank(self, namespace: str, name: str, sig: int = signal.SIGKILL) -> bool:
        """Fault injection: signal a rank's process group."""
        with self._lock:
            r = self._ranks.get((namespace, name))
            if r is None or r.proc is None or r.proc.poll() is not None:
                return False
            os.killpg(r.proc.pid, sig)
            self._wake.set()
            return True

    def _contained(self, namespace: str, name: str) -> Optional[str]:
        """<work_root>/<namespace>/<name>, or None when that path would leave the work root ('..', '/', symlinks)."""
        root = os.path.realpath(self.work_root)
        path = os.path.realpath(os.path.join(root, namespace, name))
        if os.path.commonpath([root, path]) != root or path == root:
            return None
        return path

    def logs(self, namespace: str, name: str) -> str:
        with self._lock:
            r = self._ranks.get((namespace, name))
        path = r.log_path if r else None
        if path is None:
            wd = self._contained(namespace, name)
            if wd is None:
                return ""
            path = os.path.join(wd, "log.txt")
        try:
            with open(path, "r", errors="replace") as f:
                return f.read()
        except OSError:
            return ""

    # -- driver -----------------------------------------------------------------------
    def start(self) -> None:
        if self._thread is not None:
            return
        self._stop.clear()
        self._thread = threading.Thread(target=self._loop, name="pdjob-supervisor", daemon=True)
        self._thread.start()

    def _loop(self) -> None:
        while not self._stop.is_set():
            try:
                self.sync()
            except Exception:  # noqa: BLE001
                log.exception("supervisor sync failed")
            self._wake.wait(self.poll)
            self._wake.clear()

    def stop(self, kill: bool = True) -> None:
        self._stop.set()
        self._wake.set()
        if self._thread is not None:
            self._thread.join(5)
            self._thread = None
        if kill:
            with self._lock:
                for r in self._ranks.values():
                    self._kill(r, grace=1.0)
        with self._lock:
            for key in list(self._standbys):
                self._drop_standby(key)
        self._cancel()
