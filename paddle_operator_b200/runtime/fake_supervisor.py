"""A scriptable stand-in for kubelet/scheduler, for tests and dry runs.

The reference tests its controller under envtest: a real apiserver but *no*
kubelet, so pods are API objects that never run (controllers/suite_test.go:
51-88).  With no supervisor attached the object store behaves exactly like
that.  FakeSupervisor adds the missing half in a controllable way: tests call
its methods to give pods an address, start the coord gate, let containers run,
and finish or fail them - which drives the job through Pending -> Starting ->
Running -> Completed/Failed.
"""
from __future__ import annotations

from typing import Optional

from ..controller import helpers as H
from ..controller.store import NotFound, ObjectStore


class FakeSupervisor:
    def __init__(self, store: ObjectStore, ip_prefix: str = "10.1.0."):
        self.store = store
        self.ip_prefix = ip_prefix
        self._next_ip = 1
        self.exec_log: list[tuple] = []  # (namespace, pod, container, tuple(cmd))

    # -- scheduling: Pending + address + init container running ---------------
    def schedule(self, namespace: str, name: str) -> None:
        pod = self.store.get("Pod", namespace, name)
        if H.pod_phase(pod):
            return
        ip = f"{self.ip_prefix}{self._next_ip}"
        self._next_ip += 1
        inits = [{"name": c["name"], "ready": False, "state": {"running": {"startedAt": H.now_rfc3339()}}}
                 for c in pod["spec"].get("initContainers") or []]

        def fn(st):
            st.update({"phase": "Pending", "podIP": ip, "hostIP": "127.0.0.1"})
            if inits:
                st["initContainerStatuses"] = inits

        self.store.patch_status("Pod", namespace, name, fn)

    def schedule_all(self, namespace: str = "default") -> int:
        n = 0
        for pod in self.store.list("Pod", namespace):
            if not H.pod_phase(pod):
                self.schedule(namespace, pod["metadata"]["name"])
                n += 1
        return n

    # -- the pods/exec endpoint (controllers/paddlejob_controller.go:491-518) --
    def exec_in_pod(self, namespace: str, pod_name: str, container: str, cmd: list) -> int:
        self.exec_log.append((namespace, pod_name, container, tuple(cmd)))
        if container == H.COORD_CONTAINER_NAME and list(cmd) == H.COORD_RELEASE_CMD:
            def fn(st):
                for c in st.get("initContainerStatuses") or []:
                    if c.get("name") == H.COORD_CONTAINER_NAME:
                        c["ready"] = True
                        c["state"] = {"terminated": {"exitCode": 0, "reason": "Completed"}}

            try:
                self.store.patch_status("Pod", namespace, pod_name, fn)
            except NotFound:
                return 1
        return 0

    # -- container lifecycle ----------------------------------------------------------
    def _gate_open(self, pod: dict) -> bool:
        for c in (pod.get("status") or {}).get("initContainerStatuses") or []:
            if (c.get("state") or {}).get("terminated") is None:
                return False
        return True

    def _config_ready(self, pod: dict) -> bool:
        ns = pod["metadata"]["namespace"]
        for c in pod["spec"].get("containers") or []:
            for ef in c.get("envFrom") or []:
                ref = (ef.get("configMapRef") or {}).get("name")
                if ref and self.store.try_get("ConfigMap", ns, ref) is None:
                    return False  # kubelet: CreateContainerConfigError (faq.md:18-26 of the reference)
        return True

    def start_ready(self, namespace: str = "default") -> int:
        """Moves every Pending pod whose gate is open and whose ConfigMap exists to Running."""
        n = 0
        for pod in self.store.list("Pod", namespace):
            if H.pod_phase(pod) != "Pending" or not self._gate_open(pod) or not self._config_ready(pod):
                continue
            cs = [{"name": c.get("name", "c"), "ready": True, "restartCount": 0,
                   "state": {"running": {"startedAt": H.now_rfc3339()}}} for c in pod["spec"].get("containers") or []]
            self.store.patch_status("Pod", namespace, pod["metadata"]["name"],
                                    lambda st, cs=cs: st.update({"phase": "Running", "containerStatuses": cs,
                                                                 "startTime": H.now_rfc3339()}))
            n += 1
        return n

    def finish(self, namespace: str, pod_name: str, succeeded: bool = True, exit_code: Optional[int] = None) -> None:
        code = exit_code if exit_code is not None else (0 if succeeded else 1)

        def fn(st):
            st["phase"] = "Succeeded" if succeeded else "Failed"
            for c in st.get("containerStatuses") or []:
                c["ready"] = False
                c["state"] = {"terminated": {"exitCode": code, "reason": "Completed" if succeeded else "Error"}}

        self.store.patch_status("Pod", namespace, pod_name, fn)

    def finish_all(self, namespace: str = "default", succeeded: bool = True) -> None:
        for pod in self.store.list("Pod", namespace):
            if H.pod_phase(pod) == "Running":
                self.finish(namespace, pod["metadata"]["name"], succeeded)

    def set_unready(self, namespace: str, pod_name: str) -> None:
        def fn(st):
            for c in st.get("containerStatuses") or []:
                c["ready"] = False

        self.store.patch_status("Pod", namespace, pod_name, fn)
