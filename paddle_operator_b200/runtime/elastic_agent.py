"""Launcher-side elastic membership through the etcd-compatible store.

The reference's controller contributes only the environment
(PADDLE_ELASTIC_JOB_ID / _NP / _TIMEOUT / _SERVER, controllers/
paddlejob_helper.go:333-347, controllers/paddlejob_controller.go:259-265) and the
`np` push (controllers/paddlejob_elastic.go:41-55); everything else happens in
the rank processes (Paddle's elastic manager there).  This is that other half,
following the same key layout:

    /paddle/<job_id>/np                 desired world size (created by the ranks,
                                        updated by the controller)
    /paddle/<job_id>/nodes/<host>       one key per live rank, attached to a lease
                                        that the rank keeps alive
    /paddle/<job_id>/stop/<gen>         step at which generation <gen> stops so that
                                        every rank leaves the collective together
    /paddle/<job_id>/gen                current generation (bumped on every re-form)
    /paddle/<job_id>/resume/<gen>       JSON {step, world} published by rank 0 of the
                                        new generation before training resumes

A *generation* is one fixed membership.  Ranks train in lock-step inside a
generation; when np or the node set changes they agree on a stop step, leave
the collective at that step (state stays in HBM), and re-form.
"""
from __future__ import annotations

import json
import os
import threading
import time
import uuid
from dataclasses import dataclass, field
from typing import Optional


@dataclass
class Membership:
    generation: int
    rank: int
    world: int
    hosts: list  # sorted node keys of the members
    root: Optional[str] = None   # member that holds valid state (None: start from the checkpoint / scratch)
    step: Optional[int] = None   # step at which this generation starts (None: taken from the checkpoint)
    fresh: bool = False          # no live predecessor generation: state comes from disk, not from a peer
    incs: dict = field(default_factory=dict)  # host -> process incarnation the generation was formed with


class ElasticAgent:
    def __init__(self, etcd, job_id: str, host: str, np: int, timeout: float = 60.0, ttl: int = 10):
        self.etcd, self.job_id, self.host = etcd, job_id, host
        self.prefix = f"/paddle/{job_id}"
        self.initial_np, self.timeout, self.ttl = int(np), float(timeout), int(ttl)
        self.lease: Optional[int] = None
        self.leaving = False
        self._dirty = threading.Event()
        self._stop = threading.Event()
        self._ka: Optional[threading.Thread] = None
        self._cancel_watch = None
        self._lock = threading.Lock()
        self._np: Optional[int] = None
        self._nodes: dict = {}  # host -> incarnation
        self._stops: dict = {}
        self.inc = uuid.uuid4().hex[:12]  # distinguishes a restarted process under the same host key

    @classmethod
    def from_env(cls, etcd=None, ttl: int = 10) -> "ElasticAgent":
        env = os.environ
        if etcd is None:
            from .etcd_client import EtcdClient

            etcd = EtcdClient(env["PADDLE_ELASTIC_SERVER"])
        tid = int(env.get("PADDLE_TRAINER_ID", "0"))
        # the sort key puts ranks in trainer-id order, like the endpoint lists of the reference
        host = f"{tid:05d}-{env.get('POD_IP', env.get('PDJ_POD_NAME', 'localhost'))}"
        return cls(etcd, env["PADDLE_ELASTIC_JOB_ID"], host, int(env.get("PADDLE_ELASTIC_NP", "1")),
                   float(env.get("PADDLE_ELASTIC_TIMEOUT", "60")), ttl)

    # -- registration ---------------------------------------------------------------
    def register(self) -> None:
        self.lease = self.etcd.lease_grant(self.ttl)
        self.etcd.put(f"{self.prefix}/nodes/{self.host}", json.dumps({"host": self.host, "pid": os.getpid(),
                                                                       "inc": self.inc, "ts": time.time()}),
                      lease=self.lease)
        self.etcd.put_if_absent(f"{self.prefix}/np", str(self.initial_np))
        self._refresh()
        self._cancel_watch = self.etcd.watch_prefix(self.prefix + "/", self._on_events)
        self._refresh()
        self._ka = threading.Thread(target=self._keepalive, name="elastic-keepalive", daemon=True)
        self._ka.start()

    def _keepalive(self) -> None:
        while not self._stop.wait(max(0.2, self.ttl / 3.0)):
            try:
                if self.lease is not None and self.etcd.lease_keepalive(self.lease) == 0 and not self.leaving:
                    # the lease was lost (store restart, long stall): register again
                    self.lease = self.etcd.lease_grant(self.ttl)
                    self.etcd.put(f"{self.prefix}/nodes/{self.host}", json.dumps({"host": self.host, "inc": self.inc}),
                                  lease=self.lease)
            except Exception:  # noqa: BLE001 - keep trying; expiry is the failure signal
                pass

    def _refresh(self) -> None:
        kvs = self.etcd.get_prefix(self.prefix + "/")
        with self._lock:
            self._nodes = {k.rsplit("/", 1)[1]: self._inc_of(v) for k, v in kvs
                           if k.startswith(self.prefix + "/nodes/")}
            for k, v in kvs:
                if k == self.prefix + "/np":
                    self._np = int(v)
                elif k.startswith(self.prefix + "/stop/"):
                    self._stops[int(k.rsplit("/", 1)[1])] = int(v)
        self._dirty.set()

    def _on_events(self, events: list) -> None:
        with self._lock:
            for typ, k, v in events:
                if k == self.prefix + "/np" and typ == "PUT":
                    self._np = int(v)
                elif k.startswith(self.prefix + "/nodes/"):
                    h = k.rsplit("/", 1)[1]
                    if typ == "PUT":
                        self._nodes[h] = self._inc_of(v)
                    else:
                        self._nodes.pop(h, None)
                elif k.startswith(self.prefix + "/stop/") and typ == "PUT":
                    self._stops[int(k.rsplit("/", 1)[1])] = int(v)
        self._dirty.set()

    @staticmethod
    def _inc_of(value: str) -> str:
        try:
            return json.loads(value).get("inc", "")
        except ValueError:
            return ""

    # -- views ---------------------------------------------------------------------------
    def np(self) -> int:
        with self._lock:
            return self._np if self._np is not None else self.initial_np

    def nodes(self) -> list:
        with self._lock:
            return sorted(self._nodes)

    def node_incs(self) -> dict:
        with self._lock:
            return dict(self._nodes)

    def generation(self) -> int:
        kv = self.etcd.get(f"{self.prefix}/gen")
        return int(kv[0][1]) if kv else 0

    # -- generations: who is in, who holds the state ----------------------------------------
    def _members_key(self, gen: int) -> str:
        return f"{self.prefix}/members/{gen:08d}"

    def latest_members(self) -> Optional[tuple]:
        kvs = self.etcd.get_prefix(f"{self.prefix}/members/")
        if not kvs:
            return None
        k, v = max(kvs, key=lambda kv: kv[0])
        return int(k.rsplit("/", 1)[1]), json.loads(v)

    def _membership_from(self, gen: int, rec: dict) -> Optional[Membership]:
        if self.host not in rec["hosts"] or rec["incs"].get(self.host) != self.inc:
            return None
        return Membership(gen, rec["hosts"].index(self.host), len(rec["hosts"]), list(rec["hosts"]), rec.get("root"),
                          rec.get("step"), bool(rec.get("fresh")), dict(rec.get("incs") or {}))

    def _generation_alive(self, rec: dict) -> bool:
        incs = self.node_incs()
        return any(incs.get(h) == i for h, i in rec["incs"].items() if h != self.host or i != self.inc)

    def join(self, timeout: Optional[float] = None) -> Membership:
        """Registers and returns the first generation this process is a member of.

        * no generation yet, or the last one has no live member (whole-job restart): the first np
          registered nodes form a fresh generation whose state comes from the checkpoint;
        * a live generation exists: wait until its members admit this node (scale-up, or a
          restarted rank) and pass their state on."""
        if self.lease is None:
            self.register()
        deadline = time.monotonic() + (timeout if timeout is not None else self.timeout)
        seen_gen = -1
        while time.monotonic() < deadline:
            self._refresh()
            cur = self.latest_members()
            if cur is not None:
                gen, rec = cur
                m = self._membership_from(gen, rec)
                if m is not None and gen > seen_gen:
                    return m
                if self._generation_alive(rec):
                    seen_gen = max(seen_gen, gen)
                    self._dirty.clear()
                    self._dirty.wait(0.05)
                    continue
                target = gen + 1
            else:
                target = 0
            np, nodes = self.np(), self.nodes()
            if len(nodes) >= np > 0 and self.host in nodes[:np]:
                members = nodes[:np]
                if members[0] == self.host:
                    incs = self.node_incs()
                    self.etcd.put_if_absent(self._members_key(target), json.dumps(
                        {"hosts": members, "incs": {h: incs.get(h, "") for h in members}, "root": None, "step": None,
                         "fresh": True}))
            self._dirty.clear()
            self._dirty.wait(0.05)
        raise TimeoutError(f"elastic: could not join a generation within {self.timeout:.0f}s "
                           f"(np={self.np()}, nodes={self.nodes()})")

    def next_generation(self, m: Membership, stop_step: int, timeout: Optional[float] = None) -> Optional[Membership]:
        """Called by every member of generation m after it has stopped at `stop_step`.  Returns the
        next membership, or None when this process is not part of it (it then leaves)."""
        if self.leaving:
            self.leave()
            return None
        deadline = time.monotonic() + (timeout if timeout is not None else self.timeout)
        key = self._members_key(m.generation + 1)
        while time.monotonic() < deadline:
            self._refresh()
            kv = self.etcd.get(key)
            if kv:
                nm = self._membership_from(m.generation + 1, json.loads(kv[0][1]))
                if nm is None:
                    self.leave()
                return nm
            np, nodes, incs = self.np(), self.nodes(), self.node_incs()
            if len(nodes) >= np > 0:
                members = nodes[:np]
                stay = [h for h in m.hosts if h in members and (not m.incs.get(h) or incs.get(h) == m.incs[h])]
                if not stay:
                    self.leave()
                    return None
                if stay[0] == self.host:
                    self.etcd.put_if_absent(key, json.dumps(
                        {"hosts": members, "incs": {h: incs.get(h, "") for h in members}, "root": self.host,
                         "step": int(stop_step), "fresh": False}))
                    continue
            self._dirty.clear()
            self._dirty.wait(0.05)
        raise TimeoutError("elastic: the next generation did not form in time")

    def should_stop(self, m: Membership, step: int, margin: int = 4) -> Optional[int]:
        """The step at which generation m must stop, once a re-form is due; else None."""
        s = self.stop_step(m.generation)
        if s is None and self.needs_reform(m):
            s = self.agree_stop_step(m.generation, step, margin)
        return s

    # -- forming a generation ----------------------------------------------------------------
    def wait_ready(self, generation: int, timeout: Optional[float] = None) -> Membership:
        """Blocks until exactly np nodes are registered (the first np in sort order once more are
        present) and this host is one of them.  Raises TimeoutError after PADDLE_ELASTIC_TIMEOUT."""
        deadline = time.monotonic() + (timeout if timeout is not None else self.timeout)
        while True:
            self._refresh()
            np, nodes = self.np(), self.nodes()
            if len(nodes) >= np and np > 0:
                members = nodes[:np]
                if self.host in members:
                    return Membership(generation, members.index(self.host), np, members)
            if time.monotonic() > deadline:
                raise TimeoutError(f"elastic: {len(nodes)} of {np} nodes registered after {self.timeout:.0f}s: {nodes}")
            self._dirty.clear()
            self._dirty.wait(0.2)

    def needs_reform(self, m: Membership) -> bool:
        """True when the desired world differs from the current one and can be formed now
        (scale-down: immediately; scale-up: once the newcomers have registered), or when a
        member vanished (lease expiry)."""
        if self.leaving:
            return True
        np, nodes = self.np(), set(self.nodes())
        if any(h not in nodes for h in m.hosts):
            return True  # a member's lease expired
        live = self.node_incs()
        if any(m.incs.get(h) and live.get(h) != m.incs[h] for h in m.hosts):
            return True  # a member died and was restarted under the same name: it holds no state
        if np < m.world:
            return True
        if np > m.world and len(nodes) >= np:
            return True
        return False

    def agree_stop_step(self, generation: int, my_step: int, margin: int = 4) -> int:
        """First proposer wins (create-if-absent); everybody reads the same value."""
        key = f"{self.prefix}/stop/{generation}"
        self.etcd.put_if_absent(key, str(my_step + margin))
        kv = self.etcd.get(key)
        s = int(kv[0][1])
        with self._lock:
            self._stops[generation] = s
        return s

    def stop_step(self, generation: int) -> Optional[int]:
        with self._lock:
            return self._stops.get(generation)

    def publish_resume(self, generation: int, step: int, world: int) -> None:
        self.etcd.put(f"{self.prefix}/resume/{generation}", json.dumps({"step": step, "world": world}))
        self.etcd.put(f"{self.prefix}/gen", str(generation))

    def wait_resume(self, generation: int, timeout: Optional[float] = None) -> dict:
        deadline = time.monotonic() + (timeout if timeout is not None else self.timeout)
        key = f"{self.prefix}/resume/{generation}"
        while time.monotonic() < deadline:
            kv = self.etcd.get(key)
            if kv:
                return json.loads(kv[0][1])
            time.sleep(0.02)
        raise TimeoutError(f"elastic: no resume record for generation {generation}")

    def report(self, key: str, value) -> None:
        """Progress / timing breadcrumbs for harnesses: /paddle/<job>/report/<host>/<key>."""
        try:
            self.etcd.put(f"{self.prefix}/report/{self.host}/{key}", json.dumps(value))
        except Exception:  # noqa: BLE001 - breadcrumbs are best effort; never take a rank down for one
            pass

    def leave(self) -> None:
        self.leaving = True
        self._stop.set()
        if self._cancel_watch:
            self._cancel_watch()
        if self.lease is not None:
            self.etcd.lease_revoke(self.lease)
            self.lease = None
