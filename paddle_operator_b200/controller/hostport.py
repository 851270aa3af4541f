"""Per-job host-port blocks for `intranet: Host`.

Behaviour of allocNewPort / allocHostPortForJob in the reference
(controllers/paddlejob_controller.go:407-458): blocks of HOST_PORT_NUM = 20
ports handed out from `--port-range "start,end"` with a moving cursor that wraps
at the end of the range, skipping blocks already in use; when the range is
exhausted the fixed fallback 40000 is returned.  The allocated base port is
recorded on the job as annotation `host-port`, released by the finalizer, and
re-marked from the annotation after a manager restart.  The reference keeps
all of this in one untyped map without a lock; this is a small class with one.

The stand-alone legacy allocator bundled with the reference
(third_party/hostport-allocator/pkg/core/hostportmanager.go:255-324: N ports
per TrainingJob from a k8s PortRange bitmap, annotation
`hostport-manager/hostport: "p1,p2"`, N from `hostport-manager/portnum`) is
covered by `allocate_many` / `format_annotation` / `parse_annotation`.
"""
from __future__ import annotations

import threading

HOST_PORT_NUM = 20
FALLBACK_PORT = 40000

LEGACY_PORTNUM_ANNOTATION = "hostport-manager/portnum"
LEGACY_HOSTPORT_ANNOTATION = "hostport-manager/hostport"


class HostPortAllocator:
    def __init__(self, start: int = 35000, end: int = 65000, block: int = HOST_PORT_NUM):
        if end <= start:
            raise ValueError("port range end must be greater than start")
        self.start, self.end, self.block = int(start), int(end), int(block)
        self.cur = self.start
        self._used: set[int] = set()
        self._lock = threading.Lock()

    @classmethod
    def from_flag(cls, port_range: str) -> "HostPortAllocator":
        """Parses the manager's --port-range "35000,65000" (main.go:101-112)."""
        parts = port_range.split(",")
        if len(parts) != 2:
            raise ValueError("port range must be 'start,end'")
        return cls(int(parts[0]), int(parts[1]))

    @property
    def capacity(self) -> int:
        return (self.end - self.start) // self.block

    def allocate(self) -> int:
        with self._lock:
            if len(self._used) >= self.capacity:
                return FALLBACK_PORT
            for _ in range(self.capacity + 1):
                port = self.cur + self.block
                if port + self.block > self.end + 1 or port > self.end:
                    port = self.start
                self.cur = port
                if port not in self._used:
                    self._used.add(port)
                    return port
            return FALLBACK_PORT

    def is_allocated(self, port) -> bool:
        with self._lock:
            return int(port) in self._used

    def mark(self, port) -> None:
        with self._lock:
            self._used.add(int(port))

    def release(self, port) -> bool:
        with self._lock:
            p = int(port)
            if p in self._used:
                self._used.remove(p)
                return True
            return False

    # -- legacy hostport-manager semantics ----------------------------------
    def allocate_many(self, n: int) -> list[int]:
        """N individual blocks (the legacy manager allocates N single ports
        with AllocateNext); all-or-nothing."""
        got = []
        for _ in range(n):
            p = self.allocate()
            if p == FALLBACK_PORT and (FALLBACK_PORT in got or len(self._used) >= self.capacity):
                for q in got:
                    self.release(q)
                raise RuntimeError("port range exhausted")
            got.append(p)
        return got


def format_annotation(ports: list[int]) -> str:
    return ",".join(str(p) for p in ports)


def parse_annotation(value: str) -> list[int]:
    """third_party/hostport-allocator/pkg/utils/portparse/parse.go:26-56."""
    out = []
    for tok in (value or "").split(","):
        tok = tok.strip()
        if not tok:
            continue
        p = int(tok)
        if not 0 < p < 65536:
            raise ValueError(f"port {p} out of range")
        out.append(p)
    return out


class LegacyHostPortManager:
    """The stand-alone legacy allocator as a running controller
    (third_party/hostport-allocator/pkg/core/hostportmanager.go): it watches one object kind (the reference:
    TrainingJob), and for every object that carries `hostport-manager/portnum: "N"`

      * with no `hostport-manager/hostport` yet: puts it on a work queue; the sync handler allocates N single
        ports and writes them back as the comma list `hostport-manager/hostport` (:255-324);
      * with the list already present (an informer re-list after a restart): marks those ports as taken
        (handleObject :345-385);
      * on delete: releases the ports named by the annotation (:388-418).

    Differences kept deliberate: allocation is all-or-nothing (the reference leaks the ports it had taken when
    AllocateNext fails half-way), the write-back retries on a resourceVersion conflict, and an invalid
    annotation is reported as a Warning event instead of being logged and forgotten."""

    def __init__(self, store, kind: str = "TrainingJob", start: int = 35000, end: int = 65000, recorder=None):
        import queue

        self.store, self.kind, self.recorder = store, kind, recorder
        self.alloc = HostPortAllocator(start, end, block=1)
        self._q: "queue.Queue" = queue.Queue()
        self._stop = threading.Event()
        self._thread: threading.Thread | None = None
        self._cancel = None

    # -- informer side ----------------------------------------------------------------------------
    def _on_event(self, typ: str, obj: dict) -> None:
        if obj.get("kind") != self.kind:
            return
        if typ == "DELETED":
            self._release(obj)
        else:
            self._handle(obj)

    @staticmethod
    def _anno(obj: dict) -> dict:
        return (obj.get("metadata") or {}).get("annotations") or {}

    def _portnum(self, obj: dict) -> int:
        raw = self._anno(obj).get(LEGACY_PORTNUM_ANNOTATION)
        if raw is None:
            return 0
        try:
            n = int(raw)
        except (TypeError, ValueError):
            n = -1
        if n < 0:
            self._warn(obj, "InvalidAnnotation", f"{LEGACY_PORTNUM_ANNOTATION} {raw!r} is not valid")
            return 0
        return n

    def _handle(self, obj: dict) -> None:
        need = self._portnum(obj)
        if need == 0:
            return
        have = self._anno(obj).get(LEGACY_HOSTPORT_ANNOTATION)
        if have is None:
            m = obj["metadata"]
            self._q.put((m.get("namespace", "default"), m["name"]))
            return
        try:
            for p in parse_annotation(have):
                self.alloc.mark(p)
        except ValueError as e:
            self._warn(obj, "InvalidAnnotation", f"{LEGACY_HOSTPORT_ANNOTATION}: {e}")

    def _release(self, obj: dict) -> None:
        try:
            for p in parse_annotation(self._anno(obj).get(LEGACY_HOSTPORT_ANNOTATION, "")):
                self.alloc.release(p)
        except ValueError:
            pass

    def _warn(self, obj: dict, reason: str, msg: str) -> None:
        if self.recorder is not None:
            self.recorder.event(obj, "Warning", reason, msg)

    # -- worker side (syncHandler) -------------------------------------------------------------------
    def sync(self, namespace: str, name: str) -> bool:
        """One pass of the sync handler; True when the object now carries its ports (or needs none)."""
        from .store import Conflict, NotFound

        for _ in range(10):
            try:
                obj = self.store.get(self.kind, namespace, name)
            except NotFound:
                return True  # "no longer exists": stop processing
            need = self._portnum(obj)
            if need == 0 or self._anno(obj).get(LEGACY_HOSTPORT_ANNOTATION) is not None:
                return True
            try:
                ports = self.alloc.allocate_many(need)
            except RuntimeError as e:
                self._warn(obj, "PortsExhausted", str(e))
                return False
            obj["metadata"].setdefault("annotations", {})[LEGACY_HOSTPORT_ANNOTATION] = format_annotation(ports)
            try:
                self.store.update(obj)
                return True
            except Conflict:
                for p in ports:
                    self.alloc.release(p)
            except NotFound:
                for p in ports:
                    self.alloc.release(p)
                return True
        return False

    def _run(self) -> None:
        import queue

        while not self._stop.is_set():
            try:
                ns, name = self._q.get(timeout=0.1)
            except queue.Empty:
                continue
            self.sync(ns, name)

    def start(self) -> None:
        """Lists what exists (ports already annotated are re-marked), then follows the watch."""
        self._cancel = self.store.watch(self._on_event)
        for obj in self.store.list(self.kind):
            self._handle(obj)
        self._thread = threading.Thread(target=self._run, name="hostport-manager", daemon=True)
        self._thread.start()

    def stop(self) -> None:
        self._stop.set()
        if self._cancel:
            self._cancel()
        if self._thread:
            self._thread.join(2)

    def drain(self, timeout: float = 5.0) -> None:
        """Blocks until the work queue is empty (tests)."""
        import time

        end = time.time() + timeout
        while time.time() < end and (not self._q.empty()):
            time.sleep(0.01)
        time.sleep(0.05)
