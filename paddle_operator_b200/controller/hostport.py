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
