"""Symmetric-memory communicator over NVSwitch peer mappings.

One process per GPU.  Every rank owns an *arena* (cuMemCreate allocation) that
all ranks of the job map; buffers are named by byte offset, so an offset is
valid on every rank.  Kernels in csrc/collectives.cu read and write peers
directly through those mappings.

Reference parity: this replaces the ConfigMap rendezvous of the reference
(controllers/paddlejob_controller.go:289-306, controllers/paddlejob_helper.go:
215-279).  There the rendezvous object carries "<ip>:2379" endpoint strings;
here it additionally carries one POSIX file descriptor per rank (SCM_RIGHTS
over unix sockets under a per-job directory).
"""
from __future__ import annotations

import ctypes
import os
import socket
import struct
import tempfile
import threading
import time
from dataclasses import dataclass
from typing import Optional, Sequence

from ..ops import _lib
from ..ops._lib import AR_NVLS, AR_ONESHOT, AR_TWOSHOT, PAD_BYTES, PdjComm, check

__all__ = ["Communicator", "FdExchange", "AR_TWOSHOT", "AR_ONESHOT", "AR_NVLS"]


class _RawCuda:
    """Exposes a raw device pointer through __cuda_array_interface__."""

    def __init__(self, ptr: int, nbytes: int, owner=None):
        self.__cuda_array_interface__ = {
            "shape": (nbytes,),
            "typestr": "|u1",
            "data": (ptr, False),
            "version": 2,
            "strides": None,
        }
        self._owner = owner


def _as_tensor(ptr: int, nbytes: int, device: int, owner=None):
    import torch

    return torch.as_tensor(_RawCuda(ptr, nbytes, owner), device=f"cuda:{device}")


# ---------------------------------------------------------------------------
class FdExchange:
    """All-to-all exchange of file descriptors (and small blobs) between the
    ranks of one job on one host, over unix sockets in a shared directory."""

    def __init__(self, key: str, rank: int, nranks: int, timeout: float = 120.0):
        self.rank, self.nranks, self.timeout = rank, nranks, timeout
        # The directory carries the GPU-memory file descriptors of the job: it must belong to this user and be
        # closed to everybody else.  Under the supervisor it sits next to the hosts file (PDJ_HOSTS_FILE, inside the
        # manager's work root); otherwise under the temp dir, where a pre-created directory or a symlink planted by
        # another local user is refused instead of followed.
        hosts = os.environ.get("PDJ_HOSTS_FILE")
        root = os.path.join(os.path.dirname(hosts), "fdx") if hosts else tempfile.gettempdir()
        if hosts:
            os.makedirs(root, mode=0o700, exist_ok=True)
        safe = "".join(ch if (ch.isalnum() or ch in "-_.") else "_" for ch in key)
        self.dir = os.path.join(root, f"pdj-{safe}")
        try:
            os.mkdir(self.dir, 0o700)
        except FileExistsError:
            pass
        st = os.lstat(self.dir)
        import stat as _stat

        if _stat.S_ISLNK(st.st_mode) or not _stat.S_ISDIR(st.st_mode) or st.st_uid != os.geteuid():
            raise PermissionError(f"{self.dir} is not a directory owned by uid {os.geteuid()}; refusing to exchange fds there")
        if st.st_mode & 0o077:
            os.chmod(self.dir, 0o700)
        self._round = 0

    def _path(self, rnd: int, rank: int) -> str:
        return os.path.join(self.dir, f"r{rnd}-{rank}.sock")

    def exchange(self, fd: Optional[int], blob: bytes = b"") -> list[tuple[Optional[int], bytes]]:
        """Every rank contributes (fd, blob); returns the list indexed by rank
        (own entry is (fd, blob) unchanged)."""
        rnd = self._round
        self._round += 1
        n, me = self.nranks, self.rank
        out: list = [None] * n
        out[me] = (fd, blob)
        if n == 1:
            return out
        path = self._path(rnd, me)
        if os.path.exists(path):
            os.unlink(path)
        srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
        srv.bind(path)
        srv.listen(n)
        srv.settimeout(self.timeout)
        err: list = []

        def serve():
            try:
                for _ in range(n - 1):
                    conn, _ = srv.accept()
                    with conn:
                        msg, fds, _, _ = socket.recv_fds(conn, 1 << 16, 1)
                        src, has_fd, blen = struct.unpack("<iii", msg[:12])
                        body = msg[12:]
                        while len(body) < blen:
                            more = conn.recv(blen - len(body))
                            if not more:
                                raise ConnectionError(f"rank {src} closed the fd exchange after {len(body)} of {blen} bytes")
                            body += more
                        out[src] = (fds[0] if has_fd and fds else None, bytes(body))
                        conn.sendall(b"k")
            except Exception as e:  # pragma: no cover - surfaced below
                err.append(e)

        th = threading.Thread(target=serve, daemon=True)
        th.start()
        for peer in range(n):
            if peer == me:
                continue
            deadline = time.time() + self.timeout
            while True:
                c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
                try:
                    c.connect(self._path(rnd, peer))
                    break
                except (FileNotFoundError, ConnectionRefusedError):
                    c.close()
                    if time.time() > deadline:
                        raise TimeoutError(f"rank {me}: peer {peer} never opened {self._path(rnd, peer)}")
                    time.sleep(0.01)
            with c:
                hdr = struct.pack("<iii", me, 1 if fd is not None else 0, len(blob))
                socket.send_fds(c, [hdr + blob], [fd] if fd is not None else [])
                c.settimeout(self.timeout)
                c.recv(1)
        th.join(self.timeout)
        srv.close()
        try:
            os.unlink(path)
        except OSError:
            pass
        if err:
            raise err[0]
        if any(o is None for o in out):
            raise TimeoutError(f"rank {me}: fd exchange round {rnd} incomplete")
        return out

    def barrier(self) -> None:
        self.exchange(None, b"")


# ---------------------------------------------------------------------------
@dataclass
class _Alloc:
    offset: int
    nbytes: int


class Communicator:
    """A group of ranks with mapped arenas.

    ``Communicator.emulated(n)`` puts n arenas on one device and runs every
    collective as one cooperative launch over all ranks (used by the GPU tests,
    which only get one GPU).  ``Communicator.from_env()`` is the real thing.
    """

    def __init__(self):
        self.lib = _lib.load()
        self.nranks = 1
        self.rank = 0
        self.device = 0
        self.emulate = False
        self.arena_bytes = 0
        self._arenas: list[int] = []  # PdjArena* per rank (as mapped here)
        self._mc: Optional[int] = None
        self._comm = PdjComm()
        self._cursor = PAD_BYTES
        self._bases: list[int] = []
        self._channels: dict[int, PdjComm] = {}
        self.has_multicast = False

    # -- construction -------------------------------------------------------
    @classmethod
    def emulated(cls, nranks: int, arena_bytes: int = 64 << 20, device: int = 0) -> "Communicator":
        self = cls()
        self.nranks, self.rank, self.device, self.emulate = nranks, 0, device, True
        rounded = ctypes.c_size_t()
        check(self.lib.pdj_arena_round_size(device, arena_bytes, 1, ctypes.byref(rounded)), "arena_round_size")
        self.arena_bytes = rounded.value
        for _ in range(nranks):
            h = ctypes.c_void_p()
            check(self.lib.pdj_arena_create(device, self.arena_bytes, ctypes.byref(h)), "arena_create")
            self._arenas.append(h.value)
        self._finish()
        return self

    @classmethod
    def create(cls, rank: int, nranks: int, device: int, arena_bytes: int, key: str,
               want_multicast: bool = True) -> "Communicator":
        self = cls()
        self.nranks, self.rank, self.device = nranks, rank, device
        lib = self.lib
        rounded = ctypes.c_size_t()
        check(lib.pdj_arena_round_size(device, arena_bytes, nranks, ctypes.byref(rounded)), "arena_round_size")
        self.arena_bytes = rounded.value
        mine = ctypes.c_void_p()
        check(lib.pdj_arena_create(device, self.arena_bytes, ctypes.byref(mine)), "arena_create")
        self._arenas = [None] * nranks
        self._arenas[rank] = mine.value
        if nranks > 1:
            ex = FdExchange(key, rank, nranks)
            fd = ctypes.c_int(-1)
            check(lib.pdj_arena_export_fd(mine, ctypes.byref(fd)), "arena_export_fd")
            got = ex.exchange(fd.value, struct.pack("<q", self.arena_bytes))
            for r, (pfd, blob) in enumerate(got):
                if r == rank:
                    continue
                (pbytes,) = struct.unpack("<q", blob)
                if pbytes != self.arena_bytes:
                    raise RuntimeError(f"rank {r} arena is {pbytes} bytes, mine is {self.arena_bytes}")
                h = ctypes.c_void_p()
                check(lib.pdj_arena_import_fd(device, pfd, self.arena_bytes, ctypes.byref(h)), "arena_import_fd")
                os.close(pfd)
                self._arenas[r] = h.value
            os.close(fd.value)
            if want_multicast and os.environ.get("PDJ_DISABLE_NVLS", "0") != "1":
                self._setup_multicast(ex)
            ex.barrier()
        self._finish()
        return self

    @classmethod
    def from_env(cls, arena_bytes: int = 512 << 20, key: Optional[str] = None, **kw) -> "Communicator":
        """Builds the communicator from torchrun-style (RANK/WORLD_SIZE/
        LOCAL_RANK) or PaddleJob-style (PADDLE_TRAINER_ID/PADDLE_TRAINERS_NUM)
        environment, the latter being what the reconciler injects
        (controllers/paddlejob_helper.go:319-322 in the reference)."""
        env = os.environ
        rank = int(env.get("RANK", env.get("PADDLE_TRAINER_ID", "0")))
        n = int(env.get("WORLD_SIZE", env.get("PADDLE_TRAINERS_NUM", "1")))
        dev = int(env.get("LOCAL_RANK", env.get("PDJ_LOCAL_DEVICE", str(rank))))
        if key is None:
            key = env.get("PDJ_JOB_KEY") or f"{env.get('MASTER_ADDR', 'local')}-{env.get('MASTER_PORT', '0')}"
        return cls.create(rank, n, dev, arena_bytes, key, **kw)

    def _setup_multicast(self, ex: FdExchange) -> None:
        lib = self.lib
        ok = bool(lib.pdj_multicast_supported(self.device))
        flags = ex.exchange(None, b"1" if ok else b"0")
        if not all(b == b"1" for _, b in flags):
            return
        mc = ctypes.c_void_p()
        fd = ctypes.c_int(-1)
        rc = 0
        if self.rank == 0:
            rc = lib.pdj_multicast_create(self.device, self.nranks, self.arena_bytes, ctypes.byref(mc), ctypes.byref(fd))
        got = ex.exchange(fd.value if (self.rank == 0 and rc == 0) else None, b"1" if rc == 0 else b"0")
        root_fd, root_ok = got[0]
        if root_ok != b"1" or (self.rank != 0 and root_fd is None):
            return
        if self.rank != 0:
            rc = lib.pdj_multicast_import(self.device, root_fd, self.arena_bytes, ctypes.byref(mc))
            os.close(root_fd)
        else:
            os.close(fd.value)
        if rc == 0:
            rc = lib.pdj_multicast_add_device(mc)
        oks = ex.exchange(None, b"1" if rc == 0 else b"0")
        if not all(b == b"1" for _, b in oks):
            return
        ptr = ctypes.c_void_p()
        rc = lib.pdj_multicast_bind(mc, ctypes.c_void_p(self._arenas[self.rank]), ctypes.byref(ptr))
        oks = ex.exchange(None, b"1" if rc == 0 else b"0")
        if not all(b == b"1" for _, b in oks):
            return
        self._mc = mc.value
        self._comm.mc_base = ptr.value
        self.has_multicast = True

    def _finish(self) -> None:
        lib = self.lib
        self._comm.nranks = self.nranks
        self._comm.rank = self.rank
        # how long a device-side barrier waits for a peer before trapping (0 = the 8 s default)
        self._comm.spin_timeout_ms = int(os.environ.get("PDJ_SPIN_TIMEOUT_MS", "0"))
        self._bases = [lib.pdj_arena_ptr(ctypes.c_void_p(a)) for a in self._arenas]
        for i, b in enumerate(self._bases):
            self._comm.base[i] = b
        self._cursor = PAD_BYTES
        self._channels.clear()

    # -- memory ---------------------------------------------------------------
    def alloc(self, nbytes: int, align: int = 256) -> int:
        """Bump-allocates `nbytes` in the arena; returns the byte offset.  All
        ranks must call alloc in the same order with the same sizes."""
        off = (self._cursor + align - 1) // align * align
        if off + nbytes > self.arena_bytes:
            raise MemoryError(f"symmetric arena exhausted: need {off + nbytes}, have {self.arena_bytes}")
        self._cursor = off + nbytes
        return off

    def reset_allocations(self) -> None:
        self._cursor = PAD_BYTES

    def tensor(self, offset: int, shape: Sequence[int], dtype, rank: Optional[int] = None):
        """A torch view of [offset, offset+size) in `rank`'s arena (default: my
        own).  Peer views are real peer-mapped memory."""
        import torch

        r = self.rank if rank is None else rank
        n = 1
        for s in shape:
            n *= int(s)
        nbytes = n * torch.empty((), dtype=dtype).element_size()
        if offset + nbytes > self.arena_bytes:
            raise ValueError("tensor exceeds arena")
        t = _as_tensor(self._bases[r] + offset, nbytes, self.device, owner=self)
        return t.view(dtype).view(*shape)

    # -- collectives ------------------------------------------------------------
    def _stream(self, stream) -> int:
        import torch

        return (stream or torch.cuda.current_stream()).cuda_stream

    def barrier(self, stream=None) -> None:
        check(self.lib.pdj_barrier(ctypes.byref(self._comm), int(self.emulate), self._stream(stream)), "barrier")

    def pick_algo(self, nbytes: int, in_place: bool = False) -> int:
        """The all-reduce algorithm the product runs for a message of `nbytes`, from the sweeps in profiles/
        (r1/sweep_n2.json, r1/sweep_n8.json, r2/): with a multicast object and >= 4 ranks the switch reduction wins
        or ties at every size; peer-to-peer, the one-shot form only pays up to 64 KiB (and cannot run in place)."""
        if self.nranks == 1:
            return AR_TWOSHOT
        if self.has_multicast and not self.emulate and self.nranks >= 4:
            return AR_NVLS
        if nbytes <= (64 << 10) and not in_place:
            return AR_ONESHOT
        return AR_TWOSHOT

    def pick_blocks(self, nbytes: int, algo: Optional[int] = None) -> int:
        """Thread blocks for a collective of `nbytes` (same sweeps): launch-latency-bound messages want few blocks
        (every block takes part in the device barrier), bandwidth-bound ones want the whole chip."""
        if algo == AR_NVLS:
            # 8 ranks: 32 and 148 blocks tie at every size; 4 ranks: 148 wins from 32 MiB up (r2/sweep_n4.json:
            # 613 vs 699 us at 256 MiB)
            return 148 if nbytes >= (32 << 20) else 32
        if nbytes <= (256 << 10):
            return 16
        if nbytes <= (1 << 20):
            return 64
        return 148

    def all_reduce_bf16(self, in_off: int, out_off: int, n_elems: int, scale: float = 1.0, algo: Optional[int] = None,
                        out_f32: bool = False, nblocks: int = 32, stream=None) -> None:
        if algo is None:
            algo = self.pick_algo(n_elems * 2, in_place=(in_off == out_off))
            if algo == AR_ONESHOT and out_f32:
                algo = AR_TWOSHOT
            if algo == AR_NVLS and out_f32:
                algo = AR_TWOSHOT
        if nblocks <= 0:
            nblocks = self.pick_blocks(n_elems * 2, algo)
        check(self.lib.pdj_allreduce_bf16(ctypes.byref(self._comm), in_off, out_off, n_elems, scale, algo,
                                          int(out_f32), nblocks, int(self.emulate), self._stream(stream)),
              "allreduce_bf16")

    def allreduce_sgd(self, grad_off: int, param_off: int, master_off: int, mom_off: int, first_elem: int, n_elems: int,
                      wd_split: int, hyper_ptr: int, scale: float, algo: Optional[int] = None, nblocks: int = 32,
                      stream=None) -> None:
        """K1b: reduce-scatter of the gradients + SGD-momentum on the owner's slice + broadcast of the new master and
        bf16 parameter, one kernel (csrc/collectives.cu::allreduce_sgd_kernel)."""
        if algo is None or algo == AR_ONESHOT:
            algo = AR_NVLS if (self.has_multicast and not self.emulate and self.nranks >= 4) else AR_TWOSHOT
        check(self.lib.pdj_allreduce_sgd(ctypes.byref(self._comm), grad_off, param_off, master_off, mom_off, first_elem,
                                         n_elems, wd_split, ctypes.c_void_p(hyper_ptr), scale, algo, nblocks,
                                         int(self.emulate), self._stream(stream)), "allreduce_sgd")

    def all_to_all(self, in_off: int, out_off: int, chunk_bytes: int, nblocks: int = 0, stream=None) -> None:
        if nblocks <= 0:
            nblocks = self.pick_blocks(chunk_bytes * self.nranks)
        check(self.lib.pdj_alltoall(ctypes.byref(self._comm), in_off, out_off, chunk_bytes, nblocks,
                                    int(self.emulate), self._stream(stream)), "alltoall")

    def broadcast(self, root: int, off: int, nbytes: int, nblocks: int = 32, stream=None) -> None:
        check(self.lib.pdj_broadcast(ctypes.byref(self._comm), root, off, nbytes, nblocks, int(self.emulate),
                                     self._stream(stream)), "broadcast")

    @property
    def comm_struct(self) -> PdjComm:
        return self._comm

    def channel_struct(self, channel: int) -> PdjComm:
        """A view of this communicator that uses barrier-slot channel `channel` of the signal
        pad.  Kernels that may overlap on the GPU (e.g. a gradient all-reduce on a side stream
        and an embedding push on the main stream) must not share a channel: each barrier slot's
        epoch counter is owned by one kernel at a time.  Channel 0 is the default."""
        if not 0 <= channel < _lib.NUM_CHANNELS:
            raise ValueError(f"channel must be in [0, {_lib.NUM_CHANNELS})")
        c = self._channels.get(channel)
        if c is None:
            c = PdjComm()
            ctypes.pointer(c)[0] = self._comm
            c.pad_base = channel * _lib.CHANNEL_BLOCKS
            self._channels[channel] = c
        return c

    def error_flags(self) -> list[int]:
        """Reads the timeout/error word of every local pad (debugging)."""
        import torch

        ranks = range(self.nranks) if self.emulate else [self.rank]
        out = []
        for r in ranks:
            pad = self.tensor(0, (PAD_BYTES // 4,), torch.int32, rank=r)
            out.append(int(pad[_lib.MAX_BLOCKS * 8 + _lib.MAX_BLOCKS].item()))
        return out

    def close(self) -> None:
        import torch

        if torch.cuda.is_available():
            torch.cuda.synchronize(self.device)
        if self._mc:
            self.lib.pdj_multicast_destroy(ctypes.c_void_p(self._mc))
            self._mc = None
        for a in self._arenas:
            if a:
                self.lib.pdj_arena_destroy(ctypes.c_void_p(a))
        self._arenas = []
