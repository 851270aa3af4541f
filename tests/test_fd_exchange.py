"""FdExchange (the rendezvous that carries arena file descriptors between rank
processes) with real processes on CPU."""
import multiprocessing as mp
import os
import tempfile
import uuid


def _worker(key, rank, n, q):
    from paddle_operator_b200.parallel.symm import FdExchange

    ex = FdExchange(key, rank, n, timeout=30)
    f = tempfile.TemporaryFile()
    f.write(f"hello from {rank}".encode())
    f.flush()
    got = ex.exchange(f.fileno(), f"blob{rank}".encode())
    texts = []
    for r, (fd, blob) in enumerate(got):
        assert blob == f"blob{r}".encode()
        if r == rank:
            continue
        with os.fdopen(fd, "rb") as g:
            g.seek(0)
            texts.append(g.read().decode())
    ex.barrier()
    flags = ex.exchange(None, b"x")
    assert all(fd is None for fd, _ in flags)
    q.put((rank, sorted(texts)))


def test_fd_exchange_three_processes():
    ctx = mp.get_context("spawn")
    key = "test-" + uuid.uuid4().hex[:8]
    q = ctx.Queue()
    n = 3
    ps = [ctx.Process(target=_worker, args=(key, r, n, q)) for r in range(n)]
    for p in ps:
        p.start()
    res = dict(q.get(timeout=60) for _ in range(n))
    for p in ps:
        p.join(30)
        assert p.exitcode == 0
    for r in range(n):
        assert res[r] == sorted(f"hello from {o}" for o in range(n) if o != r)


def test_native_runtime_host_side_under_sanitizers(tmp_path):
    """ASAN+UBSAN and TSAN builds of csrc/comm_runtime.cu + tma_util.cu driven by eight threads
    (tests/native/sanitize_host.cpp).  Without a GPU every entry point must fail cleanly; with one the threads
    create / export / import / destroy arenas concurrently."""
    import shutil
    import subprocess

    import pytest

    if shutil.which("nvcc") is None:
        pytest.skip("nvcc not available")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([os.path.join(root, "tests", "native", "build_sanitize.sh"), str(tmp_path)], capture_output=True,
                       text=True, timeout=900)
    assert r.returncode == 0 and "sanitize: ok" in r.stdout, (r.stdout + r.stderr)[-3000:]
    assert "ERROR: AddressSanitizer" not in r.stderr and "WARNING: ThreadSanitizer" not in r.stderr
