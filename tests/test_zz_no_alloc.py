"""pdj_bn_fwd_noalloc, called through the C ABI exactly as an FFI user would.

That entry point exists for callers that cannot allocate device memory (no
allocator of their own, or a frozen HBM budget where cudaMalloc/cudaFree would
fail or stall peers spinning in a collective), so it keeps its scratch in a
device array built into the library.  This file checks it against the regular
pdj_bn_fwd (which takes a caller-provided workspace) on two back-to-back calls
of different shapes: the second call finds the first call's leftovers in the
shared array and must still be right.  It is named so that pytest collects it
last, and it must pass both alone in a fresh process and after the other tests.
"""
import ctypes
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _lib():
    from paddle_operator_b200.ops import _lib

    lib = _lib.load()
    vp = ctypes.c_void_p
    lib.pdj_bn_fwd_noalloc.restype = ctypes.c_int
    lib.pdj_bn_fwd_noalloc.argtypes = [vp, vp, vp, ctypes.c_int, ctypes.c_int, vp, vp, ctypes.c_float,
                                       ctypes.c_float, vp, vp, vp, vp, ctypes.c_int, vp]
    return lib


def _run_both(lib, m, c, relu, with_res, seed):
    dev = torch.device("cuda", 0)
    torch.manual_seed(seed)
    x = (torch.randn(m, c, device=dev) * 1.7 + 0.4).to(torch.bfloat16)
    res = torch.randn(m, c, device=dev).to(torch.bfloat16) if with_res else None
    gamma = torch.rand(c, device=dev) + 0.5
    beta = torch.randn(c, device=dev) * 0.1
    outs = []
    for which in ("workspace", "noalloc"):
        y = torch.full((m, c), float("nan"), device=dev, dtype=torch.bfloat16)
        rm, rv = torch.zeros(c, device=dev), torch.ones(c, device=dev)
        stats = torch.zeros(4, c, device=dev)
        st = torch.cuda.current_stream().cuda_stream
        rp = res.data_ptr() if res is not None else None
        if which == "workspace":
            ws = torch.zeros(lib.pdj_bn_workspace_bytes(), dtype=torch.uint8, device=dev)
            rc = lib.pdj_bn_fwd(x.data_ptr(), rp, y.data_ptr(), m, c, gamma.data_ptr(), beta.data_ptr(), 1e-5, 0.1,
                                rm.data_ptr(), rv.data_ptr(), stats.data_ptr(), stats[2:].data_ptr(), int(relu),
                                ws.data_ptr(), st)
        else:
            rc = lib.pdj_bn_fwd_noalloc(x.data_ptr(), rp, y.data_ptr(), m, c, gamma.data_ptr(), beta.data_ptr(), 1e-5,
                                        0.1, rm.data_ptr(), rv.data_ptr(), stats.data_ptr(), stats[2:].data_ptr(),
                                        int(relu), st)
        assert rc == 0, lib.pdj_last_error()
        torch.cuda.synchronize()
        outs.append((y, rm, rv, stats))
    return outs


def test_noalloc_matches_workspace_entry_point_twice():
    lib = _lib()
    # two calls in a row; the second one reuses the array the first one dirtied
    for seed, (m, c, relu, with_res) in enumerate([(4 * 56 * 56, 256, True, True), (3 * 7 * 7, 2048, False, False)]):
        (y0, rm0, rv0, s0), (y1, rm1, rv1, s1) = _run_both(lib, m, c, relu, with_res, seed)
        assert not torch.isnan(y1.float()).any()
        torch.testing.assert_close(y1.float(), y0.float(), atol=2e-2, rtol=2e-2)
        torch.testing.assert_close(s1, s0, atol=1e-4, rtol=1e-4)
        torch.testing.assert_close(rm1, rm0, atol=1e-5, rtol=1e-4)
        torch.testing.assert_close(rv1, rv0, atol=1e-5, rtol=1e-4)


def _fp32_reference(x, res, gamma, beta, relu, eps=1e-5):
    """Plain PyTorch fp32 BatchNorm(+add)(+ReLU) over the rows of x[M, C]: an oracle that shares no code
    with either entry point."""
    xf = x.float()
    mean, var = xf.mean(0), xf.var(0, unbiased=False)
    y = (xf - mean) * torch.rsqrt(var + eps) * gamma + beta
    if res is not None:
        y = y + res.float()
    return torch.relu(y) if relu else y


def test_noalloc_against_fp32_torch_and_leaves_device_memory_alone():
    lib = _lib()
    dev = torch.device("cuda", 0)
    m, c = 8 * 28 * 28, 512
    torch.manual_seed(11)
    x = (torch.randn(m, c, device=dev) * 1.3 - 0.2).to(torch.bfloat16)
    res = torch.randn(m, c, device=dev).to(torch.bfloat16)
    gamma, beta = torch.rand(c, device=dev) + 0.5, torch.randn(c, device=dev) * 0.1
    y = torch.empty_like(x)
    rm, rv, stats = torch.zeros(c, device=dev), torch.ones(c, device=dev), torch.zeros(4, c, device=dev)
    st = torch.cuda.current_stream().cuda_stream

    def call():
        rc = lib.pdj_bn_fwd_noalloc(x.data_ptr(), res.data_ptr(), y.data_ptr(), m, c, gamma.data_ptr(), beta.data_ptr(),
                                    1e-5, 0.1, rm.data_ptr(), rv.data_ptr(), stats.data_ptr(), stats[2:].data_ptr(), 1, st)
        assert rc == 0, lib.pdj_last_error()
        torch.cuda.synchronize()

    call()  # the first call on a device does the one-time set-up
    free_before = torch.cuda.mem_get_info()[0]
    call()
    assert torch.cuda.mem_get_info()[0] == free_before  # a warmed-up call takes no device memory
    torch.testing.assert_close(y.float(), _fp32_reference(x, res, gamma, beta, True), atol=3e-2, rtol=2e-2)


_NARROW_TILES_CHILD = '''
import ctypes, sys, torch
sys.path.insert(0, %r)
sys.path.insert(0, %r)
import test_zz_no_alloc as t
lib = t._lib()
dev = torch.device("cuda", 0)
# PDJ_BN_CVT=1 asks for one channel-vector per block: C=2048 would need 256 "last block" tickets where 64
# exist, so the library must widen the tiles; PDJ_BN_TMA=0 sends pdj_bn_fwd down the same plain-load kernels
for seed, (m, c) in enumerate([(3 * 7 * 7, 2048), (4096, 2048), (2048, 1024), (512, 64)]):
    (y0, rm0, rv0, s0), (y1, rm1, rv1, s1) = t._run_both(lib, m, c, True, False, seed)
    torch.manual_seed(seed)
    x = (torch.randn(m, c, device=dev) * 1.7 + 0.4).to(torch.bfloat16)
    gamma = torch.rand(c, device=dev) + 0.5
    beta = torch.randn(c, device=dev) * 0.1
    ref = t._fp32_reference(x, None, gamma, beta, True)
    for name, y in (("pdj_bn_fwd", y0), ("pdj_bn_fwd_noalloc", y1)):
        err = float((y.float() - ref).abs().max())
        assert err < 5e-2, (name, m, c, err)
print("narrow-tiles ok")
'''


def test_narrow_channel_tiles_do_not_overrun_the_64_tickets():
    """PDJ_BN_CVT=1|2 used to give more channel tiles than ticket counters for C >= 1024 (tickets aliased the
    partial sums: silently wrong statistics).  The knob is read once per process, hence the child process."""
    for cvt in ("1", "2"):
        env = dict(os.environ, PDJ_BN_CVT=cvt, PDJ_BN_TMA="0")
        r = subprocess.run([sys.executable, "-c", _NARROW_TILES_CHILD % (ROOT, os.path.join(ROOT, "tests"))],
                           capture_output=True, text=True, timeout=600, env=env, cwd=ROOT)
        assert r.returncode == 0 and "narrow-tiles ok" in r.stdout, (r.stdout + r.stderr)[-2000:]


def test_noalloc_rejects_oversize():
    lib = _lib()
    dev = torch.device("cuda", 0)
    t = torch.zeros(8, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    p = t.data_ptr()
    for m, c in [(16, 4096), ((1 << 24) + 1, 64), (16, 24), (0, 64)]:
        rc = lib.pdj_bn_fwd_noalloc(p, None, p, m, c, p, p, 1e-5, 0.1, None, None, p, p, 1, st)
        assert rc != 0
        assert b"pdj_bn_fwd_noalloc" in lib.pdj_last_error()
    torch.cuda.synchronize()
