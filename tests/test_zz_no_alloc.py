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

import pytest
import torch

pytestmark = pytest.mark.gpu


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


def test_noalloc_rejects_oversize():
    lib = _lib()
    dev = torch.device("cuda", 0)
    t = torch.zeros(8, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    p = t.data_ptr()
    for m, c in [(16, 4096), ((1 << 24) + 1, 64), (16, 24)]:
        rc = lib.pdj_bn_fwd_noalloc(p, None, p, m, c, p, p, 1e-5, 0.1, None, None, p, p, 1, st)
        assert rc != 0
        assert b"pdj_bn_fwd_noalloc" in lib.pdj_last_error()
    torch.cuda.synchronize()
