"""Tuning sweep for the two BN reduction kernels (statistics pass, backward
reduce) over grid size / rows in flight, on ResNet-50 layer shapes at batch 256.
Each configuration runs in its own process because the knobs are read once."""
import os
import subprocess
import sys

SHAPES = [(256, 64, 112, 112), (256, 64, 56, 56), (256, 256, 56, 56), (256, 128, 28, 28), (256, 512, 28, 28),
          (256, 256, 14, 14), (256, 1024, 14, 14), (256, 512, 7, 7), (256, 2048, 7, 7)]


def child():
    import torch

    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from paddle_operator_b200.ops.bn import bn_act

    dev = torch.device("cuda", 0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    tot_f = tot_b = 0.0
    out = []
    for shp in SHAPES:
        n, c, h, w = shp
        x = torch.randn(n, c, h, w, device=dev).to(torch.bfloat16).contiguous(memory_format=torch.channels_last)
        x.requires_grad_(True)
        g = torch.ones(c, device=dev, requires_grad=True)
        b = torch.zeros(c, device=dev, requires_grad=True)
        dy = torch.randn_like(x)
        tf = tb = 0.0
        for it in range(6):
            flush.zero_()
            e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
            e0.record()
            y = bn_act(x, g, b, relu=True)
            e1.record()
            y.backward(dy)
            e2.record()
            torch.cuda.synchronize()
            if it >= 2:
                tf += e0.elapsed_time(e1) / 4
                tb += e1.elapsed_time(e2) / 4
        mb = x.numel() * 2 / 1e6
        out.append(f"{c}x{h}:{tf * 1e3:.0f}/{tb * 1e3:.0f}us")
        tot_f += tf
        tot_b += tb
    print(f"fwd {tot_f:.3f} ms bwd {tot_b:.3f} ms | " + " ".join(out), flush=True)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "child":
        child()
    else:
        for cvt in (4,):
            for bps in (2, 4):
                for us, ub in ((1, 1), (2, 1), (4, 1), (2, 2), (4, 2)):
                    env = dict(os.environ, PDJ_BN_BLOCKS_PER_SM=str(bps), PDJ_BN_STATS_UNROLL=str(us), PDJ_BN_BWD_UNROLL=str(ub),
                               PDJ_BN_MIN_ROWS="4", PDJ_BN_CVT=str(cvt))
                    r = subprocess.run([sys.executable, __file__, "child"], env=env, capture_output=True, text=True)
                    print(f"cvt={cvt} bps={bps} stats_unroll={us} bwd_unroll={ub}: {r.stdout.strip()[-400:]} {r.stderr.strip()[-200:]}", flush=True)
