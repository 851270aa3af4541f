"""Multi-GPU checks (skipped on boxes with a single GPU): the real multi-process path - fd
rendezvous, peer mappings, NVLS multicast, every collective against NCCL, and a short
data-parallel ResNet run whose ranks must agree on the loss."""
import json
import re
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NGPU = torch.cuda.device_count() if torch.cuda.is_available() else 0
needs2 = pytest.mark.skipif(NGPU < 2, reason="needs at least 2 GPUs")


def _torchrun(n, port, *mod_args, timeout=300):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr",
           "127.0.0.1", "--master-port", str(port)] + list(mod_args)
    return subprocess.run(cmd, cwd=ROOT, capture_output=True, text=True, timeout=timeout,
                          env=dict(os.environ, PYTHONPATH=ROOT))


@needs2
def test_collectives_match_nccl_across_processes():
    n = 2 if NGPU < 4 else 4
    r = _torchrun(n, 29711, "-m", "paddle_operator_b200.workloads.comm_check")
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "numerics: OK" in r.stdout


@needs2
def test_data_parallel_resnet_ranks_agree():
    r = _torchrun(2, 29712, "-m", "paddle_operator_b200.workloads.train_resnet", "--tiny", "--image", "64", "--batch",
                  "16", "--steps", "12", "--log-every", "12")
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    rows = []
    for m in re.findall(r"\{[^{}]*\}", r.stdout):  # two ranks share stdout: their lines may interleave or abut
        try:
            rows.append(json.loads(m))
        except ValueError:
            pass
    last = {x["rank"]: x for x in rows if x.get("step") == 12}
    assert len(last) == 2, r.stdout[-1500:]
    assert all(x["loss"] == x["loss"] for x in last.values())
    # every rank sees a different batch but holds bit-identical weights after the fused all-reduce
    sums = {x["param_checksum"] for x in last.values()}
    assert len(sums) == 1, f"ranks disagree on the parameters: {last}"
