"""The parts of the bench.py contract that do not need a GPU."""
import importlib.util
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _load_bench():
    spec = importlib.util.spec_from_file_location("bench_under_test", os.path.join(ROOT, "bench.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_reference_arm_reports_unavailable_and_exits_zero():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "1", "--steps", "2",
                        "--warmup", "1"], capture_output=True, text=True, timeout=120, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-500:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and "unavailable" in line and "\n" not in line["unavailable"]


def test_reference_arm_prints_once_under_torchrun_env():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2"],
                       capture_output=True, text=True, timeout=120, cwd=ROOT, env=env)
    assert r.returncode == 0 and r.stdout.strip() == ""  # only rank 0 prints the line


def test_clock_sampler_watches_only_the_gpus_in_use(monkeypatch):
    bench = _load_bench()
    monkeypatch.setenv("WORLD_SIZE", "4")
    monkeypatch.delenv("CUDA_VISIBLE_DEVICES", raising=False)
    assert bench.ClockSampler.gpus_in_use() == "0,1,2,3"
    monkeypatch.setenv("CUDA_VISIBLE_DEVICES", "4,5,6,7")
    monkeypatch.setenv("WORLD_SIZE", "2")
    assert bench.ClockSampler.gpus_in_use() == "4,5"
    monkeypatch.setenv("WORLD_SIZE", "1")
    monkeypatch.delenv("CUDA_VISIBLE_DEVICES", raising=False)
    assert bench.ClockSampler.gpus_in_use() == "0"
    # a box without nvidia-smi must not break the bench
    s = bench.ClockSampler()
    s.start()
    out = s.stop()
    assert set(out) >= {"sm_mhz", "sm_max_mhz", "reasons"}


def _reference_adagrad(deep, wide, acc_d, acc_w, rows, gd, gw, lr, eps):
    """Dense-gradient oracle: sum duplicate rows, then one Adagrad step on the touched rows."""
    import torch

    G = torch.zeros_like(deep).index_add_(0, rows, gd)
    Gw = torch.zeros_like(wide).index_add_(0, rows, gw)
    touched = torch.zeros(deep.shape[0], dtype=torch.bool)
    touched[rows] = True
    acc_d = acc_d + G * G
    acc_w = acc_w + Gw * Gw
    deep = torch.where(touched[:, None], deep - lr * G * torch.rsqrt(acc_d + eps), deep)
    wide = torch.where(touched, wide - lr * Gw * torch.rsqrt(acc_w + eps), wide)
    return deep, wide


def test_nccl_baseline_sparse_adagrad_merges_duplicates_without_host_sync():
    """The Wide&Deep baseline arm's sort-merge update equals summing duplicate rows first."""
    import torch

    from paddle_operator_b200.workloads.bench_wide_deep import NcclShardedJob

    job = NcclShardedJob(batch=64, rows=97, dim=8, rank=0, world=1, dev=torch.device("cpu"))
    torch.manual_seed(3)
    m = 500
    rows = torch.randint(0, job.rps, (m,))
    rows[::7] = job.rps  # padding slots hit the spare row
    gd, gw = torch.randn(m, 8), torch.randn(m)
    gd[::7], gw[::7] = 0.0, 0.0
    d0, w0 = job.deep.clone(), job.wide.clone()
    job._sparse_adagrad(rows, gd, gw)
    keep = rows != job.rps
    d_ref, w_ref = _reference_adagrad(d0[:-1], w0[:-1], torch.zeros(job.rps, 8), torch.zeros(job.rps), rows[keep],
                                      gd[keep], gw[keep], job.lr, job.eps)
    assert torch.allclose(job.deep[:-1], d_ref, atol=1e-6) and torch.allclose(job.wide[:-1], w_ref, atol=1e-6)
    assert torch.equal(job.deep[-1], d0[-1])  # zero gradients leave the spare row alone


def test_nccl_baseline_step_runs_and_learns_on_cpu():
    import torch

    from paddle_operator_b200.workloads.bench_wide_deep import NcclShardedJob

    job = NcclShardedJob(batch=256, rows=1000, dim=8, rank=0, world=1, dev=torch.device("cpu"))
    losses = [float(job.step_device()) for _ in range(40)]
    # (each step draws a fresh random batch, so single losses jitter: compare averages)
    assert sum(losses[-10:]) / 10 < sum(losses[:5]) / 5 and not bool(job.overflowed)


_TWO_RANK_ROUTING = '''
import sys, torch, torch.distributed as dist
sys.path.insert(0, %r)
dist.init_process_group("gloo")
r, w = dist.get_rank(), dist.get_world_size()
from paddle_operator_b200.workloads.bench_wide_deep import NcclShardedJob
job = NcclShardedJob(batch=128, rows=1000, dim=8, rank=r, world=w, dev=torch.device("cpu"))
shards = [torch.zeros(job.rps, 8) for _ in range(w)]
dist.all_gather(shards, job.deep[:job.rps].contiguous())
full = torch.empty(job.num_rows, 8)
for k in range(w):
    full[k::w] = shards[k]
flat = job.ids.reshape(-1)
order, slot, local_rows = job._route(flat)
back = job._a2a(job.deep[local_rows])
got = torch.empty(job.n, 8).index_copy_(0, order, back[slot])
assert torch.equal(got, full[flat]), "pull mismatch"
l0 = float(job.step_device())
for _ in range(20):
    l1 = float(job.step_device())
assert l1 < l0 and not bool(job.overflowed), (l0, l1)
sys.stdout.write("rank " + str(r) + " ok" + chr(10)); sys.stdout.flush()  # one write: the two ranks share a pipe
'''


def test_nccl_baseline_fixed_capacity_exchange_two_ranks(tmp_path):
    """The baseline's fixed-capacity all-to-all routing returns exactly the owners' rows (gloo, 2 ranks)."""
    script = tmp_path / "route.py"
    script.write_text(_TWO_RANK_ROUTING % ROOT)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29537", str(script)],
                       capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert r.returncode == 0, (r.stdout + r.stderr)[-1500:]
    assert "rank 0 ok" in r.stdout and "rank 1 ok" in r.stdout
