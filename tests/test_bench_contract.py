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
