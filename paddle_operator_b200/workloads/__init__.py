"""The programs a rank runs (the user images of the reference's examples) and their benchmark harnesses."""


def require_gpu(program: str, dry_run_flag: str = "") -> None:
    """GPU workloads fail with one clear line (exit code 3) on a box without a usable GPU, instead of a CUDA
    initialisation traceback in the rank's log."""
    import sys

    import torch

    if torch.cuda.is_available():
        return
    hint = f"; pass {dry_run_flag} (or PDJ_CPU_DRY_RUN=1) to exercise only the role plumbing" if dry_run_flag else ""
    print(f"{program}: no usable NVIDIA GPU on this machine - this workload runs on B200 (sm_100a) only{hint}",
          file=sys.stderr, flush=True)
    raise SystemExit(3)
