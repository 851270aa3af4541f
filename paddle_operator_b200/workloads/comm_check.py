"""Multi-GPU correctness check and bandwidth sweep of the peer-memory collectives
against NCCL (BASELINE.json config 5: all-reduce / all-to-all, 1 KB - 1 GB, at
2/4/8 GPUs, vs NCCL and vs 900 GB/s per direction).

  torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
      -m paddle_operator_b200.workloads.comm_check [--sweep] [--max-mb 1024] [--out FILE]

Device-timed with CUDA events, max over ranks.  NCCL is the oracle and the
baseline; it is not on the product path.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

from ..parallel.symm import AR_NVLS, AR_ONESHOT, AR_TWOSHOT, Communicator

ALGO_NAMES = {AR_TWOSHOT: "twoshot", AR_ONESHOT: "oneshot", AR_NVLS: "nvls"}


def _time(fn, iters, dev, warmup=3):
    for _ in range(warmup):
        fn()
    dist.barrier(device_ids=[dev.index])
    torch.cuda.synchronize(dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize(dev)
    t = torch.tensor([e0.elapsed_time(e1) / iters], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())  # ms


def check(comm: Communicator, dev, rank, world) -> list[str]:
    """Numerics of every collective against NCCL / direct construction."""
    fails = []
    comm.reset_allocations()
    n = 8 * 4099 * world
    a, b = comm.alloc(n * 2), comm.alloc(n * 4)
    x = comm.tensor(a, (n,), torch.bfloat16)
    algos = [AR_TWOSHOT, AR_ONESHOT] + ([AR_NVLS] if comm.has_multicast else [])
    for algo in algos:
        for it in range(3):
            torch.manual_seed(100 * it + rank)
            v = torch.randn(n, device=dev).to(torch.bfloat16)
            x.copy_(v)
            ref = v.float().clone()
            dist.all_reduce(ref)
            ref /= world
            out_off = a if algo != AR_ONESHOT else b
            comm.all_reduce_bf16(a, out_off, n, scale=1.0 / world, algo=algo, nblocks=16)
            torch.cuda.synchronize(dev)
            got = comm.tensor(out_off, (n,), torch.bfloat16).float()
            err = (got - ref).abs().max().item()
            if not err < 3e-2:
                fails.append(f"allreduce {ALGO_NAMES[algo]} it{it}: max err {err}")
    # fp32 output variant
    torch.manual_seed(rank)
    v = torch.randn(n, device=dev).to(torch.bfloat16)
    x.copy_(v)
    ref = v.float().clone()
    dist.all_reduce(ref)
    comm.all_reduce_bf16(a, b, n, scale=1.0, algo=AR_TWOSHOT, out_f32=True, nblocks=16)
    torch.cuda.synchronize(dev)
    err = (comm.tensor(b, (n,), torch.float32) - ref).abs().max().item()
    if not err < 1e-4:
        fails.append(f"allreduce f32-out: max err {err}")
    # all-to-all
    comm.reset_allocations()
    chunk = 16 * 1031
    a, b = comm.alloc(chunk * world), comm.alloc(chunk * world)
    src = torch.randint(0, 255, (world, chunk), dtype=torch.uint8, device=dev)
    comm.tensor(a, (world, chunk), torch.uint8).copy_(src)
    ref = torch.empty_like(src)
    dist.all_to_all_single(ref, src)
    comm.all_to_all(a, b, chunk, nblocks=8)
    torch.cuda.synchronize(dev)
    if not torch.equal(comm.tensor(b, (world, chunk), torch.uint8), ref):
        fails.append("alltoall mismatch")
    # broadcast
    root = world - 1
    t = comm.tensor(a, (world, chunk), torch.uint8)
    ref = t.clone()
    dist.broadcast(ref, root)
    comm.broadcast(root, a, chunk * world, nblocks=8)
    torch.cuda.synchronize(dev)
    if not torch.equal(t, ref):
        fails.append("broadcast mismatch")
    comm.barrier()
    torch.cuda.synchronize(dev)
    if comm.error_flags()[0] != 0:
        fails.append("signal pad error flag set")
    return fails


def allreduce_check(dev, rank, world, nbytes: int = 64 << 20, key: str = "archeck") -> dict:
    """One all-reduce per algorithm on `nbytes` of random bf16 against `dist.all_reduce` (fp32 sum of the
    same bf16 inputs).  Returns {"twoshot": "ok"|"MISMATCH", "nvls": ..., "oneshot": ..., "max_abs_err": e,
    "ok": bool}; identical on every rank (errors are max-reduced), so all ranks can act on it together.
    bench.py runs it before the timed arms: a wrong NVLS sum must not hide behind a finite loss."""
    comm = Communicator.from_env(arena_bytes=2 * nbytes + (32 << 20), key=key)
    out = {"bytes": nbytes, "nvls_available": bool(comm.has_multicast)}
    n = nbytes // 2
    a, b = comm.alloc(nbytes), comm.alloc(nbytes)
    x = comm.tensor(a, (n,), torch.bfloat16)
    torch.manual_seed(4321 + rank)
    v = torch.randn(n, device=dev).to(torch.bfloat16)
    ref = v.float()
    dist.all_reduce(ref)
    ref /= world
    worst = 0.0
    cases = [("twoshot", AR_TWOSHOT, n, a)] + ([("nvls", AR_NVLS, n, a)] if comm.has_multicast else [])
    cases.append(("oneshot", AR_ONESHOT, (256 << 10) // 2, b))  # the small-message algorithm, out of place
    for name, algo, m, out_off in cases:
        x.copy_(v)
        comm.barrier()
        comm.all_reduce_bf16(a, out_off, m, scale=1.0 / world, algo=algo, nblocks=24)
        torch.cuda.synchronize(dev)
        got = comm.tensor(out_off, (m,), torch.bfloat16).float()
        err = torch.tensor([float((got - ref[:m]).abs().max().item())], dtype=torch.float64, device=dev)
        dist.all_reduce(err, op=dist.ReduceOp.MAX)
        e = float(err.item())
        worst = max(worst, e)
        # bf16 rounding of a mean of N(0,1) samples: half an ulp at |x| < 4 is 2^-7
        out[name] = "ok" if e < 2e-2 else "MISMATCH"
        comm.barrier()
    if not comm.has_multicast:
        out["nvls"] = "not available (no multicast object on this box / world size)"
    torch.cuda.synchronize(dev)
    flags = comm.error_flags()[0]
    out["max_abs_err"] = worst
    out["ok"] = all(out[nm] == "ok" for nm, *_ in cases) and flags == 0
    dist.barrier(device_ids=[dev.index])
    comm.close()
    return out


def sweep(comm: Communicator, dev, rank, world, max_bytes, blocks_list) -> list[dict]:
    rows = []
    comm.reset_allocations()
    a, b = comm.alloc(max_bytes), comm.alloc(max_bytes)
    size = 1024
    nccl_buf = torch.zeros(max_bytes // 2, dtype=torch.bfloat16, device=dev)
    nccl_out = torch.zeros(max_bytes // 2, dtype=torch.bfloat16, device=dev)
    while size <= max_bytes:
        n = size // 2
        iters = 50 if size <= (4 << 20) else (20 if size <= (64 << 20) else 8)
        t_nccl = _time(lambda: dist.all_reduce(nccl_buf[:n]), iters, dev)
        row = {"op": "allreduce", "bytes": size, "n_gpus": world, "nccl_ms": t_nccl,
               "nccl_busbw_gbs": size / t_nccl / 1e6 * 2 * (world - 1) / world}
        best = None
        algos = [AR_TWOSHOT] + ([AR_ONESHOT] if size <= (1 << 20) else []) + ([AR_NVLS] if comm.has_multicast else [])
        for algo in algos:
            for nb in blocks_list:
                out = b if algo == AR_ONESHOT else a
                t = _time(lambda: comm.all_reduce_bf16(a, out, n, scale=1.0 / world, algo=algo, nblocks=nb), iters, dev)
                row[f"{ALGO_NAMES[algo]}_b{nb}_ms"] = t
                if best is None or t < best[0]:
                    best = (t, ALGO_NAMES[algo], nb)
        # what the PRODUCT runs for this size: pick_algo + pick_blocks, not the best cell of the table above
        p_algo = comm.pick_algo(size, in_place=True)
        p_blocks = comm.pick_blocks(size, p_algo)
        t_prod = _time(lambda: comm.all_reduce_bf16(a, a, n, scale=1.0 / world, algo=p_algo, nblocks=p_blocks), iters, dev)
        row.update({"product_ms": t_prod, "product_algo": ALGO_NAMES[p_algo], "product_blocks": p_blocks,
                    "product_busbw_gbs": size / t_prod / 1e6 * 2 * (world - 1) / world,
                    "product_speedup_vs_nccl": t_nccl / t_prod,
                    "best_ms": best[0], "best_algo": best[1], "best_blocks": best[2]})
        rows.append(row)
        # all-to-all (size = total bytes sent per rank)
        if size >= 16 * world and size % (16 * world) == 0:
            chunk = size // world
            t_nccl = _time(lambda: dist.all_to_all_single(nccl_out[:n], nccl_buf[:n]), iters, dev)
            best = None
            r2 = {"op": "alltoall", "bytes": size, "n_gpus": world, "nccl_ms": t_nccl,
                  "nccl_busbw_gbs": size / t_nccl / 1e6 * (world - 1) / world}
            for nb in blocks_list:
                t = _time(lambda: comm.all_to_all(a, b, chunk, nblocks=nb), iters, dev)
                r2[f"b{nb}_ms"] = t
                if best is None or t < best[0]:
                    best = (t, nb)
            t_prod = _time(lambda: comm.all_to_all(a, b, chunk), iters, dev)  # default block count = pick_blocks
            r2.update({"product_ms": t_prod, "product_blocks": comm.pick_blocks(size),
                       "product_busbw_gbs": size / t_prod / 1e6 * (world - 1) / world,
                       "product_speedup_vs_nccl": t_nccl / t_prod, "best_ms": best[0], "best_blocks": best[1]})
            rows.append(r2)
        size *= 4
    return rows


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--sweep", action="store_true")
    ap.add_argument("--max-mb", type=int, default=256)
    ap.add_argument("--blocks", default="16,64,148")
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    max_bytes = args.max_mb << 20
    comm = Communicator.from_env(arena_bytes=(2 * max_bytes + (64 << 20)) if args.sweep else (256 << 20))
    if rank == 0:
        print(f"[comm_check] world={world} arena={comm.arena_bytes >> 20} MiB multicast={comm.has_multicast}", flush=True)
    fails = check(comm, dev, rank, world)
    flag = torch.tensor([len(fails)], device=dev)
    dist.all_reduce(flag)
    if rank == 0:
        print(f"[comm_check] numerics: {'OK' if flag.item() == 0 else 'FAIL'} {fails}", flush=True)
    rows = []
    if args.sweep and flag.item() == 0:
        rows = sweep(comm, dev, rank, world, max_bytes, [int(b) for b in args.blocks.split(",")])
        if rank == 0:
            for r in rows:
                print(json.dumps(r), flush=True)
            if args.out:
                with open(args.out, "w") as f:
                    json.dump({"world": world, "multicast": comm.has_multicast, "rows": rows}, f, indent=1)
    torch.cuda.synchronize(dev)
    dist.barrier(device_ids=[dev.index])
    comm.close()
    dist.destroy_process_group()
    return 1 if flag.item() else 0


if __name__ == "__main__":
    sys.exit(main())
