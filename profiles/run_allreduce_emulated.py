"""All ranks of an 8-way fused all-reduce as ONE cooperative launch on a single GPU (the emulation mode the
tests use), so that ncu can profile the kernel body: peer loads / stores, system-scope flags, fp32 reduction.
Bandwidth numbers from this run are HBM-local and mean nothing for NVLink; see profiles/r1/sweep_n8.json."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from paddle_operator_b200.parallel.symm import AR_ONESHOT, AR_TWOSHOT, Communicator  # noqa: E402

comm = Communicator.emulated(8, arena_bytes=160 << 20)
n = 32 << 20  # bf16 elements: 64 MiB per rank
a, b = comm.alloc(n * 2), comm.alloc(n * 2)
for r in range(8):
    comm.tensor(a, (n,), torch.bfloat16, rank=r).normal_()
for _ in range(3):
    comm.all_reduce_bf16(a, a, n, scale=0.125, algo=AR_TWOSHOT, nblocks=16)
    comm.all_reduce_bf16(a, b, 8 * 4096, scale=0.125, algo=AR_ONESHOT, nblocks=4)
    comm.all_to_all(a, b, (n * 2) // 8, nblocks=16)
torch.cuda.synchronize()
print("ok", comm.error_flags())
comm.close()
