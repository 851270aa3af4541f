"""All ranks of an 8-way fused all-reduce as ONE cooperative launch on a single GPU (the emulation mode the
tests use), so that ncu can profile the kernel body: peer loads / stores, system-scope flags, fp32 reduction.
Bandwidth numbers from this run are HBM-local and mean nothing for NVLink; see profiles/r1/sweep_n8.json."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from paddle_operator_b200.parallel.symm import AR_ONESHOT, AR_TWOSHOT, Communicator  # noqa: E402

comm = Communicator.emulated(8, arena_bytes=288 << 20)
n = 32 << 20  # bf16 elements: 64 MiB per rank
a, b = comm.alloc(n * 2), comm.alloc(n * 2)
for r in range(8):
    comm.tensor(a, (n,), torch.bfloat16, rank=r).normal_()
for _ in range(3):
    comm.all_reduce_bf16(a, a, n, scale=0.125, algo=AR_TWOSHOT, nblocks=16)
    comm.all_reduce_bf16(a, b, 8 * 4096, scale=0.125, algo=AR_ONESHOT, nblocks=4)
    comm.all_to_all(a, b, (n * 2) // 8, nblocks=16)
# K1b: reduce-scatter + SGD-momentum + master / parameter broadcast over an 8 Mi-element bucket
m = 8 << 20
g_off, p_off, w_off, v_off = comm.alloc(m * 2), comm.alloc(m * 2), comm.alloc(m * 4), comm.alloc(m * 4)
hyper = torch.tensor([0.1, 0.9, 1e-4, 1.0], device="cuda")
for r in range(8):
    comm.tensor(g_off, (m,), torch.bfloat16, rank=r).normal_()
    comm.tensor(w_off, (m,), torch.float32, rank=r).fill_(1.0)
    comm.tensor(v_off, (m,), torch.float32, rank=r).zero_()
for _ in range(3):
    comm.allreduce_sgd(g_off, p_off, w_off, v_off, 0, m, m, hyper.data_ptr(), 0.125, nblocks=16)
torch.cuda.synchronize()
print("ok", comm.error_flags())
comm.close()
