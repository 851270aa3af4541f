// K1 / K4 / K5 / K6: collectives issued from inside one kernel over NVSwitch
// peer memory (P2P ld/st on VMM-mapped peer arenas, or NVLS multimem.*).
//
// Reference parity: Collective mode only publishes PADDLE_TRAINER_ENDPOINTS /
// PADDLE_TRAINERS_NUM / PADDLE_TRAINER_ID (controllers/paddlejob_helper.go:
// 261-265,319-322); the gradient all-reduce those imply is executed by Paddle
// through NCCL.  These kernels are that data plane, fused with 1/N scaling and
// the dtype cast, sized for one 8xB200 NVSwitch domain.
//
// All buffers are named by byte offset into the symmetric arena, so the same
// offset is valid on every rank.
#include "common.cuh"
#include "../../include/pdjob_b200.h"

namespace {

template <int NR>
struct Acc8 {
  float v[8];
  __device__ __forceinline__ void zero() {
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = 0.f;
  }
  __device__ __forceinline__ void add(const uint4& u) {
    float2 a = pdj_bf2_to_f2(u.x), b = pdj_bf2_to_f2(u.y), c = pdj_bf2_to_f2(u.z), d = pdj_bf2_to_f2(u.w);
    v[0] += a.x; v[1] += a.y; v[2] += b.x; v[3] += b.y;
    v[4] += c.x; v[5] += c.y; v[6] += d.x; v[7] += d.y;
  }
  __device__ __forceinline__ uint4 pack(float s) const {
    uint4 o;
    o.x = pdj_f2_to_bf2(v[0] * s, v[1] * s);
    o.y = pdj_f2_to_bf2(v[2] * s, v[3] * s);
    o.z = pdj_f2_to_bf2(v[4] * s, v[5] * s);
    o.w = pdj_f2_to_bf2(v[6] * s, v[7] * s);
    return o;
  }
};

// ---------------------------------------------------------------------------
// Device barrier (K4).  One block per rank is enough.
__global__ void barrier_kernel(PdjComm c) {
  const int rank = c.rank + blockIdx.y;
  PdjSignalPad* me = pdj_pad(c, rank);
  const uint32_t e = me->epoch[pdj_slot(c)];
  pdj_block_barrier(c, rank, e + 1);
  if (threadIdx.x == 0) me->epoch[pdj_slot(c)] = e + 1;
}

// ---------------------------------------------------------------------------
// Two-shot all-reduce, bf16 in, fp32 accumulate, x scale, bf16 (or fp32) out.
// Rank r owns the r-th slice: it pulls that slice from every peer, reduces,
// and pushes the result into every peer's output.  in == out is allowed
// because only the owner ever reads its slice of a peer's input.
template <int NR, bool OUT_F32>
__global__ void __launch_bounds__(512)
allreduce_twoshot_kernel(PdjComm c, size_t in_off, size_t out_off, size_t n_vec, float scale) {
  const int rank = c.rank + blockIdx.y;
  PdjSignalPad* me = pdj_pad(c, rank);
  const uint32_t e = me->epoch[pdj_slot(c)];
  pdj_block_barrier(c, rank, e + 1);  // every rank's input is complete

  const size_t per = (n_vec + NR - 1) / NR;
  const size_t beg = (size_t)rank * per;
  const size_t end = beg + per < n_vec ? beg + per : n_vec;
  const char* src[NR];
  char* dst[NR];
#pragma unroll
  for (int p = 0; p < NR; ++p) {
    const int peer = (rank + p) % NR;  // stagger so that links are used evenly
    src[p] = c.base[peer] + in_off;
    dst[p] = c.base[peer] + out_off;
  }
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = beg + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < end; i += stride) {
    uint4 u[NR];
#pragma unroll
    for (int p = 0; p < NR; ++p) u[p] = pdj_ld16(src[p] + i * 16);
    Acc8<NR> acc;
    acc.zero();
#pragma unroll
    for (int p = 0; p < NR; ++p) acc.add(u[p]);
    if (OUT_F32) {
      uint4 lo, hi;
      lo.x = __float_as_uint(acc.v[0] * scale); lo.y = __float_as_uint(acc.v[1] * scale);
      lo.z = __float_as_uint(acc.v[2] * scale); lo.w = __float_as_uint(acc.v[3] * scale);
      hi.x = __float_as_uint(acc.v[4] * scale); hi.y = __float_as_uint(acc.v[5] * scale);
      hi.z = __float_as_uint(acc.v[6] * scale); hi.w = __float_as_uint(acc.v[7] * scale);
#pragma unroll
      for (int p = 0; p < NR; ++p) {
        pdj_st16(dst[p] + i * 32, lo);
        pdj_st16(dst[p] + i * 32 + 16, hi);
      }
    } else {
      const uint4 o = acc.pack(scale);
#pragma unroll
      for (int p = 0; p < NR; ++p) pdj_st16(dst[p] + i * 16, o);
    }
  }
  pdj_block_barrier(c, rank, e + 2);  // every peer's pushes into my output have landed
  if (threadIdx.x == 0) me->epoch[pdj_slot(c)] = e + 2;
}

// One-shot all-reduce for latency-bound sizes: every rank pulls the whole
// message from every peer.  out must not alias in.  Output is local only.
template <int NR>
__global__ void __launch_bounds__(512)
allreduce_oneshot_kernel(PdjComm c, size_t in_off, size_t out_off, size_t n_vec, float scale) {
  const int rank = c.rank + blockIdx.y;
  PdjSignalPad* me = pdj_pad(c, rank);
  const uint32_t e = me->epoch[pdj_slot(c)];
  pdj_block_barrier(c, rank, e + 1);
  const char* src[NR];
#pragma unroll
  for (int p = 0; p < NR; ++p) src[p] = c.base[(rank + p) % NR] + in_off;
  char* dst = c.base[rank] + out_off;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_vec; i += stride) {
    uint4 u[NR];
#pragma unroll
    for (int p = 0; p < NR; ++p) u[p] = pdj_ld16(src[p] + i * 16);
    // Sum in rank order so that every rank produces bit-identical results.
    Acc8<NR> acc;
    acc.zero();
#pragma unroll
    for (int q = 0; q < NR; ++q) {
#pragma unroll
      for (int p = 0; p < NR; ++p)
        if ((rank + p) % NR == q) acc.add(u[p]);
    }
    pdj_st16(dst + i * 16, acc.pack(scale));
  }
  pdj_block_barrier(c, rank, e + 2);  // peers are done reading my input
  if (threadIdx.x == 0) me->epoch[pdj_slot(c)] = e + 2;
}

// NVLS all-reduce: the switch reduces (multimem.ld_reduce) and broadcasts
// (multimem.st); each rank handles 1/N of the message.
__global__ void __launch_bounds__(512)
allreduce_nvls_kernel(PdjComm c, size_t in_off, size_t out_off, size_t n_vec, float scale) {
  const int rank = c.rank + blockIdx.y;
  const int NR = c.nranks;
  PdjSignalPad* me = pdj_pad(c, rank);
  const uint32_t e = me->epoch[pdj_slot(c)];
  pdj_block_barrier(c, rank, e + 1);
  const size_t per = (n_vec + NR - 1) / NR;
  const size_t beg = (size_t)rank * per;
  const size_t end = beg + per < n_vec ? beg + per : n_vec;
  const char* src = c.mc_base + in_off;
  char* dst = c.mc_base + out_off;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = beg + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < end; i += stride) {
    uint4 u;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(u.x), "=r"(u.y), "=r"(u.z), "=r"(u.w)
                 : "l"(src + i * 16)
                 : "memory");
    float2 a = pdj_bf2_to_f2(u.x), b = pdj_bf2_to_f2(u.y), cc = pdj_bf2_to_f2(u.z), d = pdj_bf2_to_f2(u.w);
    uint4 o;
    o.x = pdj_f2_to_bf2(a.x * scale, a.y * scale);
    o.y = pdj_f2_to_bf2(b.x * scale, b.y * scale);
    o.z = pdj_f2_to_bf2(cc.x * scale, cc.y * scale);
    o.w = pdj_f2_to_bf2(d.x * scale, d.y * scale);
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(dst + i * 16),
                 "r"(o.x), "r"(o.y), "r"(o.z), "r"(o.w)
                 : "memory");
  }
  pdj_block_barrier(c, rank, e + 2);
  if (threadIdx.x == 0) me->epoch[pdj_slot(c)] = e + 2;
}

// ---------------------------------------------------------------------------
// K1b: gradient reduce-scatter + SGD-momentum + parameter broadcast in ONE kernel per bucket.
// The flat buffers (bf16 gradients and parameters, fp32 masters and momentum) all live in the arena at the same
// offsets on every rank.  Rank r owns the r-th slice of the bucket: it pulls that slice of every peer's gradient
// (or lets the switch add them: multimem.ld_reduce), applies
//     g = sum * scale * hyper[3] + wd * w;  v = mu * v + g;  w -= lr * v
// to its fp32 master and momentum, and pushes the new fp32 master and bf16 parameter into every rank's copy (peer
// stores or multimem.st).  Reduced gradients never travel back and no rank runs an optimizer pass over the whole
// model: each element is updated once, by its owner, and every rank ends up with bit-identical parameters.  Only the
// owner's momentum slice is current (see CollectiveTrainer.sync_state for checkpoints and re-forms).
// hyper = {lr, momentum, weight_decay, grad_scale} in device memory, as for pdj_sgd_momentum.
template <int NR, bool NVLS>
__global__ void __launch_bounds__(512)
allreduce_sgd_kernel(PdjComm c, size_t grad_off, size_t param_off, size_t master_off, size_t mom_off, size_t vec0,
                     size_t n_vec, size_t wd_split_vec, const float* __restrict__ hyper, float scale) {
  const int rank = c.rank + blockIdx.y;
  PdjSignalPad* me = pdj_pad(c, rank);
  const uint32_t e = me->epoch[pdj_slot(c)];
  pdj_block_barrier(c, rank, e + 1);  // every rank's gradients of this bucket are complete
  const float lr = hyper[0], mu = hyper[1], wd = hyper[2], gs = hyper[3] * scale;
  const size_t per = (n_vec + NR - 1) / NR;
  const size_t beg = (size_t)rank * per;
  const size_t end = beg + per < n_vec ? beg + per : n_vec;
  char* mine = c.base[rank];
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = beg + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < end; i += stride) {
    const size_t g8 = vec0 + i;  // index of this group of 8 elements in the flat buffers
    float g[8];
    if (NVLS) {
      uint4 u;
      asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                   : "=r"(u.x), "=r"(u.y), "=r"(u.z), "=r"(u.w)
                   : "l"(c.mc_base + grad_off + g8 * 16)
                   : "memory");
      const float2 a = pdj_bf2_to_f2(u.x), b = pdj_bf2_to_f2(u.y), cc = pdj_bf2_to_f2(u.z), d = pdj_bf2_to_f2(u.w);
      g[0] = a.x; g[1] = a.y; g[2] = b.x; g[3] = b.y; g[4] = cc.x; g[5] = cc.y; g[6] = d.x; g[7] = d.y;
    } else {
      uint4 u[NR];
#pragma unroll
      for (int p = 0; p < NR; ++p) u[p] = pdj_ld16(c.base[(rank + p) % NR] + grad_off + g8 * 16);
      // add in rank order: the sum does not depend on which rank owns the slice
      Acc8<NR> acc;
      acc.zero();
#pragma unroll
      for (int q = 0; q < NR; ++q) {
#pragma unroll
        for (int p = 0; p < NR; ++p)
          if ((rank + p) % NR == q) acc.add(u[p]);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) g[j] = acc.v[j];
    }
    float4* wp = reinterpret_cast<float4*>(mine + master_off + g8 * 32);
    float4* mp = reinterpret_cast<float4*>(mine + mom_off + g8 * 32);
    const float4 w0 = wp[0], w1 = wp[1], m0 = mp[0], m1 = mp[1];
    float ww[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
    float mm[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
    const float dec = g8 < wd_split_vec ? wd : 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float gj = fmaf(dec, ww[j], g[j] * gs);
      mm[j] = fmaf(mu, mm[j], gj);
      ww[j] = fmaf(-lr, mm[j], ww[j]);
    }
    mp[0] = make_float4(mm[0], mm[1], mm[2], mm[3]);
    mp[1] = make_float4(mm[4], mm[5], mm[6], mm[7]);
    uint4 lo, hi, pb;
    lo.x = __float_as_uint(ww[0]); lo.y = __float_as_uint(ww[1]); lo.z = __float_as_uint(ww[2]); lo.w = __float_as_uint(ww[3]);
    hi.x = __float_as_uint(ww[4]); hi.y = __float_as_uint(ww[5]); hi.z = __float_as_uint(ww[6]); hi.w = __float_as_uint(ww[7]);
    pb.x = pdj_f2_to_bf2(ww[0], ww[1]); pb.y = pdj_f2_to_bf2(ww[2], ww[3]);
    pb.z = pdj_f2_to_bf2(ww[4], ww[5]); pb.w = pdj_f2_to_bf2(ww[6], ww[7]);
    if (NVLS) {
      char* mc = c.mc_base;
      asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc + master_off + g8 * 32), "r"(lo.x),
                   "r"(lo.y), "r"(lo.z), "r"(lo.w) : "memory");
      asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc + master_off + g8 * 32 + 16), "r"(hi.x),
                   "r"(hi.y), "r"(hi.z), "r"(hi.w) : "memory");
      asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc + param_off + g8 * 16), "r"(pb.x),
                   "r"(pb.y), "r"(pb.z), "r"(pb.w) : "memory");
    } else {
#pragma unroll
      for (int p = 0; p < NR; ++p) {
        char* dst = c.base[(rank + p) % NR];
        pdj_st16(dst + master_off + g8 * 32, lo);
        pdj_st16(dst + master_off + g8 * 32 + 16, hi);
        pdj_st16(dst + param_off + g8 * 16, pb);
      }
    }
  }
  pdj_block_barrier(c, rank, e + 2);  // every owner's pushes into my parameters have landed
  if (threadIdx.x == 0) me->epoch[pdj_slot(c)] = e + 2;
}

// ---------------------------------------------------------------------------
// K5 all-to-all: rank r's input is NR chunks of chunk_vec vectors; chunk p is
// pushed to peer p's output slot r.
__global__ void __launch_bounds__(512)
alltoall_kernel(PdjComm c, size_t in_off, size_t out_off, size_t chunk_vec) {
  const int rank = c.rank + blockIdx.y;
  const int NR = c.nranks;
  PdjSignalPad* me = pdj_pad(c, rank);
  const uint32_t e = me->epoch[pdj_slot(c)];
  pdj_block_barrier(c, rank, e + 1);  // peers' outputs are free to overwrite
  const char* in = c.base[rank] + in_off;
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  const size_t total = chunk_vec * NR;
  // Four 16-byte vectors per thread per trip, all loads issued before the first store: the local loads overlap and
  // four peer stores leave back to back (one vector per trip needed 148 blocks to cover the NVLink latency).
  // Peers are walked in a rotated order so that at any instant the N ranks target N different destinations.
  constexpr int U = 4;
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i + (U - 1) * stride < total; i += U * stride) {
    uint4 v[U];
    char* dst[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t j = i + u * stride;
      const size_t within = j % chunk_vec;
      const int p = (int)((j / chunk_vec + rank) % NR);
      v[u] = pdj_ld16(in + ((size_t)p * chunk_vec + within) * 16);
      dst[u] = c.base[p] + out_off + ((size_t)rank * chunk_vec + within) * 16;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) pdj_st16(dst[u], v[u]);
  }
  for (; i < total; i += stride) {
    const size_t within = i % chunk_vec;
    const int p = (int)((i / chunk_vec + rank) % NR);
    const uint4 v = pdj_ld16(in + ((size_t)p * chunk_vec + within) * 16);
    pdj_st16(c.base[p] + out_off + ((size_t)rank * chunk_vec + within) * 16, v);
  }
  pdj_block_barrier(c, rank, e + 2);
  if (threadIdx.x == 0) me->epoch[pdj_slot(c)] = e + 2;
}

// K6 broadcast: every non-root rank pulls the root's buffer (pull keeps the
// root's SMs free and uses each puller's own ingress link).
__global__ void __launch_bounds__(512)
broadcast_kernel(PdjComm c, int root, size_t off, size_t n_vec) {
  const int rank = c.rank + blockIdx.y;
  PdjSignalPad* me = pdj_pad(c, rank);
  const uint32_t e = me->epoch[pdj_slot(c)];
  pdj_block_barrier(c, rank, e + 1);
  if (rank != root) {
    const char* src = c.base[root] + off;
    char* dst = c.base[rank] + off;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 3 * stride < n_vec; i += 4 * stride) {  // four peer loads in flight per thread
      const uint4 a = pdj_ld16(src + i * 16), b = pdj_ld16(src + (i + stride) * 16);
      const uint4 cc = pdj_ld16(src + (i + 2 * stride) * 16), d = pdj_ld16(src + (i + 3 * stride) * 16);
      pdj_st16(dst + i * 16, a);
      pdj_st16(dst + (i + stride) * 16, b);
      pdj_st16(dst + (i + 2 * stride) * 16, cc);
      pdj_st16(dst + (i + 3 * stride) * 16, d);
    }
    for (; i < n_vec; i += stride) pdj_st16(dst + i * 16, pdj_ld16(src + i * 16));
  }
  pdj_block_barrier(c, rank, e + 2);
  if (threadIdx.x == 0) me->epoch[pdj_slot(c)] = e + 2;
}

template <typename... Args>
int launch_comm(void (*kern)(Args...), int nblocks, int nthreads, int emulate, int nranks, cudaStream_t st,
                Args... args) {
  if (nblocks < 1) nblocks = 1;
  if (nblocks > PDJ_CHANNEL_BLOCKS) nblocks = PDJ_CHANNEL_BLOCKS;
  dim3 grid(nblocks, emulate ? nranks : 1);
  if (emulate) {
    // All y-slices spin on one another: they must be co-resident.
    void* argv[] = {(void*)&args...};
    PDJ_CUDA_CHECK(cudaLaunchCooperativeKernel((const void*)kern, grid, dim3(nthreads), argv, 0, st));
  } else {
    kern<<<grid, nthreads, 0, st>>>(args...);
    PDJ_CUDA_CHECK(cudaGetLastError());
  }
  return 0;
}

}  // namespace

extern "C" {

int pdj_barrier(const PdjComm* c, int emulate, cudaStream_t st) {
  return launch_comm(barrier_kernel, 1, 32, emulate, c->nranks, st, *c);
}

int pdj_allreduce_bf16(const PdjComm* c, size_t in_off, size_t out_off, size_t n_elems, float scale, int algo,
                       int out_f32, int nblocks, int emulate, cudaStream_t st) {
  if (n_elems % 8 != 0 || in_off % 16 != 0 || out_off % 16 != 0) {
    pdj_set_error("pdj_allreduce_bf16: n_elems must be a multiple of 8 and offsets 16-byte aligned");
    return -1;
  }
  const size_t n_vec = n_elems / 8;
  const int nr = c->nranks;
  if (algo == PDJ_AR_NVLS) {
    if (c->mc_base == nullptr || out_f32 || emulate) {
      pdj_set_error("pdj_allreduce_bf16: NVLS needs a multicast mapping, bf16 output and real ranks");
      return -1;
    }
    return launch_comm(allreduce_nvls_kernel, nblocks, 512, 0, nr, st, *c, in_off, out_off, n_vec, scale);
  }
  if (algo == PDJ_AR_ONESHOT) {
    if (out_f32 || in_off == out_off) {
      pdj_set_error("pdj_allreduce_bf16: one-shot needs out != in and bf16 output");
      return -1;
    }
#define ONE(NR) \
  case NR: return launch_comm(allreduce_oneshot_kernel<NR>, nblocks, 512, emulate, nr, st, *c, in_off, out_off, n_vec, scale)
    switch (nr) { ONE(1); ONE(2); ONE(3); ONE(4); ONE(5); ONE(6); ONE(7); ONE(8); }
#undef ONE
  } else {
#define TWO(NR)                                                                                                   \
  case NR:                                                                                                        \
    return out_f32 ? launch_comm(allreduce_twoshot_kernel<NR, true>, nblocks, 512, emulate, nr, st, *c, in_off,  \
                                 out_off, n_vec, scale)                                                           \
                   : launch_comm(allreduce_twoshot_kernel<NR, false>, nblocks, 512, emulate, nr, st, *c, in_off, \
                                 out_off, n_vec, scale)
    switch (nr) { TWO(1); TWO(2); TWO(3); TWO(4); TWO(5); TWO(6); TWO(7); TWO(8); }
#undef TWO
  }
  pdj_set_error("pdj_allreduce_bf16: unsupported nranks %d", nr);
  return -1;
}

int pdj_allreduce_sgd(const PdjComm* c, size_t grad_off, size_t param_off, size_t master_off, size_t mom_off,
                      size_t first_elem, size_t n_elems, size_t wd_split, const float* hyper, float scale, int algo,
                      int nblocks, int emulate, cudaStream_t st) {
  if (n_elems % 8 != 0 || first_elem % 8 != 0 || wd_split % 8 != 0 || grad_off % 16 != 0 || param_off % 16 != 0 ||
      master_off % 16 != 0 || mom_off % 16 != 0 || hyper == nullptr) {
    pdj_set_error("pdj_allreduce_sgd: element counts must be multiples of 8, offsets 16-byte aligned, hyper non-NULL");
    return -1;
  }
  const size_t vec0 = first_elem / 8, n_vec = n_elems / 8, wdv = wd_split / 8;
  const int nr = c->nranks;
  if (algo == PDJ_AR_NVLS) {
    if (c->mc_base == nullptr || emulate) {
      pdj_set_error("pdj_allreduce_sgd: NVLS needs a multicast mapping and real ranks");
      return -1;
    }
#define NV(NR) \
  case NR: return launch_comm(allreduce_sgd_kernel<NR, true>, nblocks, 512, 0, nr, st, *c, grad_off, param_off, master_off, mom_off, vec0, n_vec, wdv, hyper, scale)
    switch (nr) { NV(2); NV(3); NV(4); NV(5); NV(6); NV(7); NV(8); }
#undef NV
  } else {
#define P2P(NR) \
  case NR: return launch_comm(allreduce_sgd_kernel<NR, false>, nblocks, 512, emulate, nr, st, *c, grad_off, param_off, master_off, mom_off, vec0, n_vec, wdv, hyper, scale)
    switch (nr) { P2P(1); P2P(2); P2P(3); P2P(4); P2P(5); P2P(6); P2P(7); P2P(8); }
#undef P2P
  }
  pdj_set_error("pdj_allreduce_sgd: unsupported nranks %d", nr);
  return -1;
}

int pdj_alltoall(const PdjComm* c, size_t in_off, size_t out_off, size_t chunk_bytes, int nblocks, int emulate,
                 cudaStream_t st) {
  if (chunk_bytes % 16 != 0 || in_off % 16 != 0 || out_off % 16 != 0) {
    pdj_set_error("pdj_alltoall: chunk_bytes and offsets must be multiples of 16");
    return -1;
  }
  return launch_comm(alltoall_kernel, nblocks, 512, emulate, c->nranks, st, *c, in_off, out_off, chunk_bytes / 16);
}

int pdj_broadcast(const PdjComm* c, int root, size_t off, size_t bytes, int nblocks, int emulate, cudaStream_t st) {
  if (bytes % 16 != 0 || off % 16 != 0) {
    pdj_set_error("pdj_broadcast: bytes and offset must be multiples of 16");
    return -1;
  }
  return launch_comm(broadcast_kernel, nblocks, 512, emulate, c->nranks, st, *c, root, off, bytes / 16);
}

}  // extern "C"
