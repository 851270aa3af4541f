// K2 / K3: GPU-sharded embedding tables - the parameter-server role of PS mode
// with the all-to-all fused into the lookup (pull) and into the sparse
// optimizer (push).
//
// Reference parity: PS mode in the reference is `spec.ps` + `spec.worker`
// (controllers/paddlejob_helper.go:191-193): it creates PSERVER pods and TRAINER
// pods and publishes PADDLE_PSERVERS_IP_PORT_LIST (helper.go:258-260); what the
// roles do is described in docs/design-arch.md:7-9,49-58 - trainers pull
// parameters and push gradients, pservers "run optimize functions to update
// their parameters" - and executed by Paddle's brpc parameter server.  Here a
// pserver shard is a slice of the table resident in one GPU's arena:
//   row `id` lives on shard  id % nranks  at local row  id / nranks.
//   pull  = every trainer reads the owner's HBM directly (peer ld over NVSwitch);
//   push  = every trainer adds its row gradients into the owner's staging rows
//           with peer `red.add` and registers the row in the owner's touched
//           list; after a device barrier the owner applies SGD / Adagrad to the
//           touched rows (gradients of repeated ids, from any trainer, are summed
//           first: synchronous-SGD semantics, docs/design-arch.md:62-66);
//   async = trainers apply `w -= lr * g` on the owner immediately with peer
//           red.add and no barrier (asynchronous SGD, docs/design-arch.md:68-74).
//
// Everything is addressed by arena offsets so the same arguments are valid on
// every rank; the single-GPU emulation runs all ranks in one cooperative launch.
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "../../include/pdjob_b200.h"

namespace {

__device__ __forceinline__ float4 ld_f4(const void* p) {
  float4 v;
  asm volatile("ld.global.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ float ld_f1(const void* p) {
  float v;
  asm volatile("ld.global.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_add_f4(void* p, float a, float b, float c, float d) {
  asm volatile("red.relaxed.sys.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d)
               : "memory");
}
__device__ __forceinline__ void red_add_f1(void* p, float a) {
  asm volatile("red.relaxed.sys.global.add.f32 [%0], %1;" ::"l"(p), "f"(a) : "memory");
}
__device__ __forceinline__ uint32_t atom_exch_sys(uint32_t* p, uint32_t v) {
  uint32_t old;
  asm volatile("atom.relaxed.sys.global.exch.b32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
  return old;
}
__device__ __forceinline__ uint32_t atom_add_sys(uint32_t* p, uint32_t v) {
  uint32_t old;
  asm volatile("atom.relaxed.sys.global.add.u32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
  return old;
}

// Byte distance between consecutive rows of the deep arrays / of the wide arrays.  t.list_off = 0: separate dense
// arrays (dim * 4 and 4 bytes).  t.list_off = S > 0: ONE record of S bytes per row holds the row's weights, Adagrad
// accumulators and gradient staging (deep and wide of each next to one another; the *_off fields point at the
// parts of record 0), so a lookup touches one 128-byte line per row instead of two, and the apply pass reads and
// writes whole aligned records (256 B for dim 16) instead of six scattered sectors.
__device__ __forceinline__ size_t emb_sd(const PdjEmbTable& t) { return t.list_off ? t.list_off : (size_t)t.dim * 4; }
__device__ __forceinline__ size_t emb_sw(const PdjEmbTable& t) { return t.list_off ? t.list_off : (size_t)4; }

// Where row `id` lives.  The table is cut into P = t.list_cap shards (0 = one per rank): the PSERVER count of the
// job (spec.ps.replicas, controllers/paddlejob_helper.go:225-235,258-260).  Shard j = id % P sits on GPU j % NR in
// slot j / NR of that GPU's table, at row id / P of the slot; a slot is t.rows_per_shard rows.
__device__ __forceinline__ void emb_locate(const PdjEmbTable& t, int NR, long long id, int* owner, long long* row) {
  const int P = t.list_cap ? (int)t.list_cap : NR;
  if (((unsigned long long)id >> 32) == 0ull) {  // every table that fits a GPU: 32-bit division instead of the 64-bit routine
    const uint32_t u = (uint32_t)id, q = u / (uint32_t)P, shard = u - q * (uint32_t)P;
    const uint32_t slot = shard / (uint32_t)NR;
    *owner = (int)(shard - slot * (uint32_t)NR);
    *row = (long long)slot * t.rows_per_shard + q;
    return;
  }
  const int shard = (int)(id % P);
  *owner = shard % NR;
  *row = (long long)(shard / NR) * t.rows_per_shard + id / P;
}

__device__ __forceinline__ long long ld_id(const long long* p) {
  long long v;
  asm volatile("ld.global.nc.L1::no_allocate.b64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint2 ld_u2(const void* p) {
  uint2 v;
  asm volatile("ld.global.nc.L1::no_allocate.v2.b32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p) : "memory");
  return v;
}

// The columns of an output row past the S * D embedding values: the sample's dense features (fp32 -> bf16) and
// zeros up to the row pitch.  With out_pitch = the padded input width of the MLP the lookup writes the tower's input
// in place - no concatenation, no conversion kernel, no zero fill.
__device__ __forceinline__ void emb_write_tail(const PdjComm& c, int rank, __nv_bfloat16* out_deep, int b, int used,
                                               int out_pitch, size_t dense_off, int dense_n, int lane) {
  if (out_pitch <= used) return;
  const float* dense = dense_n > 0 ? reinterpret_cast<const float*>(c.base[rank] + dense_off) + (size_t)b * dense_n : nullptr;
  for (int k = used + lane; k < out_pitch; k += 32) {
    const int j = k - used;
    out_deep[(size_t)b * out_pitch + k] = __float2bfloat16(j < dense_n ? dense[j] : 0.f);
  }
}

// One warp per sample: S slots, D floats per row, D/4 lanes per row.
__global__ void __launch_bounds__(1024)
emb_lookup_kernel(PdjComm c, PdjEmbTable t, size_t ids_off, size_t out_deep_off, size_t out_wide_off, int B, int S,
                  int do_barrier, int out_pitch, size_t dense_off, int dense_n) {
  const int rank = c.rank + blockIdx.y;
  const int NR = c.nranks;
  PdjSignalPad* me = pdj_pad(c, rank);
  const uint32_t e = me->epoch[pdj_slot(c)];
  if (do_barrier) pdj_block_barrier(c, rank, e + 1);  // owners finished applying the previous step
  const int D = t.dim;
  const int lpr = D >> 2;          // lanes per row
  const int rpw = 32 / lpr;        // rows per warp iteration
  const int lane = threadIdx.x & 31;
  const int sub = lane / lpr, part = lane % lpr;
  const long long* ids = reinterpret_cast<const long long*>(c.base[rank] + ids_off);
  __nv_bfloat16* out_deep = reinterpret_cast<__nv_bfloat16*>(c.base[rank] + out_deep_off);
  float* out_wide = reinterpret_cast<float*>(c.base[rank] + out_wide_off);
  const int warps = (gridDim.x * blockDim.x) >> 5;
  for (int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; b < B; b += warps) {
    float wide = 0.f;
    for (int s0 = 0; s0 < S; s0 += rpw) {
      const int s = s0 + sub;
      if (s < S) {
        const long long id = ids[(size_t)b * S + s];
        int owner;
        long long row;
        emb_locate(t, NR, id, &owner, &row);
        const float4 v = ld_f4(c.base[owner] + t.deep_off + (size_t)row * emb_sd(t) + part * 16);
        uint2 o;
        o.x = pdj_f2_to_bf2(v.x, v.y);
        o.y = pdj_f2_to_bf2(v.z, v.w);
        *reinterpret_cast<uint2*>(out_deep + (size_t)b * out_pitch + (size_t)s * D + part * 4) = o;
        if (part == 0 && t.has_wide) wide += ld_f1(c.base[owner] + t.wide_off + (size_t)row * emb_sw(t));
      }
    }
    emb_write_tail(c, rank, out_deep, b, S * D, out_pitch, dense_off, dense_n, lane);
    if (t.has_wide) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) wide += __shfl_xor_sync(0xffffffffu, wide, o);
      if (lane == 0) out_wide[b] = wide;
    }
  }
  if (do_barrier && threadIdx.x == 0) me->epoch[pdj_slot(c)] = e + 1;
}

// Batched lookup (the default for up to eight passes per sample).  A random-row gather is bound by memory-level
// parallelism, not bytes, so every lane keeps PASSES row loads (16 B each, D/4 lanes per row) in flight at once -
// ids first, all addresses, all loads, then the conversions - and the ids of the warp's NEXT sample are requested
// before the current rows are touched, which takes the id load off the critical path (one DRAM / NVLink round trip
// per sample instead of 2 x passes).  The loop above issues one dependent load at a time.
template <int PASSES>
__global__ void __launch_bounds__(PASSES <= 2 ? 1024 : 768)
emb_lookup_batched_kernel(PdjComm c, PdjEmbTable t, size_t ids_off, size_t out_deep_off, size_t out_wide_off, int B, int S,
                          int do_barrier, int out_pitch, size_t dense_off, int dense_n) {
  const int rank = c.rank + blockIdx.y;
  const int NR = c.nranks;
  PdjSignalPad* me = pdj_pad(c, rank);
  const uint32_t e = me->epoch[pdj_slot(c)];
  if (do_barrier) pdj_block_barrier(c, rank, e + 1);  // owners finished applying the previous step
  const int D = t.dim;
  const int lpr = D >> 2, rpw = 32 / lpr;
  const int lane = threadIdx.x & 31;
  const int sub = lane / lpr, part = lane % lpr;
  const long long* ids = reinterpret_cast<const long long*>(c.base[rank] + ids_off);
  __nv_bfloat16* out_deep = reinterpret_cast<__nv_bfloat16*>(c.base[rank] + out_deep_off);
  float* out_wide = reinterpret_cast<float*>(c.base[rank] + out_wide_off);
  const int warps = (gridDim.x * blockDim.x) >> 5;
  const size_t sd = emb_sd(t), sw = emb_sw(t);
  const bool wide_lane = part == 0 && t.has_wide;
  long long idn[PASSES];
  auto load_ids = [&](int bb) {
#pragma unroll
    for (int p = 0; p < PASSES; ++p) {
      const int s_ = p * rpw + sub;
      idn[p] = (bb < B && s_ < S) ? ld_id(ids + (size_t)bb * S + s_) : -1ll;
    }
  };
  int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  load_ids(b);
  for (; b < B; b += warps) {
    int owner[PASSES];
    long long row[PASSES];
    bool ok[PASSES];
#pragma unroll
    for (int p = 0; p < PASSES; ++p) {
      ok[p] = idn[p] >= 0;
      owner[p] = 0;
      row[p] = 0;
      if (ok[p]) emb_locate(t, NR, idn[p], &owner[p], &row[p]);
    }
    load_ids(b + warps);
    float4 v[PASSES];
    float wv[PASSES];
#pragma unroll
    for (int p = 0; p < PASSES; ++p) {
      wv[p] = 0.f;
      if (ok[p]) {
        const char* ob = c.base[owner[p]];
        v[p] = ld_f4(ob + t.deep_off + (size_t)row[p] * sd + part * 16);
        if (wide_lane) wv[p] = ld_f1(ob + t.wide_off + (size_t)row[p] * sw);
      }
    }
    float wide = 0.f;
#pragma unroll
    for (int p = 0; p < PASSES; ++p) {
      if (ok[p]) {
        uint2 o;
        o.x = pdj_f2_to_bf2(v[p].x, v[p].y);
        o.y = pdj_f2_to_bf2(v[p].z, v[p].w);
        *reinterpret_cast<uint2*>(out_deep + (size_t)b * out_pitch + (size_t)(p * rpw + sub) * D + part * 4) = o;
        wide += wv[p];
      }
    }
    emb_write_tail(c, rank, out_deep, b, S * D, out_pitch, dense_off, dense_n, lane);
    if (t.has_wide) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) wide += __shfl_xor_sync(0xffffffffu, wide, o);
      if (lane == 0) out_wide[b] = wide;
    }
  }
  if (do_barrier && threadIdx.x == 0) me->epoch[pdj_slot(c)] = e + 1;
}

// Lookup through the bf16 read view at the head of each record (PdjEmbTable.shadow_off): one 64-byte request per
// row - [bf16 deep, 2 * dim bytes][wide weight] - instead of a 64-byte and a 32-byte one, nothing to convert, and
// a third fewer bytes over NVLink (the pull is packet-bound there: 330 - 380 GB/s effective at 2 - 4 GPUs).  Same
// batching as emb_lookup_batched_kernel.  LPR lanes per row, 16 bytes each: lanes < dim / 8 carry the deep values,
// lane dim / 8 carries the wide weight in its first word, the rest (if any) stay idle.
template <int PASSES>
__global__ void __launch_bounds__(PASSES <= 2 ? 1024 : 768)
emb_lookup_shadow_kernel(PdjComm c, PdjEmbTable t, size_t ids_off, size_t out_deep_off, size_t out_wide_off, int B, int S,
                         int do_barrier, int out_pitch, size_t dense_off, int dense_n, int lpr) {
  const int rank = c.rank + blockIdx.y;
  const int NR = c.nranks;
  PdjSignalPad* me = pdj_pad(c, rank);
  const uint32_t e = me->epoch[pdj_slot(c)];
  if (do_barrier) pdj_block_barrier(c, rank, e + 1);  // owners finished applying the previous step
  const int D = t.dim;
  const int dchunks = D >> 3;  // 16-byte chunks of bf16 deep values per row
  const int rpw = 32 / lpr;
  const int lane = threadIdx.x & 31;
  const int sub = lane / lpr, part = lane % lpr;
  const long long* ids = reinterpret_cast<const long long*>(c.base[rank] + ids_off);
  __nv_bfloat16* out_deep = reinterpret_cast<__nv_bfloat16*>(c.base[rank] + out_deep_off);
  float* out_wide = reinterpret_cast<float*>(c.base[rank] + out_wide_off);
  const int warps = (gridDim.x * blockDim.x) >> 5;
  const size_t sd = emb_sd(t);
  const bool active = part < dchunks || (part == dchunks && t.has_wide);
  long long idn[PASSES];
  auto load_ids = [&](int bb) {
#pragma unroll
    for (int p = 0; p < PASSES; ++p) {
      const int s_ = p * rpw + sub;
      idn[p] = (bb < B && s_ < S && active) ? ld_id(ids + (size_t)bb * S + s_) : -1ll;
    }
  };
  int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  load_ids(b);
  for (; b < B; b += warps) {
    int owner[PASSES];
    long long row[PASSES];
    bool ok[PASSES];
#pragma unroll
    for (int p = 0; p < PASSES; ++p) {
      ok[p] = idn[p] >= 0;
      owner[p] = 0;
      row[p] = 0;
      if (ok[p]) emb_locate(t, NR, idn[p], &owner[p], &row[p]);
    }
    load_ids(b + warps);
    uint4 v[PASSES];
#pragma unroll
    for (int p = 0; p < PASSES; ++p) {
      v[p] = make_uint4(0u, 0u, 0u, 0u);
      if (ok[p]) {
        const float4 f = ld_f4(c.base[owner[p]] + t.shadow_off + (size_t)row[p] * sd + part * 16);
        v[p] = make_uint4(__float_as_uint(f.x), __float_as_uint(f.y), __float_as_uint(f.z), __float_as_uint(f.w));
      }
    }
    float wide = 0.f;
#pragma unroll
    for (int p = 0; p < PASSES; ++p) {
      if (!ok[p]) continue;
      if (part < dchunks)
        *reinterpret_cast<uint4*>(out_deep + (size_t)b * out_pitch + (size_t)(p * rpw + sub) * D + part * 8) = v[p];
      else
        wide += __uint_as_float(v[p].x);
    }
    emb_write_tail(c, rank, out_deep, b, S * D, out_pitch, dense_off, dense_n, lane);
    if (t.has_wide) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) wide += __shfl_xor_sync(0xffffffffu, wide, o);
      if (lane == 0) out_wide[b] = wide;
    }
  }
  if (do_barrier && threadIdx.x == 0) me->epoch[pdj_slot(c)] = e + 1;
}

// TMA-fed lookup.  One warp per sample, double buffered: lane s asks the TMA engine for row
// id[b, s] with one bulk copy (cp.async.bulk, 16-byte granularity, completion counted on an
// mbarrier) straight from the owner GPU's HBM into this warp's shared-memory tile, for sample
// b + stride while sample b is being converted.  26 x 64 B rows per sample and two samples per
// warp are in flight without holding a single register, which is what a latency-bound gather
// over NVLink (~2 us per access) needs.  SASS: UBLKCP + SYNCS.
__device__ __forceinline__ uint32_t emb_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(1024)
emb_lookup_tma_kernel(PdjComm c, PdjEmbTable t, size_t ids_off, size_t out_deep_off, size_t out_wide_off, int B, int S,
                      int do_barrier, int out_pitch, size_t dense_off, int dense_n) {
  extern __shared__ __align__(128) uint8_t emb_sm[];
  __shared__ __align__(8) uint64_t bars[32][2];  // one pair per warp, up to 1024 threads
  const int rank = c.rank + blockIdx.y;
  const int NR = c.nranks;
  PdjSignalPad* me = pdj_pad(c, rank);
  const uint32_t e = me->epoch[pdj_slot(c)];
  const int D = t.dim;
  const uint32_t row_bytes = (uint32_t)D * 4u;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const uint32_t tile = emb_smem_u32(emb_sm) + (uint32_t)w * 2u * 32u * row_bytes;  // [2][32][row_bytes]
  const uint32_t bar0 = emb_smem_u32(&bars[w][0]), bar1 = emb_smem_u32(&bars[w][1]);
  if (lane == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0));
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar1));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
  if (do_barrier) pdj_block_barrier(c, rank, e + 1);  // owners finished applying the previous step
  const long long* ids = reinterpret_cast<const long long*>(c.base[rank] + ids_off);
  __nv_bfloat16* out_deep = reinterpret_cast<__nv_bfloat16*>(c.base[rank] + out_deep_off);
  float* out_wide = reinterpret_cast<float*>(c.base[rank] + out_wide_off);
  const int wpb = (int)(blockDim.x >> 5);
  const int warps = gridDim.x * wpb;

  auto issue = [&](int b, int buf) -> float {
    const uint32_t bar = buf ? bar1 : bar0;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // earlier generic reads of this buffer are done
    if (lane == 0)
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((uint32_t)S * row_bytes) : "memory");
    __syncwarp();
    float wv = 0.f;
    if (lane < S) {
      const long long id = ids[(size_t)b * S + lane];
      int owner;
      long long row;
      emb_locate(t, NR, id, &owner, &row);
      const char* src = c.base[owner] + t.deep_off + (size_t)row * emb_sd(t);
      const uint32_t dst = tile + ((uint32_t)buf * 32u + (uint32_t)lane) * row_bytes;
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                   "l"(src), "r"(row_bytes), "r"(bar)
                   : "memory");
      if (t.has_wide) wv = ld_f1(c.base[owner] + t.wide_off + (size_t)row * emb_sw(t));
    }
    return wv;
  };

  uint32_t phase0 = 0, phase1 = 0;
  int b = blockIdx.x * wpb + w;
  int buf = 0;
  float wv_cur = 0.f;
  if (b < B) wv_cur = issue(b, 0);
  for (; b < B; b += warps) {
    const int nb = b + warps;
    float wv_next = 0.f;
    if (nb < B) wv_next = issue(nb, buf ^ 1);
    {
      const uint32_t bar = buf ? bar1 : bar0;
      const uint32_t ph = buf ? phase1 : phase0;
      asm volatile(
          "{\n\t"
          ".reg .pred p;\n\t"
          "EW_%=:\n\t"
          "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
          "@p bra ED_%=;\n\t"
          "bra EW_%=;\n\t"
          "ED_%=:\n\t"
          "}" ::"r"(bar),
          "r"(ph)
          : "memory");
      if (buf) phase1 ^= 1u; else phase0 ^= 1u;
    }
    if (lane < S) {
      const uint32_t src = tile + ((uint32_t)buf * 32u + (uint32_t)lane) * row_bytes;
      __nv_bfloat16* o = out_deep + (size_t)b * out_pitch + (size_t)lane * D;
      for (int k = 0; k < D; k += 8) {
        float4 v0, v1;
        asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v0.x), "=f"(v0.y), "=f"(v0.z), "=f"(v0.w) : "r"(src + k * 4));
        asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v1.x), "=f"(v1.y), "=f"(v1.z), "=f"(v1.w) : "r"(src + k * 4 + 16));
        uint4 p;
        p.x = pdj_f2_to_bf2(v0.x, v0.y); p.y = pdj_f2_to_bf2(v0.z, v0.w);
        p.z = pdj_f2_to_bf2(v1.x, v1.y); p.w = pdj_f2_to_bf2(v1.z, v1.w);
        *reinterpret_cast<uint4*>(o + k) = p;
      }
    }
    if (t.has_wide) {
      float wide = wv_cur;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) wide += __shfl_xor_sync(0xffffffffu, wide, o);
      if (lane == 0) out_wide[b] = wide;
    }
    emb_write_tail(c, rank, out_deep, b, S * D, out_pitch, dense_off, dense_n, lane);
    __syncwarp();
    wv_cur = wv_next;
    buf ^= 1;
  }
  if (do_barrier && threadIdx.x == 0) me->epoch[pdj_slot(c)] = e + 1;
}

// Push.  mode 0 (sync): accumulate into the owner's staging rows and register
// the row; mode 1 (async SGD): apply w -= lr * g on the owner directly.
__global__ void __launch_bounds__(1024)
emb_push_kernel(PdjComm c, PdjEmbTable t, size_t ids_off, size_t gdeep_off, size_t gwide_off, int B, int S, int mode,
                float lr, float grad_scale, int g_pitch) {
  const int rank = c.rank + blockIdx.y;
  const int NR = c.nranks;
  PdjSignalPad* me = pdj_pad(c, rank);
  const uint32_t e = me->epoch[pdj_slot(c)];
  const int D = t.dim;
  const int lpr = D >> 2, rpw = 32 / lpr;
  const int lane = threadIdx.x & 31;
  const int sub = lane / lpr, part = lane % lpr;
  const long long* ids = reinterpret_cast<const long long*>(c.base[rank] + ids_off);
  const __nv_bfloat16* gdeep = reinterpret_cast<const __nv_bfloat16*>(c.base[rank] + gdeep_off);
  const float* gwide = reinterpret_cast<const float*>(c.base[rank] + gwide_off);
  const int warps = (gridDim.x * blockDim.x) >> 5;
  const float k = mode == 1 ? -lr * grad_scale : grad_scale;
  for (int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; b < B; b += warps) {
    const float gw = t.has_wide ? gwide[b] * k : 0.f;
    for (int s0 = 0; s0 < S; s0 += rpw) {
      const int s = s0 + sub;
      if (s >= S) continue;
      const long long id = ids[(size_t)b * S + s];
      int owner;
      long long row;
      emb_locate(t, NR, id, &owner, &row);
      const uint2 g = *reinterpret_cast<const uint2*>(gdeep + (size_t)b * g_pitch + (size_t)s * D + part * 4);
      const float2 g0 = pdj_bf2_to_f2(g.x), g1 = pdj_bf2_to_f2(g.y);
      char* ob = c.base[owner];
      const size_t deep_dst = (mode == 1 ? t.deep_off : t.gdeep_off) + (size_t)row * emb_sd(t) + part * 16;
      red_add_f4(ob + deep_dst, g0.x * k, g0.y * k, g1.x * k, g1.y * k);
      if (part == 0) {
        if (t.has_wide) red_add_f1(ob + (mode == 1 ? t.wide_off : t.gwide_off) + (size_t)row * emb_sw(t), gw);
        if (mode == 0) {
          // Mark the row as touched: an idempotent one-byte store.  (A first version claimed a slot in
          // a per-owner list with atom.exch + atom.add on ONE counter; 850 k returning atomics on a
          // single address cost 1.6 ms per step.  The owner now finds the rows by scanning the byte
          // map, which is 16 MB per 16 M rows - microseconds.)
          asm volatile("st.relaxed.sys.global.u8 [%0], %1;" ::"l"(ob + t.flag_off + (size_t)row), "r"(1u) : "memory");
        }
      }
    }
  }
  // All pushes of all ranks have landed once every block of every rank has
  // passed this barrier and the kernels have drained.
  pdj_block_barrier(c, rank, e + 1);
  if (threadIdx.x == 0) me->epoch[pdj_slot(c)] = e + 1;
}

// Batched push: the same memory-level-parallelism treatment as emb_lookup_batched_kernel.  The ids and gradient
// pieces of the warp's next sample are requested before the current sample's reductions are issued, so the only
// thing a warp ever waits for is the previous iteration's prefetch.
template <int PASSES>
__global__ void __launch_bounds__(PASSES <= 2 ? 1024 : 768)
emb_push_batched_kernel(PdjComm c, PdjEmbTable t, size_t ids_off, size_t gdeep_off, size_t gwide_off, int B, int S, int mode,
                        float lr, float grad_scale, int g_pitch) {
  const int rank = c.rank + blockIdx.y;
  const int NR = c.nranks;
  PdjSignalPad* me = pdj_pad(c, rank);
  const uint32_t e = me->epoch[pdj_slot(c)];
  const int D = t.dim;
  const int lpr = D >> 2, rpw = 32 / lpr;
  const int lane = threadIdx.x & 31;
  const int sub = lane / lpr, part = lane % lpr;
  const long long* ids = reinterpret_cast<const long long*>(c.base[rank] + ids_off);
  const __nv_bfloat16* gdeep = reinterpret_cast<const __nv_bfloat16*>(c.base[rank] + gdeep_off);
  const float* gwide = reinterpret_cast<const float*>(c.base[rank] + gwide_off);
  const int warps = (gridDim.x * blockDim.x) >> 5;
  const float k = mode == 1 ? -lr * grad_scale : grad_scale;
  const size_t sd = emb_sd(t), sw = emb_sw(t);
  const size_t deep_base = mode == 1 ? t.deep_off : t.gdeep_off, wide_base = mode == 1 ? t.wide_off : t.gwide_off;
  long long idn[PASSES];
  uint2 gn[PASSES];
  float gwn = 0.f;
  auto load_next = [&](int bb) {
#pragma unroll
    for (int p = 0; p < PASSES; ++p) {
      const int s_ = p * rpw + sub;
      const bool valid = bb < B && s_ < S;
      idn[p] = valid ? ld_id(ids + (size_t)bb * S + s_) : -1ll;
      gn[p] = valid ? ld_u2(gdeep + (size_t)bb * g_pitch + (size_t)s_ * D + part * 4) : make_uint2(0u, 0u);
    }
    gwn = (bb < B && t.has_wide) ? __ldg(gwide + bb) : 0.f;
  };
  int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  load_next(b);
  for (; b < B; b += warps) {
    int owner[PASSES];
    long long row[PASSES];
    uint2 g[PASSES];
    bool ok[PASSES];
    const float gw = gwn * k;
#pragma unroll
    for (int p = 0; p < PASSES; ++p) {
      ok[p] = idn[p] >= 0;
      g[p] = gn[p];
      owner[p] = 0;
      row[p] = 0;
      if (ok[p]) emb_locate(t, NR, idn[p], &owner[p], &row[p]);
    }
    load_next(b + warps);
#pragma unroll
    for (int p = 0; p < PASSES; ++p) {
      if (!ok[p]) continue;
      const float2 g0 = pdj_bf2_to_f2(g[p].x), g1 = pdj_bf2_to_f2(g[p].y);
      char* ob = c.base[owner[p]];
      red_add_f4(ob + deep_base + (size_t)row[p] * sd + part * 16, g0.x * k, g0.y * k, g1.x * k, g1.y * k);
      if (part == 0) {
        if (t.has_wide) red_add_f1(ob + wide_base + (size_t)row[p] * sw, gw);
        if (mode == 0)  // mark the row as touched: an idempotent one-byte store (see emb_push_kernel)
          asm volatile("st.relaxed.sys.global.u8 [%0], %1;" ::"l"(ob + t.flag_off + (size_t)row[p]), "r"(1u) : "memory");
      }
    }
  }
  pdj_block_barrier(c, rank, e + 1);
  if (threadIdx.x == 0) me->epoch[pdj_slot(c)] = e + 1;
}

// Owner-side optimizer over the touched rows.  opt 0 = SGD, 1 = Adagrad.
// Each warp scans windows of 512 byte-flags (one 16-byte load per lane), compacts the set rows into a
// shared-memory queue with a warp prefix sum, clears the flags, and then updates the queued rows with D/4 lanes
// per row (float4 each), four rows per lane at a time: all of their loads (gradient, weight, accumulator, and the
// wide scalars) are issued before the first is used, so a warp pays two memory round trips per 512 rows.
constexpr int kApplyWin = 512, kApplyBatch = 4;
__global__ void __launch_bounds__(256)
emb_apply_kernel(PdjComm c, PdjEmbTable t, int opt, float lr, float eps) {
  __shared__ uint32_t queue[8][kApplyWin];
  const int rank = c.rank + blockIdx.y;
  char* base = c.base[rank];
  const int D = t.dim;
  const int lpr = D >> 2, rpw = 32 / lpr;
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int sub = lane / lpr, part = lane % lpr;
  // every slot this GPU holds: ceil(P / NR) of them (the last may be unused on some ranks; its flags stay clear)
  const int P_ = t.list_cap ? (int)t.list_cap : c.nranks;
  const long long nrows = (long long)((P_ + c.nranks - 1) / c.nranks) * t.rows_per_shard;
  const long long flag_limit = (nrows + 127) / 128 * 128;  // bytes the flag map holds
  const long long nwin = (nrows + kApplyWin - 1) / kApplyWin;
  const long long gw = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long nw = ((long long)gridDim.x * blockDim.x) >> 5;
  char* flags = base + t.flag_off;
  const size_t sd = emb_sd(t), sw = emb_sw(t);
  const bool wide_lane = part == 0 && t.has_wide;
  uint32_t touched = 0;
  for (long long w = gw; w < nwin; w += nw) {
    const long long r0 = w * kApplyWin + lane * 16;
    uint4 f = make_uint4(0u, 0u, 0u, 0u);
    if (r0 < flag_limit) f = *reinterpret_cast<const uint4*>(flags + r0);
    const uint32_t fw[4] = {f.x, f.y, f.z, f.w};
    uint32_t m = 0;
#pragma unroll
    for (int q = 0; q < 4; ++q)
      m |= (((fw[q] & 0xffu) ? 1u : 0u) | ((fw[q] & 0xff00u) ? 2u : 0u) | ((fw[q] & 0xff0000u) ? 4u : 0u) |
            ((fw[q] & 0xff000000u) ? 8u : 0u)) << (4 * q);
    if (__ballot_sync(0xffffffffu, m != 0u) == 0u) continue;
    const int cnt = __popc(m);
    int pre = cnt;  // inclusive prefix sum over the warp
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, pre, o);
      if (lane >= o) pre += v;
    }
    const int total = __shfl_sync(0xffffffffu, pre, 31);
    int pos = pre - cnt;
#pragma unroll
    for (int b = 0; b < 16; ++b)
      if ((m >> b) & 1u) queue[wib][pos++] = (uint32_t)(r0 + b);
    if (m != 0u) *reinterpret_cast<uint4*>(flags + r0) = make_uint4(0u, 0u, 0u, 0u);
    __syncwarp();
    touched += (uint32_t)total;
    for (int i0 = 0; i0 < total; i0 += rpw * kApplyBatch) {
      bool ok[kApplyBatch];
      size_t ro[kApplyBatch];
      float4 g[kApplyBatch], wv[kApplyBatch], a[kApplyBatch];
      float gs[kApplyBatch], ws[kApplyBatch], as[kApplyBatch];
#pragma unroll
      for (int k = 0; k < kApplyBatch; ++k) {
        const int idx = i0 + k * rpw + sub;
        ok[k] = idx < total;
        ro[k] = ok[k] ? (size_t)queue[wib][idx] : 0;
      }
#pragma unroll
      for (int k = 0; k < kApplyBatch; ++k) {
        if (!ok[k]) continue;
        const size_t o = ro[k] * sd + part * 16;
        g[k] = *reinterpret_cast<const float4*>(base + t.gdeep_off + o);
        wv[k] = *reinterpret_cast<const float4*>(base + t.deep_off + o);
        if (opt == 1) a[k] = *reinterpret_cast<const float4*>(base + t.acc_deep_off + o);
        if (wide_lane) {
          gs[k] = *reinterpret_cast<const float*>(base + t.gwide_off + ro[k] * sw);
          ws[k] = *reinterpret_cast<const float*>(base + t.wide_off + ro[k] * sw);
          if (opt == 1) as[k] = *reinterpret_cast<const float*>(base + t.acc_wide_off + ro[k] * sw);
        }
      }
#pragma unroll
      for (int k = 0; k < kApplyBatch; ++k) {
        if (!ok[k]) continue;
        const size_t o = ro[k] * sd + part * 16;
        float4 w4 = wv[k];
        const float4 g4 = g[k];
        if (opt == 1) {
          float4 a4 = a[k];
          a4.x += g4.x * g4.x; a4.y += g4.y * g4.y; a4.z += g4.z * g4.z; a4.w += g4.w * g4.w;
          *reinterpret_cast<float4*>(base + t.acc_deep_off + o) = a4;
          w4.x -= lr * g4.x * rsqrtf(a4.x + eps); w4.y -= lr * g4.y * rsqrtf(a4.y + eps);
          w4.z -= lr * g4.z * rsqrtf(a4.z + eps); w4.w -= lr * g4.w * rsqrtf(a4.w + eps);
        } else {
          w4.x -= lr * g4.x; w4.y -= lr * g4.y; w4.z -= lr * g4.z; w4.w -= lr * g4.w;
        }
        *reinterpret_cast<float4*>(base + t.deep_off + o) = w4;
        if (t.shadow_off)  // the bf16 view lookups read
          *reinterpret_cast<uint2*>(base + t.shadow_off + ro[k] * sd + part * 8) =
              make_uint2(pdj_f2_to_bf2(w4.x, w4.y), pdj_f2_to_bf2(w4.z, w4.w));
        *reinterpret_cast<float4*>(base + t.gdeep_off + o) = make_float4(0.f, 0.f, 0.f, 0.f);
        if (wide_lane) {
          float wn = ws[k];
          if (opt == 1) {
            const float an = as[k] + gs[k] * gs[k];
            *reinterpret_cast<float*>(base + t.acc_wide_off + ro[k] * sw) = an;
            wn -= lr * gs[k] * rsqrtf(an + eps);
          } else {
            wn -= lr * gs[k];
          }
          *reinterpret_cast<float*>(base + t.wide_off + ro[k] * sw) = wn;
          *reinterpret_cast<float*>(base + t.gwide_off + ro[k] * sw) = 0.f;
        }
      }
    }
    __syncwarp();
  }
  if (lane == 0 && touched != 0u) atomicAdd(reinterpret_cast<uint32_t*>(base + t.count_off), touched);
}

__global__ void emb_reset_count_kernel(PdjComm c, PdjEmbTable t) {
  const int rank = c.rank + blockIdx.y;
  if (threadIdx.x == 0) {
    uint32_t* cnt = reinterpret_cast<uint32_t*>(c.base[rank] + t.count_off);
    cnt[1] = cnt[0];  // keep the last step's touched-row count for inspection
    cnt[0] = 0u;
  }
}

int check_table(const PdjEmbTable* t) {
  const int d = t->dim;
  if (d < 4 || d > 128 || (d & (d - 1)) != 0) {
    pdj_set_error("embedding: dim must be a power of two in 4..128 (got %d)", d);
    return -1;
  }
  return 0;
}

template <typename... Args>
int launch_emb(void (*kern)(Args...), int nblocks, int nthreads, size_t smem, int emulate, int nranks, cudaStream_t st,
               Args... args) {
  if (nblocks < 1) nblocks = 1;
  if (nblocks > PDJ_CHANNEL_BLOCKS) nblocks = PDJ_CHANNEL_BLOCKS;
  dim3 grid(nblocks, emulate ? nranks : 1);
  if (emulate) {
    void* argv[] = {(void*)&args...};
    PDJ_CUDA_CHECK(cudaLaunchCooperativeKernel((const void*)kern, grid, dim3(nthreads), argv, smem, st));
  } else {
    kern<<<grid, nthreads, smem, st>>>(args...);
    PDJ_CUDA_CHECK(cudaGetLastError());
  }
  return 0;
}

}  // namespace

extern "C" {

int pdj_emb_lookup(const PdjComm* c, const PdjEmbTable* t, size_t ids_off, size_t out_deep_off, size_t out_wide_off,
                   int B, int S, int barrier, int nblocks, int emulate, cudaStream_t st) {
  return pdj_emb_lookup2(c, t, ids_off, out_deep_off, out_wide_off, B, S, barrier, S * t->dim, 0, 0, nblocks, emulate, st);
}

int pdj_emb_lookup2(const PdjComm* c, const PdjEmbTable* t, size_t ids_off, size_t out_deep_off, size_t out_wide_off,
                    int B, int S, int barrier, int out_pitch, size_t dense_off, int dense_n, int nblocks, int emulate,
                    cudaStream_t st) {
  if (check_table(t)) return -1;
  if (out_pitch < S * t->dim || out_pitch % 8 != 0 || dense_n < 0 || S * t->dim + dense_n > out_pitch || dense_off % 4 != 0) {
    pdj_set_error("pdj_emb_lookup2: out_pitch (%d) must be a multiple of 8 and hold S * dim + dense_n = %d values", out_pitch,
                  S * t->dim + dense_n);
    return -1;
  }
  // PDJ_EMB_LOOKUP = batched (default) | tma | loop
  const char* how = getenv("PDJ_EMB_LOOKUP");
  const bool no_tma = how == nullptr || strcmp(how, "tma") != 0;
  if ((how == nullptr || strcmp(how, "batched") == 0) && t->shadow_off != 0 && t->list_off != 0 && t->dim >= 8 && t->dim <= 64) {
    // read view: dim / 8 chunks of bf16 values + one chunk holding the wide weight, rounded up to a power of two lanes
    int lpr = 1;
    while (lpr < (t->dim >> 3) + (t->has_wide ? 1 : 0)) lpr <<= 1;
    const int sp = (S + 32 / lpr - 1) / (32 / lpr);
    if (sp <= 8 && t->wide_off == t->shadow_off + (size_t)t->dim * 2) {
      const int bt = emulate ? 256 : (sp <= 2 ? 1024 : 768);
#define PDJ_LS(P) launch_emb(emb_lookup_shadow_kernel<P>, nblocks, bt, (size_t)0, emulate, c->nranks, st, *c, *t, ids_off, \
                             out_deep_off, out_wide_off, B, S, barrier, out_pitch, dense_off, dense_n, lpr)
      return sp <= 1 ? PDJ_LS(1) : sp <= 2 ? PDJ_LS(2) : sp <= 4 ? PDJ_LS(4) : PDJ_LS(8);
#undef PDJ_LS
    }
  }
  const int passes = (S + (32 / (t->dim >> 2)) - 1) / (32 / (t->dim >> 2));
  if ((how == nullptr || strcmp(how, "batched") == 0) && passes <= 8) {
    int bt = emulate ? 256 : (passes <= 2 ? 1024 : 768);
#define PDJ_LK(P) launch_emb(emb_lookup_batched_kernel<P>, nblocks, bt, (size_t)0, emulate, c->nranks, st, *c, *t, ids_off, \
                             out_deep_off, out_wide_off, B, S, barrier, out_pitch, dense_off, dense_n)
    return passes <= 1 ? PDJ_LK(1) : passes <= 2 ? PDJ_LK(2) : passes <= 4 ? PDJ_LK(4) : PDJ_LK(8);
#undef PDJ_LK
  }
  // Random 64-byte row gathers are latency-bound: what matters is rows in flight per SM.  With one block
  // per SM (the barrier-slot budget of a channel) that means many warps per block: 32 warps, two samples
  // in flight each.  The single-GPU emulation keeps small blocks so that all ranks stay co-resident.
  static const int knob_threads = getenv("PDJ_EMB_THREADS") ? atoi(getenv("PDJ_EMB_THREADS")) : 1024;
  int threads = emulate ? 256 : knob_threads;
  if (threads < 32 || threads > 1024 || threads % 32 != 0) threads = 256;
  if (!no_tma && t->dim >= 8 && t->dim <= 64 && S <= 32) {
    const size_t per_warp = (size_t)2 * 32 * t->dim * 4;
    while ((size_t)(threads / 32) * per_warp > 160 * 1024 && threads > 32) threads -= 32;
    const size_t smem = (size_t)(threads / 32) * per_warp;
    static size_t attr_set_for = 0;
    if (smem > 48 * 1024 && smem > attr_set_for) {
      PDJ_CUDA_CHECK(cudaFuncSetAttribute(emb_lookup_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      attr_set_for = smem;
    }
    return launch_emb(emb_lookup_tma_kernel, nblocks, threads, smem, emulate, c->nranks, st, *c, *t, ids_off, out_deep_off,
                      out_wide_off, B, S, barrier, out_pitch, dense_off, dense_n);
  }
  return launch_emb(emb_lookup_kernel, nblocks, threads, (size_t)0, emulate, c->nranks, st, *c, *t, ids_off, out_deep_off,
                    out_wide_off, B, S, barrier, out_pitch, dense_off, dense_n);
}

int pdj_emb_push(const PdjComm* c, const PdjEmbTable* t, size_t ids_off, size_t gdeep_off, size_t gwide_off, int B,
                 int S, int async_sgd, float lr, float grad_scale, int nblocks, int emulate, cudaStream_t st) {
  return pdj_emb_push2(c, t, ids_off, gdeep_off, gwide_off, B, S, async_sgd, lr, grad_scale, S * t->dim, nblocks, emulate, st);
}

int pdj_emb_push2(const PdjComm* c, const PdjEmbTable* t, size_t ids_off, size_t gdeep_off, size_t gwide_off, int B,
                  int S, int async_sgd, float lr, float grad_scale, int g_pitch, int nblocks, int emulate, cudaStream_t st) {
  if (check_table(t)) return -1;
  if (g_pitch < S * t->dim || g_pitch % 4 != 0 || gdeep_off % 8 != 0) {
    pdj_set_error("pdj_emb_push2: g_pitch (%d) must be a multiple of 4 and at least S * dim = %d", g_pitch, S * t->dim);
    return -1;
  }
  static const int knob_threads = getenv("PDJ_EMB_THREADS") ? atoi(getenv("PDJ_EMB_THREADS")) : 1024;
  int threads = emulate ? 256 : knob_threads;
  if (threads < 32 || threads > 1024 || threads % 32 != 0) threads = 256;
  const char* how = getenv("PDJ_EMB_PUSH");  // batched (default) | loop
  const int passes = (S + (32 / (t->dim >> 2)) - 1) / (32 / (t->dim >> 2));
  if ((how == nullptr || strcmp(how, "batched") == 0) && passes <= 8) {
    // 512 threads (41 k registers) leave room for a weight-gradient CTA of the tower on the same SM: the push runs
    // next to those kernels (ops/_sink.py, side stream).  PDJ_EMB_PUSH_THREADS overrides.
    const char* pt = getenv("PDJ_EMB_PUSH_THREADS");
    int bt = emulate ? 256 : (pt ? atoi(pt) : 512);
    if (bt < 32 || bt > (passes <= 2 ? 1024 : 768) || bt % 32 != 0) bt = 512;
#define PDJ_PS(P) launch_emb(emb_push_batched_kernel<P>, nblocks, bt, (size_t)0, emulate, c->nranks, st, *c, *t, ids_off, \
                             gdeep_off, gwide_off, B, S, async_sgd ? 1 : 0, lr, grad_scale, g_pitch)
    return passes <= 1 ? PDJ_PS(1) : passes <= 2 ? PDJ_PS(2) : passes <= 4 ? PDJ_PS(4) : PDJ_PS(8);
#undef PDJ_PS
  }
  return launch_emb(emb_push_kernel, nblocks, threads, (size_t)0, emulate, c->nranks, st, *c, *t, ids_off, gdeep_off, gwide_off,
                    B, S, async_sgd ? 1 : 0, lr, grad_scale, g_pitch);
}

int pdj_emb_apply(const PdjComm* c, const PdjEmbTable* t, int optimizer, float lr, float eps, int nblocks,
                  int emulate, cudaStream_t st) {
  if (check_table(t)) return -1;
  if (nblocks < 1) nblocks = 1;
  dim3 grid(nblocks, emulate ? c->nranks : 1);
  emb_apply_kernel<<<grid, 256, 0, st>>>(*c, *t, optimizer, lr, eps);
  PDJ_CUDA_CHECK(cudaGetLastError());
  emb_reset_count_kernel<<<dim3(1, emulate ? c->nranks : 1), 32, 0, st>>>(*c, *t);
  PDJ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // extern "C"
