// Fused Linear (+bias)(+ReLU)(+mask) on the 5th-generation tensor cores:
//   out[M,N] = epilogue( A[M,K] . B[N,K]^T )       bf16 in, fp32 accumulate in TMEM, bf16 out
//
// Blackwell-native structure (sm_100a only):
//   * persistent: one CTA per SM walks the output tiles; operands arrive by TMA
//     (cp.async.bulk.tensor.2d, 128-byte swizzle) into a multi-stage shared-memory
//     ring guarded by mbarriers;
//   * one elected thread issues tcgen05.mma (cta_group::1, kind::f16, M=128,
//     N=block_n, K=16) into one of two accumulator buffers in tensor memory;
//     tcgen05.commit releases ring slots and signals the epilogue, which drains
//     buffer i while the MMAs of tile i + 1 fill the other one;
//   * four epilogue warps read the accumulator with tcgen05.ld (32 lanes x 32 bit
//     x 16 columns), add the bias, apply ReLU and/or a ReLU-derivative mask, pack
//     to bf16, stage the tile in shared memory and write it with one TMA store.
// Out-of-range rows / columns / K tail are handled by TMA zero fill and guarded
// stores, so M, N, K need no padding (only 16-byte aligned row pitches).
//
// Used by the Wide&Deep MLP (models/wide_deep.py) forward and input-gradient
// GEMMs.  Reference parity: none (the reference has no compute code); workload of
// deploy/examples/wide_and_deep.yaml.
#include <cuda.h>
#include <stdlib.h>

#include "common.cuh"
#include "../../include/pdjob_b200.h"

namespace {

constexpr int BM = 128;          // rows of A per tile = UMMA M
constexpr int BK = 64;           // K per stage: 64 bf16 = one 128-byte swizzle row
constexpr int UK = 16;           // K per tcgen05.mma for 16-bit inputs
constexpr int MAX_BN = 256;
constexpr int MAX_STAGES = 8;   // barrier slots; the operand ring of a launch uses up to this many
constexpr int A_STAGE_BYTES = BM * BK * 2;  // 16 KiB
constexpr int TMEM_COLS = 512;              // two accumulator buffers of up to 256 columns
constexpr int EPI_WARPS = 8;                // two warps per TMEM lane quarter, each takes half of the columns
constexpr int THREADS = 64 + 32 * EPI_WARPS; // warp 0: TMA, warp 1: MMA + TMEM owner, warps 2..9: epilogue
constexpr int SMEM_LIMIT = 227 * 1024;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t"
      "}" ::"r"(bar),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst),
      "l"(map), "r"(c0), "r"(c1), "r"(bar)
      : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, uint32_t src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(map), "r"(src), "r"(c0),
               "r"(c1)
               : "memory");
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, int c0, int c1, int c2, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];" ::"r"(dst),
      "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(bar)
      : "memory");
}
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* map, uint32_t src, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(map), "r"(src), "r"(c0),
               "r"(c1), "r"(c2)
               : "memory");
}
// Shared-memory matrix descriptor: K-major tile, 128-byte swizzle, rows 128 B apart,
// 8-row groups 1024 B apart (SBO), descriptor version 1 (sm_100).
__device__ __forceinline__ uint64_t umma_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);   // start address, bits [0,14)
  d |= (uint64_t)1 << 16;                        // leading byte offset (unused for swizzled K-major), bits [16,30)
  d |= (uint64_t)(1024 >> 4) << 32;              // stride byte offset, bits [32,46)
  d |= (uint64_t)1 << 46;                        // version
  d |= (uint64_t)2 << 61;                        // SWIZZLE_128B
  return d;
}
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "l"(da), "l"(db), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

// Shared-memory ring + accumulator hand-off barriers of one CTA.
struct Pipe {
  uint32_t a_smem, b_smem, b_stage_bytes;
  uint32_t full0, empty0;      // per ring slot: TMA -> MMA, MMA -> TMA
  uint32_t tfull0, tempty0;    // per accumulator buffer: MMA -> epilogue, epilogue -> MMA
  int stages;
};

// TMA producer (one thread): walks this CTA's tiles and keeps the ring full.
__device__ __forceinline__ void producer_loop(const CUtensorMap* tma_a, const CUtensorMap* tma_b, const Pipe& p, int total,
                                              int tiles_n, int block_n, int num_kb) {
  const uint32_t stage_bytes = (uint32_t)A_STAGE_BYTES + p.b_stage_bytes;
  uint32_t it = 0;
  for (int t = blockIdx.x; t < total; t += gridDim.x) {
    const int m0 = (t / tiles_n) * BM, n0 = (t % tiles_n) * block_n;
    for (int kb = 0; kb < num_kb; ++kb, ++it) {
      const uint32_t s = it % (uint32_t)p.stages, ph = (it / (uint32_t)p.stages) & 1u;
      mbar_wait(p.empty0 + 8 * s, ph ^ 1u);
      mbar_expect_tx(p.full0 + 8 * s, stage_bytes);
      tma_load_2d(p.a_smem + s * A_STAGE_BYTES, tma_a, kb * BK, m0, p.full0 + 8 * s);
      tma_load_2d(p.b_smem + s * p.b_stage_bytes, tma_b, kb * BK, n0, p.full0 + 8 * s);
    }
  }
}

// MMA issuer (one thread): tile i accumulates into TMEM buffer i & 1.
__device__ __forceinline__ void mma_loop(const Pipe& p, uint32_t tmem_base, int total, int block_n, int num_kb) {
  // instruction descriptor: D = f32, A = B = bf16, both K-major, N >> 3, M >> 4
  const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(block_n >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
  uint32_t it = 0, i = 0;
  for (int t = blockIdx.x; t < total; t += gridDim.x, ++i) {
    const uint32_t buf = i & 1u, aph = (i >> 1) & 1u;
    mbar_wait(p.tempty0 + 8 * buf, aph ^ 1u);  // the epilogue has drained this accumulator
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_d = tmem_base + buf * MAX_BN;
    for (int kb = 0; kb < num_kb; ++kb, ++it) {
      const uint32_t s = it % (uint32_t)p.stages, ph = (it / (uint32_t)p.stages) & 1u;
      mbar_wait(p.full0 + 8 * s, ph);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
      for (int k = 0; k < BK / UK; ++k) {
        const uint64_t da = umma_desc(p.a_smem + s * A_STAGE_BYTES + k * UK * 2);
        const uint64_t db = umma_desc(p.b_smem + s * p.b_stage_bytes + k * UK * 2);
        umma_f16(tmem_d, da, db, idesc, (uint32_t)((kb | k) != 0));
      }
      umma_commit(p.empty0 + 8 * s);  // slot is free once these MMAs have read it
    }
    umma_commit(p.tfull0 + 8 * buf);  // accumulator complete
  }
}

// Persistent: one CTA per SM walks tiles t = blockIdx.x, blockIdx.x + gridDim.x, ... (column tile
// fastest, so CTAs running at the same time share A rows in L2).  The accumulator is double
// buffered in TMEM: the epilogue of tile i overlaps the MMAs of tile i + 1.
template <bool TMA_STORE>
__global__ void __launch_bounds__(THREADS, 1)
linear_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                      const __grid_constant__ CUtensorMap tma_o, __nv_bfloat16* __restrict__ out,
                      const float* __restrict__ bias, const __nv_bfloat16* __restrict__ mask, int M, int N, int K,
                      int ldo, int block_n, int stages, int relu) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;  // SWIZZLE_128B wants 1024-byte alignment
  const uint32_t b_stage_bytes = (uint32_t)block_n * BK * 2;
  const uint32_t a_smem = base;
  const uint32_t b_smem = a_smem + stages * A_STAGE_BYTES;
  const uint32_t stage_out = b_smem + stages * b_stage_bytes;                    // [128][block_n] bf16 staging
  const uint32_t bias_smem = stage_out + (TMA_STORE ? (uint32_t)BM * block_n * 2 : 0u);      // fp32 [block_n]
  const uint32_t bars = bias_smem + MAX_BN * 4;
  const uint32_t full0 = bars, empty0 = bars + 8 * MAX_STAGES;
  const uint32_t tfull0 = bars + 16 * MAX_STAGES, tempty0 = tfull0 + 16, tmem_slot = tempty0 + 16;
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles_n = (N + block_n - 1) / block_n;
  const int total = ((M + BM - 1) / BM) * tiles_n;
  const int num_kb = (K + BK - 1) / BK;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_b) : "memory");
    if (TMA_STORE) asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_o) : "memory");
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < stages; ++s) {
        mbar_init(full0 + 8 * s, 1);
        mbar_init(empty0 + 8 * s, 1);
      }
      for (int b = 0; b < 2; ++b) {
        mbar_init(tfull0 + 8 * b, 1);
        mbar_init(tempty0 + 8 * b, EPI_WARPS);  // one arrival per epilogue warp
      }
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot_ptr;
  const Pipe pipe{a_smem, b_smem, b_stage_bytes, full0, empty0, tfull0, tempty0, stages};

  if (warp == 0) {
    if (lane == 0) producer_loop(&tma_a, &tma_b, pipe, total, tiles_n, block_n, num_kb);
  } else if (warp == 1) {
    if (lane == 0) mma_loop(pipe, tmem_base, total, block_n, num_kb);
  } else {
    // ===== epilogue: warps 2..9.  Warp w reads TMEM lane quarter (w % 4) - a hardware rule -
    // and the two warps of a quarter split the columns in halves. =====
    const int q = warp & 3;
    const int half = (warp - 2) >> 2;           // 0: lower column half, 1: upper
    const int et = threadIdx.x - 64;            // 0..255 within the epilogue group
    const int ncols_half = ((block_n / 16 + 1) / 2) * 16;
    const int c_begin = half * ncols_half;
    const int c_end = (c_begin + ncols_half < block_n) ? c_begin + ncols_half : block_n;
    float* bias_s = reinterpret_cast<float*>(smem_raw + (bias_smem - raw));
    uint32_t i = 0;
    for (int t = blockIdx.x; t < total; t += gridDim.x, ++i) {
      const int m0 = (t / tiles_n) * BM, n0 = (t % tiles_n) * block_n;
      const uint32_t buf = i & 1u, aph = (i >> 1) & 1u;
      // bias slice of this column tile -> shared memory (overlaps the wait for the accumulator)
      if (TMA_STORE && et == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      asm volatile("bar.sync 1, %0;" ::"n"(32 * EPI_WARPS) : "memory");  // previous tile's bias / staging are free
      for (int j = et; j < block_n; j += 32 * EPI_WARPS) bias_s[j] = (bias != nullptr && n0 + j < N) ? __ldg(bias + n0 + j) : 0.f;
      asm volatile("bar.sync 1, %0;" ::"n"(32 * EPI_WARPS) : "memory");
      mbar_wait(tfull0 + 8 * buf, aph);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const int row = m0 + q * 32 + lane;
      const uint32_t taddr_row = tmem_base + buf * MAX_BN + ((uint32_t)(q * 32) << 16);
      for (int c0 = c_begin; c0 < c_end; c0 += 32) {
        // two 16-column loads in flight, one wait
        uint32_t r[32];
        const bool second = c0 + 16 < c_end;
        asm volatile(
            "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
            : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
              "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
            : "r"(taddr_row + (uint32_t)c0)
            : "memory");
        if (second) {
          asm volatile(
              "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
              : "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
              : "r"(taddr_row + (uint32_t)c0 + 16u)
              : "memory");
        }
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          if (h == 1 && !second) break;
          const int cc = c0 + 16 * h;
          const int col = n0 + cc;
          float v[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            v[j] = __uint_as_float(r[16 * h + j]) + bias_s[cc + j];
            if (relu) v[j] = fmaxf(v[j], 0.f);
          }
          if (mask != nullptr && row < M) {
            const __nv_bfloat16* mk = mask + (size_t)row * ldo + col;
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (col + j < N && !(__bfloat162float(mk[j]) > 0.f)) v[j] = 0.f;
          }
          uint4 p0, p1;
          p0.x = pdj_f2_to_bf2(v[0], v[1]); p0.y = pdj_f2_to_bf2(v[2], v[3]);
          p0.z = pdj_f2_to_bf2(v[4], v[5]); p0.w = pdj_f2_to_bf2(v[6], v[7]);
          p1.x = pdj_f2_to_bf2(v[8], v[9]); p1.y = pdj_f2_to_bf2(v[10], v[11]);
          p1.z = pdj_f2_to_bf2(v[12], v[13]); p1.w = pdj_f2_to_bf2(v[14], v[15]);
          if (TMA_STORE) {
            const uint32_t dst = stage_out + (uint32_t)((q * 32 + lane) * block_n + cc) * 2;
            asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(dst), "r"(p0.x), "r"(p0.y), "r"(p0.z), "r"(p0.w) : "memory");
            asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(dst + 16), "r"(p1.x), "r"(p1.y), "r"(p1.z), "r"(p1.w) : "memory");
          } else if (row < M && col < N) {
            __nv_bfloat16* o = out + (size_t)row * ldo + col;
            if (col + 16 <= N && (((size_t)row * ldo + col) & 7) == 0) {
              reinterpret_cast<uint4*>(o)[0] = p0;
              reinterpret_cast<uint4*>(o)[1] = p1;
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j)
                if (col + j < N) o[j] = __float2bfloat16(v[j]);
            }
          }
        }
      }
      // this accumulator buffer may be overwritten by the MMA warp now
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty0 + 8 * buf);
      if (TMA_STORE) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> visible to the TMA engine
        asm volatile("bar.sync 1, %0;" ::"n"(32 * EPI_WARPS) : "memory");
        if (et == 0) {
          tma_store_2d(&tma_o, stage_out, n0, m0);  // rows / columns beyond M, N are clipped by the tensor map
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
      }
    }
    if (TMA_STORE && et == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
  }
}

// ---------------------------------------------------------------------------------------------
// 1x1 convolution (NHWC) = GEMM over [N*H*W, Cin] x [Cout, Cin]^T, fused with the per-channel
// sum / sum-of-squares that the following training-mode BatchNorm needs, so the activation is not
// read a second time for its statistics.
//
// These layers are HBM-bound (K is as small as 64), so the epilogue is built for store bandwidth:
//   * the 8 epilogue warps form two groups of four (one warp per TMEM lane quarter); group g drains
//     accumulator buffer g, i.e. the groups work on alternate tiles in parallel;
//   * a tile leaves as 64-column boxes: TMEM -> registers -> bf16 -> 128-byte-swizzled shared
//     memory (conflict-free 16-byte stores) -> one TMA store per box, two boxes in flight per group;
//   * while the TMA engine reads a box, the group sums its columns from the same staging buffer
//     (values exactly as stored, i.e. bf16-rounded) and folds them into per-thread accumulators.
// The grid is a multiple of the column-tile count, so a CTA always works on the same channel
// range and the sums stay in registers until the end; each (CTA, group) then writes one row
//   partial[p][0][c] = sum, partial[p][1][c] = sum of squares,   p = (cta / tiles_n) * 2 + group
// which is the layout bn_act.cu's finalisation consumes.  Fixed schedule => bit-reproducible.
constexpr int BOX_N = 64;                       // columns per staged box: 64 bf16 = one swizzle row
constexpr int BOX_BYTES = BM * BOX_N * 2;       // 16 KiB
constexpr int SW_STAGING_BYTES = 4 * BOX_BYTES; // 2 groups x 2 boxes
constexpr int SW_SCRATCH_BYTES = 2 * 2 * 4 * BOX_N * 4;  // [group][sum|sq][warp][64] fp32

// MODE_ADD: out = addend + A.B^T instead (no statistics).  The addend box is fetched by TMA into the
// staging buffer one box ahead, summed in place with the accumulator, and stored from there - the
// residual-gradient add of a bottleneck block folded into the input-gradient GEMM of its first conv.
constexpr int MODE_PLAIN = 0, MODE_STATS = 1, MODE_ADD = 2;
constexpr int WIN_CHUNK = 28;  // WIN: output rows per work unit (3 halo input rows per unit)

//
// WIN (the ResNet stem): the same pipeline as a 4x4 / stride-1 convolution over the space-to-depth input
// [n][Hi][Wi][16].  One tile = one output row (n, y) of up to 128 pixels; its A operand for K block jy is
// the 3-D TMA box {64 elements, 128 pixels, row n*Hi + y + jy} of a tensor map whose pixel stride is 16
// elements - consecutive pixels' 64-element windows (4 taps x 16 channels) overlap in memory, so no
// im2col buffer exists anywhere.  Pixels beyond Wo are zero-filled on load and clipped on store.
template <int MODE, bool WIN>
__global__ void __launch_bounds__(THREADS, 1)
conv1x1_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                       const __grid_constant__ CUtensorMap tma_o, const __grid_constant__ CUtensorMap tma_c,
                       float* __restrict__ partial, int M, int N, int K, int block_n, int stages, int win_ho,
                       int win_hi) {
  constexpr bool STATS = MODE == MODE_STATS;
  constexpr bool ADD = MODE == MODE_ADD;
  static_assert(!(WIN && ADD), "the stem has no addend mode");
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t b_stage_bytes = (uint32_t)block_n * BK * 2;
  const uint32_t a_smem = base;
  const uint32_t b_smem = a_smem + stages * A_STAGE_BYTES;
  const uint32_t staging = b_smem + (WIN ? 4u : (uint32_t)stages) * b_stage_bytes;  // 1024-aligned (WIN: all 4 weight k-blocks resident)
  const uint32_t scratch = staging + SW_STAGING_BYTES;
  const uint32_t bars = scratch + SW_SCRATCH_BYTES;
  const uint32_t full0 = bars, empty0 = bars + 8 * MAX_STAGES;
  const uint32_t tfull0 = bars + 16 * MAX_STAGES, tempty0 = tfull0 + 16, tmem_slot = tempty0 + 16;
  const uint32_t abar0 = tmem_slot + 16;  // [group][staging buffer]: addend box has landed (MODE_ADD)
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - raw));
  float* scratch_f = reinterpret_cast<float*>(smem_raw + (scratch - raw));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles_n = N / block_n;
  const int total = WIN ? M : ((M + BM - 1) / BM) * tiles_n;  // WIN: M counts output rows, one tile each
  const int win_upi = WIN ? (win_ho + WIN_CHUNK - 1) / WIN_CHUNK : 1;   // work units (row chunks) per image
  const int win_units = WIN ? (M / win_ho) * win_upi : 0;
  const int num_kb = (K + BK - 1) / BK;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_b) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_o) : "memory");
    if (ADD) asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_c) : "memory");
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < stages; ++s) {
        mbar_init(full0 + 8 * s, 1);
        mbar_init(empty0 + 8 * s, 1);
      }
      for (int b = 0; b < 2; ++b) {
        mbar_init(tfull0 + 8 * b, 1);
        mbar_init(tempty0 + 8 * b, 4);  // the four warps of the group that owns the buffer
      }
      for (int b = 0; b < 4; ++b) mbar_init(abar0 + 8 * b, 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot_ptr;
  const Pipe pipe{a_smem, b_smem, b_stage_bytes, full0, empty0, tfull0, tempty0, stages};

  if (warp == 0) {
    if (lane == 0) {
      if (WIN) {
        // weights: all four k-blocks once
        mbar_expect_tx(abar0, 4u * b_stage_bytes);
        for (int kb = 0; kb < 4; ++kb) tma_load_2d(b_smem + kb * b_stage_bytes, &tma_b, kb * BK, 0, abar0);
        // input rows: a ring; output row y needs rows y .. y+3, so one new row per tile within a unit
        uint32_t ix = 0;
        for (int u = blockIdx.x; u < win_units; u += gridDim.x) {
          const int n = u / win_upi, y0 = (u % win_upi) * WIN_CHUNK;
          const int rows = (win_ho - y0) < WIN_CHUNK ? (win_ho - y0) : WIN_CHUNK;
          for (int r = 0; r < rows + 3; ++r, ++ix) {
            const uint32_t s = ix % (uint32_t)stages, ph = (ix / (uint32_t)stages) & 1u;
            mbar_wait(empty0 + 8 * s, ph ^ 1u);
            mbar_expect_tx(full0 + 8 * s, A_STAGE_BYTES);
            tma_load_3d(a_smem + s * A_STAGE_BYTES, &tma_a, 0, 0, n * win_hi + y0 + r, full0 + 8 * s);
          }
        }
      } else {
        producer_loop(&tma_a, &tma_b, pipe, total, tiles_n, block_n, num_kb);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      if (WIN) {
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(block_n >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
        const uint32_t S = (uint32_t)stages;
        mbar_wait(abar0, 0);
        uint32_t ix = 0, i = 0;
        for (int u = blockIdx.x; u < win_units; u += gridDim.x) {
          const int y0 = (u % win_upi) * WIN_CHUNK;
          const int rows = (win_ho - y0) < WIN_CHUNK ? (win_ho - y0) : WIN_CHUNK;
          for (int r = 0; r < 3; ++r) mbar_wait(full0 + 8 * ((ix + r) % S), ((ix + r) / S) & 1u);
          for (int j = 0; j < rows; ++j, ++i) {
            const uint32_t buf = i & 1u, aph = (i >> 1) & 1u;
            mbar_wait(tempty0 + 8 * buf, aph ^ 1u);
            mbar_wait(full0 + 8 * ((ix + j + 3) % S), ((ix + j + 3) / S) & 1u);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t tmem_d = tmem_base + buf * MAX_BN;
#pragma unroll
            for (int kb = 0; kb < 4; ++kb) {
              const uint32_t slot = (ix + j + kb) % S;
#pragma unroll
              for (int k = 0; k < BK / UK; ++k) {
                const uint64_t da = umma_desc(a_smem + slot * A_STAGE_BYTES + k * UK * 2);
                const uint64_t db = umma_desc(b_smem + kb * b_stage_bytes + k * UK * 2);
                umma_f16(tmem_d, da, db, idesc, (uint32_t)((kb | k) != 0));
              }
            }
            umma_commit(empty0 + 8 * ((ix + j) % S));  // input row y0 + j has served its last output row
            if (j == rows - 1)
              for (int r = 1; r <= 3; ++r) umma_commit(empty0 + 8 * ((ix + j + r) % S));
            umma_commit(tfull0 + 8 * buf);
          }
          ix += rows + 3;
        }
      } else {
        mma_loop(pipe, tmem_base, total, block_n, num_kb);
      }
    }
  } else {
    const int g = (warp - 2) >> 2;               // epilogue group = accumulator buffer it drains
    const int q = warp & 3;                      // TMEM lane quarter this warp may read
    const int gw = (warp - 2) & 3;               // warp index within the group
    const int gt = gw * 32 + lane;               // thread index within the group, 0..127
    const int row = q * 32 + lane;               // accumulator row handled by this thread
    const uint32_t bar_id = 1 + g;
    const uint32_t stg0 = staging + (uint32_t)g * 2 * BOX_BYTES;
    float* sc_sum = scratch_f + g * (2 * 4 * BOX_N);
    float* sc_sq = sc_sum + 4 * BOX_N;
    const int nboxes = block_n / BOX_N;
    const int n0 = ((int)blockIdx.x % tiles_n) * block_n;   // fixed per CTA: gridDim.x % tiles_n == 0
    float acc_s[MAX_BN / BOX_N], acc_q[MAX_BN / BOX_N];
#pragma unroll
    for (int b = 0; b < MAX_BN / BOX_N; ++b) acc_s[b] = acc_q[b] = 0.f;
    uint32_t bc = 0;  // boxes staged by this group so far
    if (ADD && gt == 0) {
      const long long t0 = (long long)blockIdx.x + (long long)g * gridDim.x;
      if (t0 < total) {  // addend of the group's first box
        mbar_expect_tx(abar0 + 8 * (g * 2), BOX_BYTES);
        tma_load_2d(stg0, &tma_c, n0, (int)(t0 / tiles_n) * BM, abar0 + 8 * (g * 2));
      }
    }
    int wu = blockIdx.x, wj = 0;  // WIN: unit and row within it of CTA-local tile i
    for (uint32_t i = 0;; ++i) {
      long long t;
      if (WIN) {
        if (wu >= win_units) break;
        const int y0 = (wu % win_upi) * WIN_CHUNK;
        const int rows = (win_ho - y0) < WIN_CHUNK ? (win_ho - y0) : WIN_CHUNK;
        t = (long long)(wu / win_upi) * win_ho + y0 + wj;  // output row index
        if (++wj == rows) { wj = 0; wu += gridDim.x; }
      } else {
        t = (long long)blockIdx.x + (long long)i * gridDim.x;
        if (t >= total) break;
      }
      if ((i & 1u) != (uint32_t)g) continue;  // the other group's tile
      const int m0 = (int)(t / tiles_n) * BM;
      const uint32_t aph = (i >> 1) & 1u;
      mbar_wait(tfull0 + 8 * g, aph);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t taddr_row = tmem_base + (uint32_t)g * MAX_BN + ((uint32_t)(q * 32) << 16);
#pragma unroll
      for (int b = 0; b < MAX_BN / BOX_N; ++b) {
        if (b >= nboxes) break;
        const uint32_t sbi = bc & 1u, use = bc >> 1;
        const uint32_t stg = stg0 + sbi * BOX_BYTES;
        ++bc;
        if (ADD) {
          // the addend box was requested one box ago, after the last store out of this buffer had been read
          mbar_wait(abar0 + 8 * (g * 2 + sbi), use & 1u);
        } else {
          // (A) the TMA store that last read this staging buffer (two boxes ago) has finished reading
          if (gt == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
          asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          uint32_t r[32];
          const uint32_t ta = taddr_row + (uint32_t)(b * BOX_N + h * 32);
          asm volatile(
              "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
              : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
              : "r"(ta)
              : "memory");
          asm volatile(
              "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
              : "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
              : "r"(ta + 16u)
              : "memory");
          asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
          for (int c = 0; c < 4; ++c) {  // four 16-byte chunks = 32 columns
            const uint32_t chunk = (uint32_t)(h * 4 + c);
            const uint32_t dst = stg + (uint32_t)row * 128u + ((chunk ^ ((uint32_t)row & 7u)) << 4);
            float v[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(r[8 * c + j]);
            if (ADD) {
              uint4 cv;
              asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(cv.x), "=r"(cv.y), "=r"(cv.z), "=r"(cv.w) : "r"(dst));
              const float2 a0 = pdj_bf2_to_f2(cv.x), a1 = pdj_bf2_to_f2(cv.y), a2 = pdj_bf2_to_f2(cv.z), a3 = pdj_bf2_to_f2(cv.w);
              v[0] += a0.x; v[1] += a0.y; v[2] += a1.x; v[3] += a1.y;
              v[4] += a2.x; v[5] += a2.y; v[6] += a3.x; v[7] += a3.y;
            }
            const uint32_t p0 = pdj_f2_to_bf2(v[0], v[1]);
            const uint32_t p1 = pdj_f2_to_bf2(v[2], v[3]);
            const uint32_t p2 = pdj_f2_to_bf2(v[4], v[5]);
            const uint32_t p3 = pdj_f2_to_bf2(v[6], v[7]);
            asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(dst), "r"(p0), "r"(p1), "r"(p2), "r"(p3) : "memory");
          }
        }
        if (b == nboxes - 1) {
          // accumulator fully read: hand the TMEM buffer back to the MMA warp
          asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
          __syncwarp();
          if (lane == 0) mbar_arrive(tempty0 + 8 * g);
        }
        // (B) box staged: make the generic-proxy writes visible to the TMA engine, then store
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
        if (gt == 0) {
          if (WIN) tma_store_3d(&tma_o, stg, b * BOX_N, 0, (int)t);  // pixels beyond Wo are clipped
          else tma_store_2d(&tma_o, stg, n0 + b * BOX_N, m0);        // rows beyond M are clipped by the tensor map
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          if (ADD) {
            // request the next box's addend into the other buffer once the store out of it (the previous
            // box) has been read; the store just issued may still be in flight
            int nrow = m0, ncol = n0 + (b + 1) * BOX_N;
            bool more = b + 1 < nboxes;
            if (!more) {
              const long long t2 = t + 2ll * gridDim.x;
              more = t2 < total;
              nrow = (int)(t2 / tiles_n) * BM;
              ncol = n0;
            }
            if (more) {
              asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
              const uint32_t nb_ = abar0 + 8 * (g * 2 + (sbi ^ 1u));
              mbar_expect_tx(nb_, BOX_BYTES);
              tma_load_2d(stg0 + (sbi ^ 1u) * BOX_BYTES, &tma_c, ncol, nrow, nb_);
            }
          }
        }
        if (STATS) {
          // column pair `lane` of rows gw*32 .. gw*32+31, read back exactly as stored
          float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
          const uint32_t chunk = (uint32_t)lane >> 2, word = ((uint32_t)lane & 3u) << 2;
#pragma unroll 8
          for (int rr = 0; rr < 32; ++rr) {
            const uint32_t r2 = (uint32_t)(gw * 32 + rr);
            uint32_t v;
            asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(stg + r2 * 128u + ((chunk ^ (r2 & 7u)) << 4) + word));
            const float2 f = pdj_bf2_to_f2(v);
            s0 += f.x;
            s1 += f.y;
            q0 = fmaf(f.x, f.x, q0);
            q1 = fmaf(f.y, f.y, q1);
          }
          sc_sum[gw * BOX_N + 2 * lane] = s0;
          sc_sum[gw * BOX_N + 2 * lane + 1] = s1;
          sc_sq[gw * BOX_N + 2 * lane] = q0;
          sc_sq[gw * BOX_N + 2 * lane + 1] = q1;
          // (C) the four row-quarter partials of this box are in shared memory
          asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
          if (gt < BOX_N) {
            acc_s[b] += (sc_sum[gt] + sc_sum[BOX_N + gt]) + (sc_sum[2 * BOX_N + gt] + sc_sum[3 * BOX_N + gt]);
            acc_q[b] += (sc_sq[gt] + sc_sq[BOX_N + gt]) + (sc_sq[2 * BOX_N + gt] + sc_sq[3 * BOX_N + gt]);
          }
        }
      }
    }
    if (STATS && gt < BOX_N) {
      const size_t p = (size_t)((int)blockIdx.x / tiles_n) * 2 + g;
#pragma unroll
      for (int b = 0; b < MAX_BN / BOX_N; ++b) {
        if (b >= nboxes) break;
        partial[(p * 2 + 0) * N + n0 + b * BOX_N + gt] = acc_s[b];
        partial[(p * 2 + 1) * N + n0 + b * BOX_N + gt] = acc_q[b];
      }
    }
    if (gt == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
  }
}

// ---------------------------------------------------------------------------------------------
// Weight gradient of the ResNet stem (4x4 / stride 1, 16 -> 64 channels over the space-to-depth input):
//   dW[o][jy][jx][c] = sum over (n, y, x) of dY[n][y][x][o] * X[n][y + jy][x + jx][c]
// i.e. D[64, 256] = dY^T . im2col(X) with a reduction over every output pixel (3.2 M at batch 256) -
// a shape the library handles with 48 CTAs.  Here it is one persistent tcgen05 kernel:
//   * a k-step is one output row: the A operand is the dY row tile [pixels][64 channels], the B operand
//     for tap row jy is the X window tile of input row y + jy ([pixels][64 = 4 taps x 16 channels], the
//     same overlapping TMA windows as the forward).  Both are MN-major for the MMA (the contiguous
//     dimension is M / N, the reduction runs over the rows of the tile), 128-byte swizzle;
//   * X row tiles live in a ring and are reused by the four output rows that need them, so each step
//     fetches one X tile and one dY tile;
//   * tcgen05.mma M=64, N=64, four accumulators (one per jy) of 64 TMEM columns, fp32, kept for the whole
//     kernel; at the end each CTA writes its [64][256] partial and a second kernel adds the partials in
//     a fixed order (bit-reproducible) and casts to bf16.
constexpr int WG_RX = 8, WG_RY = 4;   // ring depths: X row tiles, dY row tiles
constexpr int WG_CHUNK = 28;          // output rows per work unit (3 halo X rows per unit)
constexpr int WG_THREADS = 192;       // warp 0: TMA, warp 1: MMA, warps 2..5: epilogue

// MN-major operand tile: rows = reduction index (128 B each, 64 elements of M or N), 8-row groups 1024 B apart.
__device__ __forceinline__ uint64_t umma_desc_mn(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)1 << 16;                 // leading byte offset: distance between 64-element groups (a single group here)
  d |= (uint64_t)(1024 >> 4) << 32;       // stride byte offset: next group of 8 reduction rows
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;                 // SWIZZLE_128B
  return d;
}

__global__ void __launch_bounds__(WG_THREADS, 1)
stem_wgrad_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_x, const __grid_constant__ CUtensorMap tma_dy,
                          float* __restrict__ partial, int n_img, int Ho, int Hi, int kp /* pixels per tile, % 16 == 0 */) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t tile = (uint32_t)kp * 128u;
  const uint32_t x_smem = base, dy_smem = x_smem + WG_RX * tile;
  const uint32_t bars = dy_smem + WG_RY * tile;
  const uint32_t xfull = bars, xempty = xfull + 8 * WG_RX, yfull = xempty + 8 * WG_RX, yempty = yfull + 8 * WG_RY;
  const uint32_t done = yempty + 8 * WG_RY, tmem_slot = done + 8;
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - raw));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int upi = (Ho + WG_CHUNK - 1) / WG_CHUNK;  // units per image
  const int units = n_img * upi;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_x) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_dy) : "memory");
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int i = 0; i < WG_RX; ++i) { mbar_init(xfull + 8 * i, 1); mbar_init(xempty + 8 * i, 1); }
      for (int i = 0; i < WG_RY; ++i) { mbar_init(yfull + 8 * i, 1); mbar_init(yempty + 8 * i, 1); }
      mbar_init(done, 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(256) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == 0) {
    if (lane == 0) {
      uint32_t ix = 0, iy = 0;
      for (int u = blockIdx.x; u < units; u += gridDim.x) {
        const int n = u / upi, y0 = (u % upi) * WG_CHUNK;
        const int rows = (Ho - y0) < WG_CHUNK ? (Ho - y0) : WG_CHUNK;
        for (int r = 0; r < rows + 3; ++r, ++ix) {
          const uint32_t s = ix % WG_RX, ph = (ix / WG_RX) & 1u;
          mbar_wait(xempty + 8 * s, ph ^ 1u);
          mbar_expect_tx(xfull + 8 * s, tile);
          tma_load_3d(x_smem + s * tile, &tma_x, 0, 0, n * Hi + y0 + r, xfull + 8 * s);
          if (r >= 3) {  // the dY row that becomes computable with this X row
            const uint32_t sy = iy % WG_RY, phy = (iy / WG_RY) & 1u;
            mbar_wait(yempty + 8 * sy, phy ^ 1u);
            mbar_expect_tx(yfull + 8 * sy, tile);
            tma_load_3d(dy_smem + sy * tile, &tma_dy, 0, 0, n * Ho + y0 + (r - 3), yfull + 8 * sy);
            ++iy;
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // D = f32, A = B = bf16, both MN-major (bits 15, 16), N = 64, M = 64
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((64u >> 3) << 17) | ((64u >> 4) << 24);
      const int ksteps = kp / 16;
      uint32_t ix = 0, iy = 0;
      for (int u = blockIdx.x; u < units; u += gridDim.x) {
        const int y0 = (u % upi) * WG_CHUNK;
        const int rows = (Ho - y0) < WG_CHUNK ? (Ho - y0) : WG_CHUNK;
        for (int r = 0; r < 3; ++r) mbar_wait(xfull + 8 * ((ix + r) % WG_RX), ((ix + r) / WG_RX) & 1u);
        for (int j = 0; j < rows; ++j, ++iy) {
          mbar_wait(xfull + 8 * ((ix + j + 3) % WG_RX), ((ix + j + 3) / WG_RX) & 1u);
          const uint32_t sy = iy % WG_RY;
          mbar_wait(yfull + 8 * sy, (iy / WG_RY) & 1u);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
          for (int jy = 0; jy < 4; ++jy) {
            const uint32_t xs = (ix + j + jy) % WG_RX;
            for (int k = 0; k < ksteps; ++k) {
              const uint64_t da = umma_desc_mn(dy_smem + sy * tile + (uint32_t)k * 2048u);
              const uint64_t db = umma_desc_mn(x_smem + xs * tile + (uint32_t)k * 2048u);
              umma_f16(tmem_base + (uint32_t)jy * 64u, da, db, idesc, (uint32_t)((iy | (uint32_t)k) != 0u));
            }
          }
          umma_commit(yempty + 8 * sy);
          umma_commit(xempty + 8 * ((ix + j) % WG_RX));       // input row y0 + j has served its last output row
          if (j == rows - 1)
            for (int r = 1; r <= 3; ++r) umma_commit(xempty + 8 * ((ix + j + r) % WG_RX));  // bottom halo rows
        }
        ix += rows + 3;
      }
      umma_commit(done);
    }
  } else {
    // epilogue: M = 64 accumulators occupy lanes 0..15 of each 32-lane TMEM quarter: row 16 * q + lane
    const int q = warp & 3;
    mbar_wait(done, 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    float* dst = partial + ((size_t)blockIdx.x * 64 + (size_t)(16 * q + lane)) * 256;
    const bool has_work = (int)blockIdx.x < units;
    for (int c0 = 0; c0 < 256; c0 += 16) {
      uint32_t r[16];
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
          : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
            "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
          : "r"(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c0)
          : "memory");
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      if (lane < 16) {
#pragma unroll
        for (int v = 0; v < 4; ++v) {
          float4 o;
          o.x = has_work ? __uint_as_float(r[4 * v + 0]) : 0.f;
          o.y = has_work ? __uint_as_float(r[4 * v + 1]) : 0.f;
          o.z = has_work ? __uint_as_float(r[4 * v + 2]) : 0.f;
          o.w = has_work ? __uint_as_float(r[4 * v + 3]) : 0.f;
          reinterpret_cast<float4*>(dst + c0)[v] = o;
        }
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(256) : "memory");
  }
}

// dw[i] = sum over the CTAs' partials, fixed order, cast to bf16.
__global__ void __launch_bounds__(256)
stem_wgrad_reduce_kernel(const float* __restrict__ partial, int nparts, __nv_bfloat16* __restrict__ dw) {
  const int i = blockIdx.x * 256 + threadIdx.x;  // 64 * 256 outputs
  float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
  int p = 0;
  for (; p + 3 < nparts; p += 4) {
    a0 += partial[(size_t)p * 16384 + i];
    a1 += partial[(size_t)(p + 1) * 16384 + i];
    a2 += partial[(size_t)(p + 2) * 16384 + i];
    a3 += partial[(size_t)(p + 3) * 16384 + i];
  }
  for (; p < nparts; ++p) a0 += partial[(size_t)p * 16384 + i];
  dw[i] = __float2bfloat16((a0 + a1) + (a2 + a3));
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn g_encode = nullptr;

// cuTensorMapEncodeTiled is a driver call and needs a current context.  A thread that has only made
// context-free runtime calls so far (an autograd worker entering backward) has none: bind the primary
// context of its current device.
int ensure_current_context() {
  typedef CUresult (*GetCur)(CUcontext*);
  typedef CUresult (*SetCur)(CUcontext);
  typedef CUresult (*Retain)(CUcontext*, CUdevice);
  typedef CUresult (*DevGet)(CUdevice*, int);
  static GetCur get_cur = nullptr;
  static SetCur set_cur = nullptr;
  static Retain retain = nullptr;
  static DevGet dev_get = nullptr;
  if (get_cur == nullptr) {
    void *a = nullptr, *b = nullptr, *c = nullptr, *d = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuCtxGetCurrent", &a, cudaEnableDefault, &q) != cudaSuccess || a == nullptr ||
        cudaGetDriverEntryPoint("cuCtxSetCurrent", &b, cudaEnableDefault, &q) != cudaSuccess || b == nullptr ||
        cudaGetDriverEntryPoint("cuDevicePrimaryCtxRetain", &c, cudaEnableDefault, &q) != cudaSuccess || c == nullptr ||
        cudaGetDriverEntryPoint("cuDeviceGet", &d, cudaEnableDefault, &q) != cudaSuccess || d == nullptr) {
      pdj_set_error("driver context entry points are unavailable");
      return -1;
    }
    set_cur = reinterpret_cast<SetCur>(b);
    retain = reinterpret_cast<Retain>(c);
    dev_get = reinterpret_cast<DevGet>(d);
    get_cur = reinterpret_cast<GetCur>(a);
  }
  CUcontext ctx = nullptr;
  if (get_cur(&ctx) == CUDA_SUCCESS && ctx != nullptr) return 0;
  int dev = 0;
  PDJ_CUDA_CHECK(cudaGetDevice(&dev));
  CUdevice cudev;
  CUresult r = dev_get(&cudev, dev);
  if (r == CUDA_SUCCESS) r = retain(&ctx, cudev);
  if (r == CUDA_SUCCESS) r = set_cur(ctx);
  if (r != CUDA_SUCCESS) {
    pdj_set_error("could not bind the primary context of device %d to this thread (CUresult %d)", dev, (int)r);
    return (int)r;
  }
  return 0;
}

int make_map(CUtensorMap* map, const void* ptr, int rows, int cols, int ld_elems, int box_cols, int box_rows, bool swizzle) {
  if (g_encode == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    if (e != cudaSuccess || p == nullptr || q != cudaDriverEntryPointSuccess) {
      pdj_set_error("cuTensorMapEncodeTiled is unavailable: %s", cudaGetErrorString(e));
      return -1;
    }
    g_encode = reinterpret_cast<EncodeTiledFn>(p);
  }
  if (int r = ensure_current_context()) return r;
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};       // innermost first
  const cuuint64_t strides[1] = {(cuuint64_t)ld_elems * 2};              // bytes, for dim 1
  const cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    pdj_set_error("cuTensorMapEncodeTiled failed with CUresult %d (rows=%d cols=%d ld=%d box=%dx%d ptr=%p)", (int)r, rows, cols,
                  ld_elems, box_rows, box_cols, ptr);
    return (int)r;
  }
  return 0;
}

// 3-D map {d0 contiguous elements, d1 with stride s1 bytes, d2 with stride s2 bytes}, 128-byte swizzle.
int make_map3(CUtensorMap* map, const void* ptr, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t s1, uint64_t s2, int b0,
              int b1) {
  if (g_encode == nullptr) {
    CUtensorMap dummy;
    static __nv_bfloat16 probe[64];
    (void)make_map(&dummy, probe, 1, 64, 64, 64, 1, false);  // resolves the driver entry point
    if (g_encode == nullptr) return -1;
  }
  if (int r = ensure_current_context()) return r;
  const cuuint64_t dims[3] = {d0, d1, d2};
  const cuuint64_t strides[2] = {s1, s2};
  const cuuint32_t box[3] = {(cuuint32_t)b0, (cuuint32_t)b1, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    pdj_set_error("cuTensorMapEncodeTiled (3-D) failed with CUresult %d (dims %llu x %llu x %llu, strides %llu, %llu)", (int)r,
                  (unsigned long long)d0, (unsigned long long)d1, (unsigned long long)d2, (unsigned long long)s1,
                  (unsigned long long)s2);
    return (int)r;
  }
  return 0;
}

bool g_attr_set = false;
int g_sms = 0;

int ensure_init() {
  if (g_attr_set) return 0;
  PDJ_CUDA_CHECK(cudaFuncSetAttribute(linear_tcgen05_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
  PDJ_CUDA_CHECK(cudaFuncSetAttribute(linear_tcgen05_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
  PDJ_CUDA_CHECK(cudaFuncSetAttribute(conv1x1_tcgen05_kernel<MODE_PLAIN, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
  PDJ_CUDA_CHECK(cudaFuncSetAttribute(conv1x1_tcgen05_kernel<MODE_STATS, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
  PDJ_CUDA_CHECK(cudaFuncSetAttribute(conv1x1_tcgen05_kernel<MODE_ADD, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
  PDJ_CUDA_CHECK(cudaFuncSetAttribute(conv1x1_tcgen05_kernel<MODE_PLAIN, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
  PDJ_CUDA_CHECK(cudaFuncSetAttribute(conv1x1_tcgen05_kernel<MODE_STATS, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
  int dev = 0;
  PDJ_CUDA_CHECK(cudaGetDevice(&dev));
  PDJ_CUDA_CHECK(cudaDeviceGetAttribute(&g_sms, cudaDevAttrMultiProcessorCount, dev));
  g_attr_set = true;
  return 0;
}

// Column tile of the 1x1-convolution kernel: the widest of 256/192/128/64 that divides N and still
// fills the last wave of the persistent grid to 85 % (the 7x7 layers have too few row tiles for
// 256-wide tiles); failing that, the candidate with the fullest last wave.
int conv_block_n(int M, int N, int sms) {
  const int tiles_m = (M + BM - 1) / BM;
  int best = 0;
  double best_eff = -1.0;
  for (int bn = MAX_BN; bn >= BOX_N; bn -= BOX_N) {
    if (N % bn != 0) continue;
    const long long tiles = (long long)tiles_m * (N / bn);
    const long long waves = (tiles + sms - 1) / sms;
    const double eff = (double)tiles / (double)(waves * sms);
    if (eff >= 0.85) return bn;
    if (eff > best_eff) { best_eff = eff; best = bn; }
  }
  return best;
}

int sm_count_cached() {
  if (g_sms == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&g_sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || g_sms <= 0)
      g_sms = 148;
  }
  return g_sms;
}

// Persistent grid: as many CTAs as SMs (or tiles), rounded down to a multiple of the column-tile
// count so that every CTA stays on one channel range.
int conv_grid(int M, int N, int bn, int sms) {
  const int tiles_n = N / bn;
  const long long total = (long long)((M + BM - 1) / BM) * tiles_n;
  long long g = total < sms ? total : sms;
  g -= g % tiles_n;
  if (g < tiles_n) g = tiles_n;
  return (int)g;
}

}  // namespace

int pdj_tma_map_2d_bf16(void* map_out, const void* ptr, int rows, int cols, int ld_elems, int box_cols, int box_rows,
                        int swizzle128) {
  return make_map(reinterpret_cast<CUtensorMap*>(map_out), ptr, rows, cols, ld_elems, box_cols, box_rows, swizzle128 != 0);
}

extern "C" {

int pdj_linear_block_n(int N) {
  // smallest multiple of 16 (<= 256) that covers N in the fewest column tiles
  const int tiles = (N + MAX_BN - 1) / MAX_BN;
  int bn = ((N + tiles - 1) / tiles + 15) / 16 * 16;
  if (bn < 16) bn = 16;
  if (bn > MAX_BN) bn = MAX_BN;
  return bn;
}

int pdj_linear_bf16(const void* a, const void* b, void* out, const float* bias, const void* mask, int M, int N, int K,
                    int lda, int ldb, int ldo, int relu, cudaStream_t st) {
  if (M < 1 || N < 1 || K < 1 || lda % 8 != 0 || ldb % 8 != 0 || ((uintptr_t)a & 15) || ((uintptr_t)b & 15)) {
    pdj_set_error("pdj_linear_bf16: need M,N,K >= 1, 16-byte aligned operands and row pitches that are multiples of 8 elements "
                  "(lda=%d ldb=%d)", lda, ldb);
    return -1;
  }
  const int bn = pdj_linear_block_n(N);
  // TMA store needs a 16-byte aligned output pitch; otherwise the epilogue stores directly.
  // Measured on B200 (profiles/gemm_bench.py): staging + TMA store wins for narrow outputs (it frees
  // the epilogue warps early), direct 32-byte stores win once the tile is 256 wide, where the staging
  // buffer would cost a pipeline stage.  PDJ_GEMM_DIRECT_STORE / PDJ_GEMM_TMA_STORE override.
  static const bool force_direct = getenv("PDJ_GEMM_DIRECT_STORE") != nullptr;
  static const bool force_tma = getenv("PDJ_GEMM_TMA_STORE") != nullptr;
  const bool tma_ok = (ldo % 8 == 0) && (((uintptr_t)out & 15) == 0);
  const bool tma_store = tma_ok && !force_direct && (force_tma || bn < MAX_BN);
  CUtensorMap ma, mb, mo;
  if (int r = make_map(&ma, a, M, K, lda, BK, BM, true)) return r;
  if (int r = make_map(&mb, b, N, K, ldb, BK, bn, true)) return r;
  if (tma_store) {
    if (int r = make_map(&mo, out, M, N, ldo, bn, BM, false)) return r;
  } else {
    mo = ma;
  }
  if (int r = ensure_init()) return r;
  const int stage_bytes = A_STAGE_BYTES + bn * BK * 2;
  const int fixed = 1024 /*alignment*/ + 256 /*barriers*/ + MAX_BN * 4 /*bias slice*/ + (tma_store ? BM * bn * 2 : 0);
  int stages = (SMEM_LIMIT - fixed) / stage_bytes;
  if (stages > MAX_STAGES) stages = MAX_STAGES;
  if (stages < 2) {
    pdj_set_error("pdj_linear_bf16: tile does not fit in shared memory (block_n=%d)", bn);
    return -1;
  }
  const int smem = fixed + stages * stage_bytes;
  const int total = ((M + BM - 1) / BM) * ((N + bn - 1) / bn);
  const int grid = total < g_sms ? total : g_sms;
  if (tma_store)
    linear_tcgen05_kernel<true><<<grid, THREADS, smem, st>>>(ma, mb, mo, (__nv_bfloat16*)out, bias, (const __nv_bfloat16*)mask,
                                                            M, N, K, ldo, bn, stages, relu);
  else
    linear_tcgen05_kernel<false><<<grid, THREADS, smem, st>>>(ma, mb, mo, (__nv_bfloat16*)out, bias,
                                                             (const __nv_bfloat16*)mask, M, N, K, ldo, bn, stages, relu);
  PDJ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int pdj_conv1x1_partial_rows(int M, int N) {
  if (M < 1 || N < BOX_N || N % BOX_N != 0) return 0;
  const int sms = sm_count_cached();
  const int bn = conv_block_n(M, N, sms);
  return 2 * conv_grid(M, N, bn, sms) / (N / bn);
}

static int launch_conv1x1(const void* a, const void* b, const void* addend, void* out, float* partial, int M, int N, int K,
                          int lda, int ldb, int ldc, int ldo, cudaStream_t st) {
  const int bn = (M >= 1 && N >= BOX_N && N % BOX_N == 0) ? conv_block_n(M, N, sm_count_cached()) : 0;
  if (M < 1 || K < 1 || bn == 0 || lda % 8 != 0 || ldb % 8 != 0 || ldo % 8 != 0 || ((uintptr_t)a & 15) ||
      ((uintptr_t)b & 15) || ((uintptr_t)out & 15) || (addend != nullptr && (ldc % 8 != 0 || ((uintptr_t)addend & 15)))) {
    pdj_set_error("pdj_conv1x1: need M,K >= 1, N a multiple of 64, 16-byte aligned operands and row pitches that are "
                  "multiples of 8 elements (N=%d lda=%d ldb=%d ldc=%d ldo=%d)", N, lda, ldb, ldc, ldo);
    return -1;
  }
  if (int r = ensure_init()) return r;
  CUtensorMap ma, mb, mo, mc;
  if (int r = make_map(&ma, a, M, K, lda, BK, BM, true)) return r;
  if (int r = make_map(&mb, b, N, K, ldb, BK, bn, true)) return r;
  if (int r = make_map(&mo, out, M, N, ldo, BOX_N, BM, true)) return r;
  if (addend != nullptr) {
    if (int r = make_map(&mc, addend, M, N, ldc, BOX_N, BM, true)) return r;
  } else {
    mc = mo;
  }
  const int stage_bytes = A_STAGE_BYTES + bn * BK * 2;
  const int fixed = 1024 /*alignment*/ + 256 /*barriers*/ + SW_STAGING_BYTES + SW_SCRATCH_BYTES;
  int stages = (SMEM_LIMIT - fixed) / stage_bytes;
  if (stages > MAX_STAGES) stages = MAX_STAGES;
  if (stages < 2) {
    pdj_set_error("pdj_conv1x1: tile does not fit in shared memory (block_n=%d)", bn);
    return -1;
  }
  const int smem = fixed + stages * stage_bytes;
  const int grid = conv_grid(M, N, bn, g_sms);
  if (addend != nullptr)
    conv1x1_tcgen05_kernel<MODE_ADD, false><<<grid, THREADS, smem, st>>>(ma, mb, mo, mc, nullptr, M, N, K, bn, stages, 0, 0);
  else if (partial != nullptr)
    conv1x1_tcgen05_kernel<MODE_STATS, false><<<grid, THREADS, smem, st>>>(ma, mb, mo, mc, partial, M, N, K, bn, stages, 0, 0);
  else
    conv1x1_tcgen05_kernel<MODE_PLAIN, false><<<grid, THREADS, smem, st>>>(ma, mb, mo, mc, nullptr, M, N, K, bn, stages, 0, 0);
  PDJ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

static int stem_units(int N, int Ho) { return N * ((Ho + WIN_CHUNK - 1) / WIN_CHUNK); }

int pdj_stem_s2d_partial_rows(int N, int Ho) {
  if (N < 1 || Ho < 1) return 0;
  const int sms = sm_count_cached(), units = stem_units(N, Ho);
  return 2 * (units < sms ? units : sms);
}

// ResNet stem: 4x4 / stride-1 convolution, 16 -> 64 channels, over the space-to-depth input
// x[N][Ho+3][Wo+3][16] (ops/optim.py), weights w[64][4][4][16] (channels-last memory of [64,16,4,4]),
// out[N][Ho][Wo][64]; optional BN partial sums as in pdj_conv1x1_bf16 (pdj_stem_s2d_partial_rows(N*Ho) rows).
int pdj_stem_s2d_fwd(const void* x, const void* w, void* out, float* partial, int N, int Ho, int Wo, cudaStream_t st) {
  if (N < 1 || Ho < 1 || Wo < 1 || Wo > BM || ((uintptr_t)x & 15) || ((uintptr_t)w & 15) || ((uintptr_t)out & 15) ||
      (long long)N * (Ho + 3) >= (1ll << 31)) {
    pdj_set_error("pdj_stem_s2d_fwd: need 1 <= Wo <= 128 and 16-byte aligned tensors (N=%d Ho=%d Wo=%d)", N, Ho, Wo);
    return -1;
  }
  if (int r = ensure_init()) return r;
  const int Hi = Ho + 3, Wi = Wo + 3, C = 16, O = 64, Kc = 4 * 4 * C;
  CUtensorMap ma, mb, mo;
  // pixel stride 16 elements, window 64 elements: overlapping rows of the same buffer
  if (int r = make_map3(&ma, x, 64, (uint64_t)Wo, (uint64_t)N * Hi, (uint64_t)C * 2, (uint64_t)Wi * C * 2, 64, BM)) return r;
  if (int r = make_map(&mb, w, O, Kc, Kc, BK, O, true)) return r;
  if (int r = make_map3(&mo, out, 64, (uint64_t)Wo, (uint64_t)N * Ho, (uint64_t)O * 2, (uint64_t)Wo * O * 2, 64, BM)) return r;
  const int bn = O;
  const int stages = 7;  // input-row ring: four rows in use by a tile, three being fetched
  const int smem = 1024 + 256 + SW_STAGING_BYTES + SW_SCRATCH_BYTES + stages * A_STAGE_BYTES + 4 * bn * BK * 2;
  const int rows = N * Ho;
  const int units = stem_units(N, Ho);
  const int grid = units < g_sms ? units : g_sms;
  if (partial != nullptr)
    conv1x1_tcgen05_kernel<MODE_STATS, true><<<grid, THREADS, smem, st>>>(ma, mb, mo, mo, partial, rows, O, Kc, bn, stages, Ho, Hi);
  else
    conv1x1_tcgen05_kernel<MODE_PLAIN, true><<<grid, THREADS, smem, st>>>(ma, mb, mo, mo, nullptr, rows, O, Kc, bn, stages, Ho, Hi);
  PDJ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int pdj_stem_s2d_wgrad_ws_bytes(void) { return sm_count_cached() * 64 * 256 * (int)sizeof(float); }

// dw[64][4][4][16] (bf16, channels-last memory of the [64,16,4,4] weight) = gradient of pdj_stem_s2d_fwd with
// respect to w, given x[N][Ho+3][Wo+3][16] and dy[N][Ho][Wo][64].  `ws`: pdj_stem_s2d_wgrad_ws_bytes() bytes.
int pdj_stem_s2d_wgrad(const void* x, const void* dy, void* dw, void* ws, int N, int Ho, int Wo, cudaStream_t st) {
  if (N < 1 || Ho < 1 || Wo < 1 || Wo > BM || ((uintptr_t)x & 15) || ((uintptr_t)dy & 15) || ws == nullptr ||
      (long long)N * (Ho + 3) >= (1ll << 31)) {
    pdj_set_error("pdj_stem_s2d_wgrad: need 1 <= Wo <= 128, 16-byte aligned tensors and a workspace (N=%d Ho=%d Wo=%d)", N, Ho, Wo);
    return -1;
  }
  if (int r = ensure_init()) return r;
  static bool attr = false;
  if (!attr) {
    PDJ_CUDA_CHECK(cudaFuncSetAttribute(stem_wgrad_tcgen05_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
    attr = true;
  }
  const int Hi = Ho + 3, Wi = Wo + 3, C = 16, O = 64;
  const int kp = (Wo + 15) / 16 * 16;
  CUtensorMap mx, my;
  if (int r = make_map3(&mx, x, 64, (uint64_t)Wo, (uint64_t)N * Hi, (uint64_t)C * 2, (uint64_t)Wi * C * 2, 64, kp)) return r;
  if (int r = make_map3(&my, dy, 64, (uint64_t)Wo, (uint64_t)N * Ho, (uint64_t)O * 2, (uint64_t)Wo * O * 2, 64, kp)) return r;
  const int units = N * ((Ho + WG_CHUNK - 1) / WG_CHUNK);
  const int grid = units < g_sms ? units : g_sms;
  const int smem = 1024 + 512 + (WG_RX + WG_RY) * kp * 128;
  stem_wgrad_tcgen05_kernel<<<grid, WG_THREADS, smem, st>>>(mx, my, (float*)ws, N, Ho, Hi, kp);
  PDJ_CUDA_CHECK(cudaGetLastError());
  stem_wgrad_reduce_kernel<<<64, 256, 0, st>>>((const float*)ws, grid, (__nv_bfloat16*)dw);
  PDJ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int pdj_conv1x1_bf16(const void* a, const void* b, void* out, float* partial, int M, int N, int K, int lda, int ldb,
                     int ldo, cudaStream_t st) {
  return launch_conv1x1(a, b, nullptr, out, partial, M, N, K, lda, ldb, 0, ldo, st);
}

int pdj_conv1x1_add_bf16(const void* a, const void* b, const void* addend, void* out, int M, int N, int K, int lda,
                         int ldb, int ldc, int ldo, cudaStream_t st) {
  if (addend == nullptr) {
    pdj_set_error("pdj_conv1x1_add_bf16: addend is NULL");
    return -1;
  }
  return launch_conv1x1(a, b, addend, out, nullptr, M, N, K, lda, ldb, ldc, ldo, st);
}

}  // extern "C"
