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
//   * eight epilogue warps (two per TMEM lane quarter) read the accumulator with
//     tcgen05.ld (32 lanes x 32 bit x 16 columns), add the bias, apply ReLU and/or a
//     ReLU-derivative mask, pack to bf16 and either stage the tile in shared memory
//     for one TMA store or store it directly.
// Out-of-range rows / columns / K tail are handled by TMA zero fill and guarded
// stores, so M, N, K need no padding (only 16-byte aligned row pitches).
//
// The pipeline pieces (PTX wrappers, operand ring, producer / MMA loops) live in
// tcgen05_common.cuh and are shared with the convolution kernels of conv_tcgen05.cu.
// Used by the Wide&Deep MLP (models/wide_deep.py) forward and input-gradient
// GEMMs.  Reference parity: none (the reference has no compute code); workload of
// deploy/examples/wide_and_deep.yaml.
#include <cuda.h>
#include <stdlib.h>

#include "common.cuh"
#include "tcgen05_common.cuh"
#include "../../include/pdjob_b200.h"

using namespace pdj_tc;

namespace {

// Persistent: one CTA per SM walks tiles t = blockIdx.x, blockIdx.x + gridDim.x, ... (column tile
// fastest, so CTAs running at the same time share A rows in L2).  The accumulator is double
// buffered in TMEM: the epilogue of tile i overlaps the MMAs of tile i + 1.
//
// PAIR2: the same pipeline on CTA pairs (cluster of two, tcgen05 cta_group::2).  A pair owns a 256-row tile: each CTA
// loads its own 128 rows of A and HALF of the B tile (block_n / 2 rows), the leader's elected thread issues M = 256
// MMAs that read both halves of B out of the two CTAs' shared memory and write 128 accumulator rows into each CTA's
// tensor memory, and tcgen05.commit multicasts the ring-slot / accumulator barriers to both CTAs.  Every byte of B
// is fetched from L2 once per 256 rows instead of once per 128: skinny layers (K, N ~ 400) are bound by exactly
// that traffic (150 MB L2 -> SM per 32768 x 400 x 432 layer against a measured ~12 TB/s ceiling).
template <bool TMA_STORE, bool PAIR2>
__global__ void __launch_bounds__(THREADS, 1)
linear_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                      const __grid_constant__ CUtensorMap tma_o, const __grid_constant__ CUtensorMap tma_m,
                      __nv_bfloat16* __restrict__ out,
                      const float* __restrict__ bias, const __nv_bfloat16* __restrict__ mask, int M, int N, int K,
                      int ldo, int block_n, int stages, int relu, int flags) {
  const int b_resident = flags & 1, b_mn = (flags >> 1) & 3;  // b_mn: B given as [K][N] (single-CTA path only), see Pipe
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;  // SWIZZLE_128B wants 1024-byte alignment
  const uint32_t b_stage_bytes = PAIR2 ? (uint32_t)(block_n / 2) * BK * 2 : b_stage_bytes_for(block_n, b_mn);  // PAIR2: this CTA's half of the B tile
  const uint32_t a_smem = base;
  const uint32_t b_smem = a_smem + stages * A_STAGE_BYTES;
  // PAIR2 with b_resident: this CTA's half of the whole B panel (every k-block of its column tile) stays in shared
  // memory for the life of the kernel and only A streams through the ring - possible because a pair always works
  // on the same column tile (the host checks that the persistent stride is a multiple of the column-tile count).
  const int num_kb_ = (K + BK - 1) / BK;
  const uint32_t stage_out = b_smem + (uint32_t)((PAIR2 && b_resident) ? num_kb_ : stages) * b_stage_bytes;  // [128][block_n] bf16 staging
  const uint32_t bias_smem = stage_out + (TMA_STORE ? (uint32_t)BM * block_n * 2 : 0u);      // fp32 [block_n]
  const uint32_t bars = bias_smem + MAX_BN * 4;
  const uint32_t full0 = bars, empty0 = bars + 8 * MAX_STAGES;
  const uint32_t tfull0 = bars + 16 * MAX_STAGES, tempty0 = tfull0 + 16, tmem_slot = tempty0 + 16;
  const uint32_t bres = tmem_slot + 16;  // resident B panel has landed (PAIR2 + b_resident)
  const uint32_t mbar_m = bres + 8;      // the tile's ReLU mask has landed in the staging tile (TMA_STORE + mask)
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - raw));

  const int dbg = relu >> 8;  // diagnostic bits (PDJ_LINEAR_DBG): 1 = no TMA store, 2 = no staging writes, 4 = no drain
  relu &= 0xff;
  const int warp = (int)warp_uniform(threadIdx.x >> 5), lane = threadIdx.x & 31;
  const int tiles_n = (N + block_n - 1) / block_n;
  constexpr int TILE_M = PAIR2 ? 2 * BM : BM;          // rows per work item (a CTA or a CTA pair)
  const int total = ((M + TILE_M - 1) / TILE_M) * tiles_n;
  const int num_kb = (K + BK - 1) / BK;
  const uint32_t crank = PAIR2 ? warp_uniform(cluster_ctarank()) : 0u;
  const int unit = PAIR2 ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;     // which persistent worker this CTA belongs to
  const int units = PAIR2 ? (int)(gridDim.x >> 1) : (int)gridDim.x;

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
        mbar_init(tempty0 + 8 * b, PAIR2 ? 2 * EPI_WARPS : EPI_WARPS);  // one arrival per epilogue warp (of both CTAs)
      }
      mbar_init(bres, 1);
      mbar_init(mbar_m, 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    if (PAIR2) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(TMEM_COLS)
                   : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(TMEM_COLS)
                   : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (PAIR2) cluster_sync_all();  // the peer's barriers exist before anything arrives on them
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = warp_uniform(*tmem_slot_ptr);
  const Pipe pipe{a_smem, b_smem, b_stage_bytes, full0, empty0, tfull0, tempty0, stages, b_mn};

  if (warp == 0) {
    if (lane == 0) {
      if (PAIR2) {
        // both CTAs load; the bytes of both land on the leader's barrier, which the leader arms for the pair
        const uint32_t per_cta = (uint32_t)A_STAGE_BYTES + (b_resident ? 0u : b_stage_bytes);
        if (b_resident && unit < total) {
          if (crank == 0) mbar_expect_tx(bres, 2u * (uint32_t)num_kb * b_stage_bytes);
          const uint32_t lb = mapa_rank(bres, 0);
          const int nb0 = (unit % tiles_n) * block_n + (int)crank * (block_n / 2);
          for (int kb = 0; kb < num_kb; ++kb) tma_load_2d_pair(b_smem + kb * b_stage_bytes, &tma_b, kb * BK, nb0, lb);
        }
        uint32_t it = 0;
        for (int t = unit; t < total; t += units) {
          const int m0 = (t / tiles_n) * TILE_M + (int)crank * BM;
          const int n0 = (t % tiles_n) * block_n + (int)crank * (block_n / 2);
          for (int kb = 0; kb < num_kb; ++kb, ++it) {
            const uint32_t s_ = it % (uint32_t)stages, ph = (it / (uint32_t)stages) & 1u;
            mbar_wait(empty0 + 8 * s_, ph ^ 1u);
            if (crank == 0) mbar_expect_tx(full0 + 8 * s_, 2u * per_cta);
            const uint32_t lbar = mapa_rank(full0 + 8 * s_, 0);
            tma_load_2d_pair(a_smem + s_ * A_STAGE_BYTES, &tma_a, kb * BK, m0, lbar);
            if (!b_resident) tma_load_2d_pair(b_smem + s_ * b_stage_bytes, &tma_b, kb * BK, n0, lbar);
          }
        }
      } else {
        producer_loop(&tma_a, &tma_b, pipe, total, tiles_n, block_n, num_kb);
      }
    }
  } else if (warp == 1) {
    if (PAIR2) {
      if (crank == 0) {  // the leader issues for the pair (whole warp, one elected lane)
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(block_n >> 3) << 17) | ((uint32_t)(TILE_M >> 4) << 24);
        const uint64_t d0 = umma_desc(0);
        uint32_t it = 0, i = 0;
        if (b_resident && unit < total) mbar_wait(bres, 0);
        for (int t = unit; t < total; t += units, ++i) {
          const uint32_t buf = i & 1u, aph = (i >> 1) & 1u;
          mbar_wait(tempty0 + 8 * buf, aph ^ 1u);  // both CTAs' epilogues have drained this accumulator
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t tmem_d = tmem_base + buf * MAX_BN;
          for (int kb = 0; kb < num_kb; ++kb, ++it) {
            const uint32_t s_ = it % (uint32_t)stages, ph = (it / (uint32_t)stages) & 1u;
            mbar_wait(full0 + 8 * s_, ph);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint64_t da = d0 + (uint64_t)((a_smem + s_ * A_STAGE_BYTES) >> 4);
            const uint64_t db = d0 + (uint64_t)((b_smem + (uint32_t)(b_resident ? kb : (int)s_) * b_stage_bytes) >> 4);
#pragma unroll
            for (int k = 0; k < BK / UK; ++k)
              umma_f16_pair_elect(tmem_d, da + (uint64_t)(k * UK * 2 >> 4), db + (uint64_t)(k * UK * 2 >> 4), idesc,
                                  (uint32_t)((kb | k) != 0));
            umma_commit_pair_elect(empty0 + 8 * s_);
          }
          umma_commit_pair_elect(tfull0 + 8 * buf);
        }
      }
    } else {
      mma_loop(pipe, tmem_base, total, block_n, num_kb);  // whole warp, one elected lane issues
    }
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
    int bias_n0 = -1;
    for (int t = unit; t < total; t += units, ++i) {
      const int m0 = (t / tiles_n) * TILE_M + (int)crank * BM, n0 = (t % tiles_n) * block_n;
      const uint32_t buf = i & 1u, aph = (i >> 1) & 1u;
      // bias slice of this column tile -> shared memory; reloaded only when the column tile changes (never, when the
      // persistent stride is a multiple of the column-tile count): a global load + a second barrier per tile otherwise
      if (TMA_STORE && et == 0) {
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        if (mask != nullptr) {
          // the ReLU mask of the tile goes INTO the staging tile by TMA (a lane-per-row global read costs one L1 tag
          // cycle per lane and instruction: ~2 us per tile); each thread later reads its elements there and writes
          // the masked result over them
          mbar_expect_tx(mbar_m, (uint32_t)BM * block_n * 2);
          tma_load_2d(stage_out, &tma_m, n0, m0, mbar_m);
        }
      }
      asm volatile("bar.sync 1, %0;" ::"n"(32 * EPI_WARPS) : "memory");  // previous tile's bias / staging are free
      if (n0 != bias_n0) {
        for (int j = et; j < block_n; j += 32 * EPI_WARPS) bias_s[j] = (bias != nullptr && n0 + j < N) ? __ldg(bias + n0 + j) : 0.f;
        asm volatile("bar.sync 1, %0;" ::"n"(32 * EPI_WARPS) : "memory");
        bias_n0 = n0;
      }
      const int row = m0 + q * 32 + lane;
      if (!TMA_STORE && mask != nullptr && row < M) {
        const char* mrow = reinterpret_cast<const char*>(mask + (size_t)row * ldo + n0 + c_begin);
        const int nb = ((n0 + c_end < N ? c_end : N - n0) - c_begin) * 2;
        for (int b = 0; b < nb; b += 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(mrow + b));
      }
      // ReLU mask, one chunk (32 columns) ahead: one lane = one row, 16-byte loads (a full 32-byte sector per 16
      // columns).  The first chunk is requested before the accumulator is waited for, chunk i + 1 before chunk i is
      // processed, so a load has a whole chunk's worth of work (or the accumulator wait) to arrive in.
      uint4 mnext[4] = {};
      auto load_mask = [&](int c0_) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int cc_ = c0_ + 16 * h, col = n0 + cc_;
          if (!TMA_STORE && mask != nullptr && row < M && cc_ < c_end && col + 16 <= N && (((size_t)row * ldo + col) & 7) == 0) {
            const uint4* mp = reinterpret_cast<const uint4*>(mask + (size_t)row * ldo + col);
            mnext[2 * h] = __ldg(mp);
            mnext[2 * h + 1] = __ldg(mp + 1);
          }
        }
      };
      load_mask(c_begin);
      mbar_wait(tfull0 + 8 * buf, aph);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint32_t taddr_row = tmem_base + buf * MAX_BN + ((uint32_t)(q * 32) << 16);
      if (TMA_STORE && mask != nullptr) mbar_wait(mbar_m, i & 1u);  // one mask tile per tile of this CTA
      for (int c0 = c_begin; c0 < ((dbg & 4) ? c_begin : c_end); c0 += 32) {
        // two 16-column loads in flight, one wait
        uint32_t r[32];
        const bool second = c0 + 16 < c_end;
        uint4 mvec[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) mvec[j] = mnext[j];
        load_mask(c0 + 32);
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
          if (mask != nullptr && (TMA_STORE || row < M)) {
            const __nv_bfloat16* mk = mask + (size_t)row * ldo + col;
            if (TMA_STORE || (col + 16 <= N && (((size_t)row * ldo + col) & 7) == 0)) {
              uint4 m0v = mvec[2 * h], m1v = mvec[2 * h + 1];
              if (TMA_STORE) {  // rows / columns beyond M, N were zero-filled by the load: masked out, clipped on store
                const uint32_t src = stage_out + (uint32_t)((q * 32 + lane) * block_n + cc) * 2;
                asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(m0v.x), "=r"(m0v.y), "=r"(m0v.z), "=r"(m0v.w) : "r"(src));
                asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(m1v.x), "=r"(m1v.y), "=r"(m1v.z), "=r"(m1v.w) : "r"(src + 16));
              }
              const uint32_t mw[8] = {m0v.x, m0v.y, m0v.z, m0v.w, m1v.x, m1v.y, m1v.z, m1v.w};
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                // bf16 > 0  <=>  bits in [0x0001, 0x7f80] (sign clear, not zero, not NaN)
                const uint32_t lo = mw[j] & 0xffffu, hi = mw[j] >> 16;
                if (lo - 1u >= 0x7f80u) v[2 * j] = 0.f;
                if (hi - 1u >= 0x7f80u) v[2 * j + 1] = 0.f;
              }
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j)
                if (col + j < N && !(__bfloat162float(mk[j]) > 0.f)) v[j] = 0.f;
            }
          }
          uint4 p0, p1;
          p0.x = pdj_f2_to_bf2(v[0], v[1]); p0.y = pdj_f2_to_bf2(v[2], v[3]);
          p0.z = pdj_f2_to_bf2(v[4], v[5]); p0.w = pdj_f2_to_bf2(v[6], v[7]);
          p1.x = pdj_f2_to_bf2(v[8], v[9]); p1.y = pdj_f2_to_bf2(v[10], v[11]);
          p1.z = pdj_f2_to_bf2(v[12], v[13]); p1.w = pdj_f2_to_bf2(v[14], v[15]);
          if (TMA_STORE && (dbg & 2)) {
          } else if (TMA_STORE) {
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
      if (lane == 0) {
        if (PAIR2) mbar_arrive_cluster(mapa_rank(tempty0 + 8 * buf, 0));  // the leader's MMA warp waits for both CTAs
        else mbar_arrive(tempty0 + 8 * buf);
      }
      if (TMA_STORE) {
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> visible to the TMA engine
        asm volatile("bar.sync 1, %0;" ::"n"(32 * EPI_WARPS) : "memory");
        if (et == 0) {
          if (!(dbg & 1)) tma_store_2d(&tma_o, stage_out, n0, m0);  // rows / columns beyond M, N are clipped by the tensor map
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
      }
    }
    if (TMA_STORE && et == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (PAIR2) cluster_sync_all();  // neither CTA leaves (or frees tensor memory) while the pair's MMAs may still touch it
  if (warp == 1) {
    if (PAIR2) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
    else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
  }
}

bool g_attr_set = false;

int ensure_init() {
  if (g_attr_set) return 0;
  PDJ_CUDA_CHECK(cudaFuncSetAttribute(linear_tcgen05_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
  PDJ_CUDA_CHECK(cudaFuncSetAttribute(linear_tcgen05_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
  PDJ_CUDA_CHECK(cudaFuncSetAttribute(linear_tcgen05_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
  PDJ_CUDA_CHECK(cudaFuncSetAttribute(linear_tcgen05_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
  g_attr_set = true;
  return 0;
}

}  // namespace

extern "C" {

int pdj_linear_block_n(int N) {
  // smallest multiple of 16 (<= 256) that covers N in the fewest column tiles
  const int tiles = (N + MAX_BN - 1) / MAX_BN;
  int bn = ((N + tiles - 1) / tiles + 15) / 16 * 16;
  if (bn < 16) bn = 16;
  if (bn > MAX_BN) bn = MAX_BN;
  return bn;
}

static int launch_linear(const void* a, const void* b, void* out, const float* bias, const void* mask, int M, int N, int K,
                         int lda, int ldb, int ldo, int relu, int b_kn, cudaStream_t st) {
  if (M < 1 || N < 1 || K < 1 || lda % 8 != 0 || ldb % 8 != 0 || ((uintptr_t)a & 15) || ((uintptr_t)b & 15)) {
    pdj_set_error("pdj_linear_bf16: need M,N,K >= 1, 16-byte aligned operands and row pitches that are multiples of 8 elements "
                  "(lda=%d ldb=%d)", lda, ldb);
    return -1;
  }
  const int bn = pdj_linear_block_n(N);
  if (const char* d = getenv("PDJ_LINEAR_DBG")) relu |= atoi(d) << 8;  // diagnostics only (profiles/r2/linear_pair.md)
  // TMA store needs a 16-byte aligned output pitch; otherwise the epilogue stores directly.
  // Measured on B200 (profiles/gemm_bench.py): staging + TMA store wins for narrow outputs (it frees
  // the epilogue warps early), direct 32-byte stores win once the tile is 256 wide, where the staging
  // buffer would cost a pipeline stage.  PDJ_GEMM_DIRECT_STORE / PDJ_GEMM_TMA_STORE override.
  static const bool force_direct = getenv("PDJ_GEMM_DIRECT_STORE") != nullptr;
  static const bool force_tma = getenv("PDJ_GEMM_TMA_STORE") != nullptr;
  const bool tma_ok = (ldo % 8 == 0) && (((uintptr_t)out & 15) == 0);
  const bool tma_store = tma_ok && !force_direct && (force_tma || bn < MAX_BN);
  // CTA pairs (tcgen05 cta_group::2): each byte of B leaves L2 once per 256 rows.  PDJ_LINEAR_PAIR=0/1 overrides.
  const char* pair_env = getenv("PDJ_LINEAR_PAIR");
  const int sms = pdj_tc_sm_count();
  // Measured (profiles/r2/linear_pair.md): +10..15 % where the B stream matters (8192 x 4096 x 4096: 1.55 -> 1.72 PFLOP/s,
  // above the library's 1.56), a loss on the skinny Wide&Deep layers, whose limit is the A stream and the epilogue.
  const bool pair_auto = K >= 1024 && N >= 256 && M >= 4 * BM;
  const bool pair = (pair_env ? atoi(pair_env) != 0 : pair_auto) && bn % 16 == 0 && M >= 2 * BM && sms >= 2 && !b_kn;
  CUtensorMap ma, mb, mo;
  if (int r = pdj_tc_make_map(&ma, a, M, K, lda, BK, BM, true)) return r;
  const int kn_mode = !b_kn ? 0 : ((N % 64 == 0 && bn % 64 == 0) ? 2 : 1);
  if (kn_mode == 2) {  // b is [K][N]: every 64-column group of a stage in one 4-D box
    if (int r = pdj_tc_make_map_kn(&mb, b, K, N, ldb, BK, bn / 64)) return r;
  } else if (kn_mode == 1) {  // boxes of 64 columns x BK rows, one per group
    if (int r = pdj_tc_make_map(&mb, b, K, N, ldb, 64, BK, true)) return r;
  } else if (int r = pdj_tc_make_map(&mb, b, N, K, ldb, BK, pair ? bn / 2 : bn, true)) {
    return r;
  }
  CUtensorMap mm;
  if (tma_store) {
    if (int r = pdj_tc_make_map(&mo, out, M, N, ldo, bn, BM, false)) return r;
  } else {
    mo = ma;
  }
  mm = mo;
  if (tma_store && mask != nullptr) {  // same geometry as the output (checked by the caller), loaded into the staging tile
    if (((uintptr_t)mask & 15) != 0) {
      pdj_set_error("pdj_linear_bf16: the mask must be 16-byte aligned");
      return -1;
    }
    if (int r = pdj_tc_make_map(&mm, mask, M, N, ldo, bn, BM, false)) return r;
  }
  if (int r = ensure_init()) return r;
  const int tile_m = pair ? 2 * BM : BM;
  const int tiles_n_h = (N + bn - 1) / bn;
  const int total = ((M + tile_m - 1) / tile_m) * tiles_n_h;
  const int fixed = 1024 /*alignment*/ + 256 /*barriers*/ + MAX_BN * 4 /*bias slice*/ + (tma_store ? BM * bn * 2 : 0);
  // CTA pairs that always work on the same column tile keep their half of the whole B panel in shared memory
  const int pair_grid = 2 * total < sms ? 2 * total : (sms & ~1);
  const int num_kb_h = (K + BK - 1) / BK;
  const int panel = num_kb_h * (bn / 2) * BK * 2;
  const char* res_env = getenv("PDJ_LINEAR_BRES");
  const bool resident = pair && (res_env ? atoi(res_env) != 0 : true) && (pair_grid / 2) % tiles_n_h == 0 &&
                        fixed + panel + 3 * A_STAGE_BYTES <= SMEM_LIMIT;
  const int stage_bytes = resident ? A_STAGE_BYTES : A_STAGE_BYTES + (pair ? (bn / 2) * BK * 2 : (int)b_stage_bytes_for(bn, b_kn));
  int stages = (SMEM_LIMIT - fixed - (resident ? panel : 0)) / stage_bytes;
  if (stages > MAX_STAGES) stages = MAX_STAGES;
  if (stages < 2) {
    pdj_set_error("pdj_linear_bf16: tile does not fit in shared memory (block_n=%d)", bn);
    return -1;
  }
  const int smem = fixed + (resident ? panel : 0) + stages * stage_bytes;
  if (pair) {
    const int grid = pair_grid;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(THREADS);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    __nv_bfloat16* outp = (__nv_bfloat16*)out;
    const __nv_bfloat16* maskp = (const __nv_bfloat16*)mask;
    if (tma_store)
      PDJ_CUDA_CHECK(cudaLaunchKernelEx(&cfg, linear_tcgen05_kernel<true, true>, ma, mb, mo, mm, outp, bias, maskp, M, N, K, ldo, bn,
                                        stages, relu, resident ? 1 : 0));
    else
      PDJ_CUDA_CHECK(cudaLaunchKernelEx(&cfg, linear_tcgen05_kernel<false, true>, ma, mb, mo, mm, outp, bias, maskp, M, N, K, ldo, bn,
                                        stages, relu, resident ? 1 : 0));
    return 0;
  }
  const int grid = total < sms ? total : sms;
  if (tma_store)
    linear_tcgen05_kernel<true, false><<<grid, THREADS, smem, st>>>(ma, mb, mo, mm, (__nv_bfloat16*)out, bias,
                                                                   (const __nv_bfloat16*)mask, M, N, K, ldo, bn, stages, relu, kn_mode << 1);
  else
    linear_tcgen05_kernel<false, false><<<grid, THREADS, smem, st>>>(ma, mb, mo, mm, (__nv_bfloat16*)out, bias,
                                                                    (const __nv_bfloat16*)mask, M, N, K, ldo, bn, stages, relu, kn_mode << 1);
  PDJ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int pdj_linear_bf16(const void* a, const void* b, void* out, const float* bias, const void* mask, int M, int N, int K,
                    int lda, int ldb, int ldo, int relu, cudaStream_t st) {
  return launch_linear(a, b, out, bias, mask, M, N, K, lda, ldb, ldo, relu, 0, st);
}

// The same with B given as b_kn[K][N] (row pitch ldb >= N elements): out = epilogue(a . b_kn).  A weight [out, in] used
// for the input gradient dX = dZ . W is exactly that, so no transposed copy is made (MN-major B operand).
int pdj_linear_bf16_kn(const void* a, const void* b_kn, void* out, const float* bias, const void* mask, int M, int N, int K,
                       int lda, int ldb, int ldo, int relu, cudaStream_t st) {
  return launch_linear(a, b_kn, out, bias, mask, M, N, K, lda, ldb, ldo, relu, 1, st);
}

}  // extern "C"
