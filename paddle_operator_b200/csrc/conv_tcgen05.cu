// tcgen05 convolution kernels of the ResNet path (sm_100a only), on the persistent TMA + tcgen05 pipeline of
// tcgen05_common.cuh:
//   * conv1x1_tcgen05_kernel - NHWC 1x1 convolution as a GEMM with a store-bandwidth-oriented epilogue
//     (BN statistics / addend / plain), and in WIN mode the 4x4 space-to-depth stem through overlapping
//     TMA windows;
//   * stem_wgrad_tcgen05_kernel - the stem's weight gradient with MN-major operands.
// Reference parity: none (the reference has no compute code); the ResNet workload of
// deploy/examples/resnet.yaml:13-19.
#include <cuda.h>
#include <stdlib.h>

#include "common.cuh"
#include "tcgen05_common.cuh"
#include "../../include/pdjob_b200.h"

using namespace pdj_tc;

namespace {

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
//
// IM2COL (3x3 and strided convolutions): the A operand of k-block (tap, channel block) is one im2col-mode TMA load
// (cuTensorMapEncodeIm2col) of 128 consecutive output pixels x 64 input channels of that filter tap - the TMA unit
// walks the pixels across image rows and images and zero-fills the padding, so M tiles are dense whatever the image
// size.  K runs over (ky, kx, c) exactly as the channels-last weight [Cout][kh][kw][Cin] is laid out.  The input
// gradient of a stride-1 convolution is the same kernel on dY with the flipped, transposed weight.
constexpr int SRC_ROWS = 0, SRC_WIN = 1, SRC_IM2COL = 2;
struct ConvGeom {
  int Q, PQ;        // output width, output pixels per image
  int stride, pad;  // traversal stride and symmetric padding
  int ksize;        // filter is ksize x ksize
  int cblocks;      // Cin / 64
  int b_mn;         // SRC_ROWS only: B is given as [K][N] (a weight used transposed) - MN-major operand, no transposed copy
};

template <int MODE, int SRC>
__global__ void __launch_bounds__(THREADS, 1)
conv1x1_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                       const __grid_constant__ CUtensorMap tma_o, const __grid_constant__ CUtensorMap tma_c,
                       float* __restrict__ partial, int M, int N, int K, int block_n, int stages, int win_ho,
                       int win_hi, const ConvGeom geom) {
  constexpr bool STATS = MODE == MODE_STATS;
  constexpr bool ADD = MODE == MODE_ADD;
  constexpr bool WIN = SRC == SRC_WIN;
  constexpr bool IM2COL = SRC == SRC_IM2COL;
  static_assert(!(WIN && ADD), "the stem has no addend mode");
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t b_stage_bytes = b_stage_bytes_for(block_n, (SRC == SRC_ROWS) ? geom.b_mn : 0);
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

  const int warp = (int)warp_uniform(threadIdx.x >> 5), lane = threadIdx.x & 31;
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
  const uint32_t tmem_base = warp_uniform(*tmem_slot_ptr);
  const Pipe pipe{a_smem, b_smem, b_stage_bytes, full0, empty0, tfull0, tempty0, stages, (SRC == SRC_ROWS) ? geom.b_mn : 0};

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
      } else if (IM2COL) {
        const uint32_t stage_bytes = (uint32_t)A_STAGE_BYTES + b_stage_bytes;
        uint32_t it = 0;
        for (int t = blockIdx.x; t < total; t += gridDim.x) {
          const int m0 = (t / tiles_n) * BM, n0 = (t % tiles_n) * block_n;
          // first output pixel of the tile -> (image, row, column); the TMA coordinates name the input position of
          // the top-left filter tap for that pixel
          const int img = m0 / geom.PQ, rem = m0 % geom.PQ;
          const int h0 = (rem / geom.Q) * geom.stride - geom.pad, w0 = (rem % geom.Q) * geom.stride - geom.pad;
          int kb = 0;
          for (int ky = 0; ky < geom.ksize; ++ky)
            for (int kx = 0; kx < geom.ksize; ++kx)
              for (int cb = 0; cb < geom.cblocks; ++cb, ++kb, ++it) {
                const uint32_t s = it % (uint32_t)stages, ph = (it / (uint32_t)stages) & 1u;
                mbar_wait(empty0 + 8 * s, ph ^ 1u);
                mbar_expect_tx(full0 + 8 * s, stage_bytes);
                tma_load_im2col_4d(a_smem + s * A_STAGE_BYTES, &tma_a, cb * BK, w0, h0, img, (uint16_t)kx, (uint16_t)ky,
                                   full0 + 8 * s);
                tma_load_2d(b_smem + s * b_stage_bytes, &tma_b, kb * BK, n0, full0 + 8 * s);
              }
        }
      } else {
        producer_loop(&tma_a, &tma_b, pipe, total, tiles_n, block_n, num_kb);
      }
    }
  } else if (warp == 1) {
    {  // the whole warp runs the issue loops; one elected lane issues (umma_*_elect)
      if (WIN) {
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(block_n >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
        const uint32_t S = (uint32_t)stages;
        const uint64_t d0 = umma_desc(0);
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
              const uint64_t da = d0 + (uint64_t)((a_smem + slot * A_STAGE_BYTES) >> 4);
              const uint64_t db = d0 + (uint64_t)((b_smem + kb * b_stage_bytes) >> 4);
#pragma unroll
              for (int k = 0; k < BK / UK; ++k)
                umma_f16_elect(tmem_d, da + (uint64_t)(k * UK * 2 >> 4), db + (uint64_t)(k * UK * 2 >> 4), idesc, (uint32_t)((kb | k) != 0));
            }
            umma_commit_elect(empty0 + 8 * ((ix + j) % S));  // input row y0 + j has served its last output row
            if (j == rows - 1)
              for (int r = 1; r <= 3; ++r) umma_commit_elect(empty0 + 8 * ((ix + j + r) % S));
            umma_commit_elect(tfull0 + 8 * buf);
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
  const int warp = (int)warp_uniform(threadIdx.x >> 5), lane = threadIdx.x & 31;
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
  const uint32_t tmem_base = warp_uniform(*tmem_slot_ptr);

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
    {  // the whole warp runs the issue loop; one elected lane issues
      // D = f32, A = B = bf16, both MN-major (bits 15, 16), N = 64, M = 64
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((64u >> 3) << 17) | ((64u >> 4) << 24);
      const int ksteps = kp / 16;
      const uint64_t d0 = umma_desc_mn(0);
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
          const uint64_t da0 = d0 + (uint64_t)((dy_smem + sy * tile) >> 4);
#pragma unroll
          for (int jy = 0; jy < 4; ++jy) {
            const uint32_t xs = (ix + j + jy) % WG_RX;
            const uint64_t db0 = d0 + (uint64_t)((x_smem + xs * tile) >> 4);
            for (int k = 0; k < ksteps; ++k)
              umma_f16_elect(tmem_base + (uint32_t)jy * 64u, da0 + (uint64_t)(k * 128), db0 + (uint64_t)(k * 128), idesc,
                             (uint32_t)((iy | (uint32_t)k) != 0u));
          }
          umma_commit_elect(yempty + 8 * sy);
          umma_commit_elect(xempty + 8 * ((ix + j) % WG_RX));       // input row y0 + j has served its last output row
          if (j == rows - 1)
            for (int r = 1; r <= 3; ++r) umma_commit_elect(xempty + 8 * ((ix + j + r) % WG_RX));  // bottom halo rows
        }
        ix += rows + 3;
      }
      umma_commit_elect(done);
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

bool g_attr_set = false;
int g_sms = 0;

int ensure_init() {
  if (g_attr_set) return 0;
  PDJ_CUDA_CHECK(cudaFuncSetAttribute(conv1x1_tcgen05_kernel<MODE_PLAIN, SRC_ROWS>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
  PDJ_CUDA_CHECK(cudaFuncSetAttribute(conv1x1_tcgen05_kernel<MODE_STATS, SRC_ROWS>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
  PDJ_CUDA_CHECK(cudaFuncSetAttribute(conv1x1_tcgen05_kernel<MODE_ADD, SRC_ROWS>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
  PDJ_CUDA_CHECK(cudaFuncSetAttribute(conv1x1_tcgen05_kernel<MODE_PLAIN, SRC_WIN>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
  PDJ_CUDA_CHECK(cudaFuncSetAttribute(conv1x1_tcgen05_kernel<MODE_STATS, SRC_WIN>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
  PDJ_CUDA_CHECK(cudaFuncSetAttribute(stem_wgrad_tcgen05_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
  PDJ_CUDA_CHECK(cudaFuncSetAttribute(conv1x1_tcgen05_kernel<MODE_PLAIN, SRC_IM2COL>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
  PDJ_CUDA_CHECK(cudaFuncSetAttribute(conv1x1_tcgen05_kernel<MODE_STATS, SRC_IM2COL>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
  g_sms = pdj_tc_sm_count();
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

int sm_count_cached() { return pdj_tc_sm_count(); }

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

extern "C" {

int pdj_conv1x1_partial_rows(int M, int N) {
  if (M < 1 || N < BOX_N || N % BOX_N != 0) return 0;
  const int sms = sm_count_cached();
  const int bn = conv_block_n(M, N, sms);
  return 2 * conv_grid(M, N, bn, sms) / (N / bn);
}

static int launch_conv1x1(const void* a, const void* b, const void* addend, void* out, float* partial, int M, int N, int K,
                          int lda, int ldb, int ldc, int ldo, cudaStream_t st, int b_kn = 0) {
  const int bn = (M >= 1 && N >= BOX_N && N % BOX_N == 0) ? conv_block_n(M, N, sm_count_cached()) : 0;
  if (M < 1 || K < 1 || bn == 0 || lda % 8 != 0 || ldb % 8 != 0 || ldo % 8 != 0 || ((uintptr_t)a & 15) ||
      ((uintptr_t)b & 15) || ((uintptr_t)out & 15) || (addend != nullptr && (ldc % 8 != 0 || ((uintptr_t)addend & 15)))) {
    pdj_set_error("pdj_conv1x1: need M,K >= 1, N a multiple of 64, 16-byte aligned operands and row pitches that are "
                  "multiples of 8 elements (N=%d lda=%d ldb=%d ldc=%d ldo=%d)", N, lda, ldb, ldc, ldo);
    return -1;
  }
  if (int r = ensure_init()) return r;
  CUtensorMap ma, mb, mo, mc;
  if (int r = pdj_tc_make_map(&ma, a, M, K, lda, BK, BM, true)) return r;
  if (b_kn) {  // b is [K][N], N a multiple of 64: one 4-D box per stage
    if (int r = pdj_tc_make_map_kn(&mb, b, K, N, ldb, BK, bn / 64)) return r;
  } else if (int r = pdj_tc_make_map(&mb, b, N, K, ldb, BK, bn, true)) {
    return r;
  }
  if (int r = pdj_tc_make_map(&mo, out, M, N, ldo, BOX_N, BM, true)) return r;
  if (addend != nullptr) {
    if (int r = pdj_tc_make_map(&mc, addend, M, N, ldc, BOX_N, BM, true)) return r;
  } else {
    mc = mo;
  }
  const int stage_bytes = A_STAGE_BYTES + (int)b_stage_bytes_for(bn, b_kn);
  const int fixed = 1024 /*alignment*/ + 256 /*barriers*/ + SW_STAGING_BYTES + SW_SCRATCH_BYTES;
  int stages = (SMEM_LIMIT - fixed) / stage_bytes;
  ConvGeom geom{};
  geom.b_mn = b_kn ? 2 : 0;
  if (stages > MAX_STAGES) stages = MAX_STAGES;
  if (stages < 2) {
    pdj_set_error("pdj_conv1x1: tile does not fit in shared memory (block_n=%d)", bn);
    return -1;
  }
  const int smem = fixed + stages * stage_bytes;
  const int grid = conv_grid(M, N, bn, g_sms);
  if (addend != nullptr)
    conv1x1_tcgen05_kernel<MODE_ADD, SRC_ROWS><<<grid, THREADS, smem, st>>>(ma, mb, mo, mc, nullptr, M, N, K, bn, stages, 0, 0, geom);
  else if (partial != nullptr)
    conv1x1_tcgen05_kernel<MODE_STATS, SRC_ROWS><<<grid, THREADS, smem, st>>>(ma, mb, mo, mc, partial, M, N, K, bn, stages, 0, 0, geom);
  else
    conv1x1_tcgen05_kernel<MODE_PLAIN, SRC_ROWS><<<grid, THREADS, smem, st>>>(ma, mb, mo, mc, nullptr, M, N, K, bn, stages, 0, 0, geom);
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
  if (int r = pdj_tc_make_map3(&ma, x, 64, (uint64_t)Wo, (uint64_t)N * Hi, (uint64_t)C * 2, (uint64_t)Wi * C * 2, 64, BM)) return r;
  if (int r = pdj_tc_make_map(&mb, w, O, Kc, Kc, BK, O, true)) return r;
  if (int r = pdj_tc_make_map3(&mo, out, 64, (uint64_t)Wo, (uint64_t)N * Ho, (uint64_t)O * 2, (uint64_t)Wo * O * 2, 64, BM)) return r;
  const int bn = O;
  const int stages = 7;  // input-row ring: four rows in use by a tile, three being fetched
  const int smem = 1024 + 256 + SW_STAGING_BYTES + SW_SCRATCH_BYTES + stages * A_STAGE_BYTES + 4 * bn * BK * 2;
  const int rows = N * Ho;
  const int units = stem_units(N, Ho);
  const int grid = units < g_sms ? units : g_sms;
  if (partial != nullptr)
    conv1x1_tcgen05_kernel<MODE_STATS, SRC_WIN><<<grid, THREADS, smem, st>>>(ma, mb, mo, mo, partial, rows, O, Kc, bn, stages, Ho, Hi, ConvGeom{});
  else
    conv1x1_tcgen05_kernel<MODE_PLAIN, SRC_WIN><<<grid, THREADS, smem, st>>>(ma, mb, mo, mo, nullptr, rows, O, Kc, bn, stages, Ho, Hi, ConvGeom{});
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
  if (int r = pdj_tc_make_map3(&mx, x, 64, (uint64_t)Wo, (uint64_t)N * Hi, (uint64_t)C * 2, (uint64_t)Wi * C * 2, 64, kp)) return r;
  if (int r = pdj_tc_make_map3(&my, dy, 64, (uint64_t)Wo, (uint64_t)N * Ho, (uint64_t)O * 2, (uint64_t)Wo * O * 2, 64, kp)) return r;
  const int units = N * ((Ho + WG_CHUNK - 1) / WG_CHUNK);
  const int grid = units < g_sms ? units : g_sms;
  const int smem = 1024 + 512 + (WG_RX + WG_RY) * kp * 128;
  stem_wgrad_tcgen05_kernel<<<grid, WG_THREADS, smem, st>>>(mx, my, (float*)ws, N, Ho, Hi, kp);
  PDJ_CUDA_CHECK(cudaGetLastError());
  stem_wgrad_reduce_kernel<<<64, 256, 0, st>>>((const float*)ws, grid, (__nv_bfloat16*)dw);
  PDJ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int pdj_conv_im2col_partial_rows(int N, int H, int W, int Cout, int ksize, int pad, int stride) {
  const int P = (H + 2 * pad - ksize) / stride + 1, Q = (W + 2 * pad - ksize) / stride + 1;
  return pdj_conv1x1_partial_rows(N * P * Q, Cout);
}

// out[N][P][Q][Cout] = conv2d(x[N][H][W][Cin], w[Cout][k][k][Cin]) (NHWC bf16, square filter, symmetric padding,
// P = (H + 2 pad - k) / stride + 1) as an implicit GEMM on the tcgen05 pipeline: the A operand comes from im2col-mode
// TMA loads, nothing is materialised.  Optional BN partial sums as in pdj_conv1x1_bf16
// (pdj_conv_im2col_partial_rows rows).  Cin % 64 == 0, Cout % 64 == 0.
int pdj_conv_im2col_fwd(const void* x, const void* w, void* out, float* partial, int N, int H, int W, int Cin, int Cout,
                        int ksize, int pad, int stride, cudaStream_t st) {
  const int P = (H + 2 * pad - ksize) / stride + 1, Q = (W + 2 * pad - ksize) / stride + 1;
  const long long Mll = (long long)N * P * Q;
  if (N < 1 || P < 1 || Q < 1 || Cin % BK != 0 || Cout % BOX_N != 0 || ksize < 1 || ksize > 7 || stride < 1 || stride > 8 ||
      pad < 0 || pad >= ksize || Mll >= (1ll << 31) || ((uintptr_t)x & 15) || ((uintptr_t)w & 15) || ((uintptr_t)out & 15)) {
    pdj_set_error("pdj_conv_im2col_fwd: need Cin, Cout multiples of 64, 1 <= k <= 7, pad < k, 16-byte aligned tensors "
                  "(N=%d H=%d W=%d Cin=%d Cout=%d k=%d pad=%d stride=%d)", N, H, W, Cin, Cout, ksize, pad, stride);
    return -1;
  }
  if (int r = ensure_init()) return r;
  const int M = (int)Mll, K = ksize * ksize * Cin;
  const int bn = conv_block_n(M, Cout, sm_count_cached());
  CUtensorMap ma, mb, mo;
  if (int r = pdj_tc_make_im2col4(&ma, x, Cin, W, H, N, ksize, pad, stride, BK, BM)) return r;
  if (int r = pdj_tc_make_map(&mb, w, Cout, K, K, BK, bn, true)) return r;
  if (int r = pdj_tc_make_map(&mo, out, M, Cout, Cout, BOX_N, BM, true)) return r;
  const int stage_bytes = A_STAGE_BYTES + bn * BK * 2;
  const int fixed = 1024 + 256 + SW_STAGING_BYTES + SW_SCRATCH_BYTES;
  int stages = (SMEM_LIMIT - fixed) / stage_bytes;
  if (stages > MAX_STAGES) stages = MAX_STAGES;
  const int smem = fixed + stages * stage_bytes;
  const int grid = conv_grid(M, Cout, bn, g_sms);
  const ConvGeom g{Q, P * Q, stride, pad, ksize, Cin / BK};
  if (partial != nullptr)
    conv1x1_tcgen05_kernel<MODE_STATS, SRC_IM2COL><<<grid, THREADS, smem, st>>>(ma, mb, mo, mo, partial, M, Cout, K, bn, stages, 0, 0, g);
  else
    conv1x1_tcgen05_kernel<MODE_PLAIN, SRC_IM2COL><<<grid, THREADS, smem, st>>>(ma, mb, mo, mo, nullptr, M, Cout, K, bn, stages, 0, 0, g);
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

// The two GEMMs above with B given as b_kn[K][N] (row pitch ldb >= N): the input gradient dX = (dskip +) dY . W reads the
// weight W[Cout][Cin] as it is stored - no transposed copy.  `addend` may be NULL (plain product).
int pdj_conv1x1_kn_bf16(const void* a, const void* b_kn, const void* addend, void* out, int M, int N, int K, int lda, int ldb,
                        int ldc, int ldo, cudaStream_t st) {
  return launch_conv1x1(a, b_kn, addend, out, nullptr, M, N, K, lda, ldb, ldc, ldo, st, 1);
}

}  // extern "C"
