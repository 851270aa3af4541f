// Weight gradients of NHWC convolutions on tcgen05 (sm_100a only).
//
//   dW[o][ky][kx][c] = sum over (n, y, x) of dY[n][y][x][o] * X[n][y + ky - 1][x + kx - 1][c]        (3x3, pad 1, stride 1)
//
// is a GEMM whose reduction runs over every output pixel (0.8 M at 56x56, batch 256) and whose output is small.
// Both operands are read exactly as they sit in memory: channels are contiguous, so for the MMA they are
// MN-major tiles whose rows are the reduction index (pixels).  What makes the kernel is how the nine taps are fed:
//
//   * one pixel block = R image rows of one image, linearised with a row pitch of P = W + 2 pixels.  The dY tile is
//     the TMA box {64 channels, P pixels from x = 0, R rows}: the columns x >= W are out of range and arrive as
//     zeros.  The X tile is the box {32 channels, P pixels from x = -1, R + 2 rows from y0 - 1}: the halo arrives as
//     zeros too.  With that, the input pixel of tap (ky, kx) for dY row k is X row k + ky * P + kx;
//   * so all nine taps read the SAME shared-memory X tile through matrix descriptors whose start address is moved
//     by ky * P + kx rows.  tcgen05 applies the shared-memory swizzle to absolute address bits, so a descriptor may
//     start at any row of a tile that TMA wrote (profiles/r2/desc_shift_experiment.log).  X is fetched once per
//     block instead of nine times - the difference between a kernel bound by the L2 -> SM path and one bound by
//     the tensor pipe.  The zero columns of dY cancel the products whose X row wrapped into the next image row;
//   * a CTA owns [Mo output channels] x [9 taps x 32 input channels] fp32 accumulators in tensor memory (288 of the
//     512 columns) for its whole life and walks its share of the pixel blocks (split-K across CTAs); at the end it
//     writes its partial tile and a second kernel adds the partials in a fixed order (bit-reproducible) and casts
//     to bf16 in the [Cout][3][3][Cin] layout of a channels-last weight.
//
// Reference parity: none (the reference has no compute code); the ResNet workload of
// deploy/examples/resnet.yaml:13-19.
#include <cuda.h>
#include <stdlib.h>

#include "common.cuh"
#include "tcgen05_common.cuh"
#include "../../include/pdjob_b200.h"

using namespace pdj_tc;

namespace {

constexpr int WG_THREADS = 192;  // warp 0: TMA, warp 1: MMA + TMEM owner, warps 2..5: epilogue
constexpr int WG_NC = 32;        // input channels per CTA tile (64-byte rows, 64-byte swizzle)
constexpr int WG_TAPS = 9;
constexpr int WG_COLS = WG_TAPS * WG_NC;  // accumulator columns in use

struct Wg3Cfg {
  int n_img, H, W, Cin, Cout;
  int P, R, kpad, chunks;   // row pitch, image rows per block, padded reduction length of a block, blocks per image
  int mo, groups;           // output channels per CTA (64 or 128) = groups x 64
  int tiles_c, tiles, S;    // (o, c) tiles and pixel splits; grid = tiles * S
  int stages;
  int debug;                // PDJ_WG_DEBUG: 1 = loads only (no MMAs), 2 = MMAs only (no loads); timing experiments
  uint32_t dy_group_bytes, dy_slot_bytes, x_slot_bytes, x_box_bytes, dy_box_bytes;
};

__global__ void __launch_bounds__(WG_THREADS, 1)
wgrad3x3_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_x, const __grid_constant__ CUtensorMap tma_dy,
                        float* __restrict__ partial, const Wg3Cfg c) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t slot_bytes = c.dy_slot_bytes + c.x_slot_bytes;
  const uint32_t bars = base + (uint32_t)c.stages * slot_bytes;
  const uint32_t full0 = bars, empty0 = bars + 8 * MAX_STAGES, done = empty0 + 8 * MAX_STAGES, tmem_slot = done + 8;
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - raw));
  const int warp = (int)warp_uniform(threadIdx.x >> 5), lane = threadIdx.x & 31;

  const int tile = (int)blockIdx.x % c.tiles, split = (int)blockIdx.x / c.tiles;
  const int o0 = (tile / c.tiles_c) * c.mo, c0 = (tile % c.tiles_c) * WG_NC;
  const int total_blocks = c.n_img * c.chunks;

  // Rows that TMA never writes (the tail of a block whose pixel count is not a multiple of 16, and the X rows past
  // the halo that the last taps reach) must read as zero, not as whatever the previous kernel left: 0 x NaN = NaN.
  {
    const uint32_t n16 = (uint32_t)c.stages * slot_bytes / 16u;
    for (uint32_t i = threadIdx.x; i < n16; i += WG_THREADS)
      asm volatile("st.shared.v4.b32 [%0], {%1,%1,%1,%1};" ::"r"(base + i * 16u), "r"(0u) : "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_x) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_dy) : "memory");
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < c.stages; ++s) {
        mbar_init(full0 + 8 * s, 1);
        mbar_init(empty0 + 8 * s, 1);
      }
      mbar_init(done, 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = warp_uniform(*tmem_slot_ptr);

  if (warp == 0) {
    if (lane == 0) {
      uint32_t it = 0;
      for (int b = split; b < total_blocks; b += c.S, ++it) {
        const uint32_t s = it % (uint32_t)c.stages, ph = (it / (uint32_t)c.stages) & 1u;
        const int n = b / c.chunks, y0 = (b % c.chunks) * c.R;
        const uint32_t dy_smem = base + s * slot_bytes, x_smem = dy_smem + c.dy_slot_bytes;
        mbar_wait(empty0 + 8 * s, ph ^ 1u);
        if (c.debug == 2) { mbar_arrive(full0 + 8 * s); continue; }
        mbar_expect_tx(full0 + 8 * s, c.x_box_bytes + (uint32_t)c.groups * c.dy_box_bytes);
        tma_load_4d(x_smem, &tma_x, c0, -1, y0 - 1, n, full0 + 8 * s);
        for (int g = 0; g < c.groups; ++g)
          tma_load_4d(dy_smem + (uint32_t)g * c.dy_group_bytes, &tma_dy, o0 + 64 * g, 0, y0, n, full0 + 8 * s);
      }
    }
  } else if (warp == 1) {
    {  // the whole warp runs this loop; one elected lane issues (see elect_one)
      // D = f32, A = B = bf16, both MN-major (bits 15, 16), M = mo, N = 96: one instruction covers the three taps of
      // a filter row.  Its B operand is three 32-channel column groups whose leading-dimension offset is ONE ROW
      // (64 bytes): group kx is the X tile moved by kx pixels.  Three instructions per 16 pixels instead of nine
      // N = 32 ones: a short MMA re-reads the 4 KiB A slice from shared memory every time and is bound by that.
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(3 * WG_NC >> 3) << 17) |
                             ((uint32_t)(c.mo >> 4) << 24);
      const uint64_t da0 = umma_desc_ex(0, c.dy_group_bytes, 1024, 2);
      const uint64_t db0 = umma_desc_ex(0, 64, 512, 4);
      const uint32_t row_step = (uint32_t)c.P * 4u;  // one filter row = P X rows of 64 B, in 16-byte units
      const int ksteps = c.kpad / 16;
      uint32_t it = 0;
      for (int b = split; b < total_blocks; b += c.S, ++it) {
        const uint32_t s = it % (uint32_t)c.stages, ph = (it / (uint32_t)c.stages) & 1u;
        const uint32_t dy_smem = base + s * slot_bytes, x_smem = dy_smem + c.dy_slot_bytes;
        mbar_wait(full0 + 8 * s, ph);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        uint64_t da = da0 + (uint64_t)(dy_smem >> 4);
        uint64_t db = db0 + (uint64_t)(x_smem >> 4);
        uint32_t acc = (uint32_t)(it != 0u);
        if (c.debug == 1) { if (elect_one()) mbar_arrive(empty0 + 8 * s); __syncwarp(); continue; }
        for (int ks = 0; ks < ksteps; ++ks) {
          umma_f16_elect(tmem_base, da, db, idesc, acc);
          umma_f16_elect(tmem_base + 3 * WG_NC, da, db + row_step, idesc, acc);
          umma_f16_elect(tmem_base + 6 * WG_NC, da, db + 2 * row_step, idesc, acc);
          da += 2048 >> 4;  // 16 dY rows of 128 B
          db += 1024 >> 4;  // 16 X rows of 64 B
          acc = 1u;
        }
        umma_commit_elect(empty0 + 8 * s);
      }
      umma_commit_elect(done);
    }
  } else {
    // epilogue: the CTA's [mo][9 x 32] tile -> partial[split][o][tap][c]
    const int q = warp & 3;
    mbar_wait_warp_sleepy(done, 0, 512);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const bool has_work = split < total_blocks;
    // M = 128: accumulator row = TMEM lane; M = 64: rows sit in lanes 0..15 of each 32-lane quarter.
    // tcgen05.ld hands each thread one row; written straight out that would be 32 scattered 64-byte pieces per
    // instruction (rows are 9 * Cin floats apart).  The warp instead parks its rows in the (now idle) operand ring
    // and writes them out with eight lanes per 128-byte (row, tap) segment: four full segments per instruction.
    const int rows_w = c.mo == 128 ? 32 : 16;             // rows this warp owns
    const int row0 = c.mo == 128 ? q * 32 : q * 16;
    const bool row_ok = lane < rows_w;
    constexpr int PITCH = WG_COLS + 4;                      // floats; 292 % 32 = 4: conflict-free float4 columns
    const uint32_t stg = base + (uint32_t)(warp - 2) * (32u * PITCH * 4u);
    for (int col = 0; col < WG_COLS; col += 16) {
      uint32_t r[16];
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
          : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
            "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
          : "r"(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)col)
          : "memory");
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      if (row_ok) {
#pragma unroll
        for (int v = 0; v < 4; ++v)
          asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(stg + (uint32_t)(lane * PITCH + col + 4 * v) * 4u),
                       "r"(has_work ? r[4 * v] : 0u), "r"(has_work ? r[4 * v + 1] : 0u), "r"(has_work ? r[4 * v + 2] : 0u),
                       "r"(has_work ? r[4 * v + 3] : 0u)
                       : "memory");
      }
    }
    __syncwarp();
    const int rsub = lane >> 3, j = lane & 7;
    for (int rr = 0; rr < rows_w; rr += 4) {
      const int rl = rr + rsub;
      float* drow = partial + (((size_t)split * c.Cout + (size_t)(o0 + row0 + rl)) * WG_TAPS) * c.Cin + c0 + 4 * j;
#pragma unroll
      for (int t = 0; t < WG_TAPS; ++t) {
        float4 v;
        asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                     : "r"(stg + (uint32_t)(rl * PITCH + t * WG_NC + 4 * j) * 4u));
        *reinterpret_cast<float4*>(drow + (size_t)t * c.Cin) = v;
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
  }
}

// out[i] = bf16(sum over parts of partial[p][i]), i < n (n % 4 == 0), in a fixed order.  J lanes share one float4
// column and take the parts p = j, j + J, ...; the lanes are then added in order.  J = 8 for many parts (few
// columns: parallelism comes from the parts), J = 1 for few parts.
template <int J>
__global__ void __launch_bounds__(256)
wgrad_reduce_kernel(const float* __restrict__ partial, int nparts, size_t n, __nv_bfloat16* __restrict__ out) {
  constexpr int COLS = 256 / J;
  __shared__ float4 sm[J][COLS];
  const int col = threadIdx.x % COLS, j = threadIdx.x / COLS;
  const size_t v = (size_t)blockIdx.x * COLS + col;  // float4 index
  float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
  if (v * 4 < n) {
    int p = j;
    for (; p + J < nparts; p += 2 * J) {
      const float4 x = *reinterpret_cast<const float4*>(partial + (size_t)p * n + v * 4);
      const float4 y = *reinterpret_cast<const float4*>(partial + (size_t)(p + J) * n + v * 4);
      a.x += x.x; a.y += x.y; a.z += x.z; a.w += x.w;
      b.x += y.x; b.y += y.y; b.z += y.z; b.w += y.w;
    }
    if (p < nparts) {
      const float4 x = *reinterpret_cast<const float4*>(partial + (size_t)p * n + v * 4);
      a.x += x.x; a.y += x.y; a.z += x.z; a.w += x.w;
    }
    a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
  }
  if (J > 1) {
    sm[j][col] = a;
    __syncthreads();
    if (j != 0) return;
#pragma unroll
    for (int k = 1; k < J; ++k) {
      const float4 x = sm[k][col];
      a.x += x.x; a.y += x.y; a.z += x.z; a.w += x.w;
    }
  }
  if (v * 4 < n) {
    uint2 o;
    o.x = pdj_f2_to_bf2(a.x, a.y);
    o.y = pdj_f2_to_bf2(a.z, a.w);
    *reinterpret_cast<uint2*>(out + v * 4) = o;
  }
}

void launch_reduce(const float* partial, int nparts, size_t n, __nv_bfloat16* out, cudaStream_t st) {
  if (nparts >= 16)
    wgrad_reduce_kernel<8><<<(unsigned)((n / 4 + 31) / 32), 256, 0, st>>>(partial, nparts, n, out);
  else
    wgrad_reduce_kernel<1><<<(unsigned)((n / 4 + 255) / 256), 256, 0, st>>>(partial, nparts, n, out);
}

// ---------------------------------------------------------------------------------------------------------------
// dW[o][c] = sum over rows r of dY[r][o] * X[r][c]: the weight gradient of a 1x1 convolution / a linear layer.  Both
// operands are row-major [rows][channels], i.e. MN-major for the MMA with the reduction running over tile rows.  A CTA
// owns [128 o] x [NC <= 256 c] fp32 accumulators for its whole life (one N = NC instruction per 16 rows: operand reads
// 4 KiB + NC * 32 B per instruction stay under the tensor pipe's time from NC = 128 up), walks its share of the
// 128-row blocks (split-K across CTAs) and writes a partial tile; wgrad_reduce_kernel adds them in a fixed order.
// Out-of-range rows / channels of a box arrive as zeros, so no shape needs padding.
struct WgMnCfg {
  int M, No, Nc;            // rows (reduction), output channels (dY columns), input channels (X columns)
  int nc_tile, groups_c;    // channels per CTA tile (multiple of 64, <= 256) = groups_c x 64
  int tiles_c, tiles, S, stages, blocks;
  uint32_t slot_bytes;
  int with_db;              // also produce the column sums of dY (the bias gradient) with one extra N = 16 MMA per step
};
constexpr int WM_KB = 128;                     // rows per block
constexpr int WM_GROUP_BYTES = WM_KB * 128;    // one 64-channel group of a block: 16 KiB

__global__ void __launch_bounds__(WG_THREADS, 1)
wgrad_mn_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_x, const __grid_constant__ CUtensorMap tma_dy,
                        float* __restrict__ partial, float* __restrict__ db_partial, const WgMnCfg c) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  // with_db: a [128 k][64] tile of bf16 ones.  D2[o][0..15] += dY^T . ones = the column sums of dY - the bias gradient of
  // the layer - from the dY tile that is in shared memory anyway: no separate pass over dY (26 MB and ~14 us per layer
  // of the Wide&Deep tower).  Only the CTAs of column tile 0 issue the extra N = 16 instruction.
  const uint32_t ones = base + (uint32_t)c.stages * c.slot_bytes;
  const uint32_t bars = ones + (c.with_db ? (uint32_t)WM_GROUP_BYTES : 0u);
  const uint32_t full0 = bars, empty0 = bars + 8 * MAX_STAGES, done = empty0 + 8 * MAX_STAGES, tmem_slot = done + 8;
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - raw));
  const int warp = (int)warp_uniform(threadIdx.x >> 5), lane = threadIdx.x & 31;
  const int tile = (int)blockIdx.x % c.tiles, split = (int)blockIdx.x / c.tiles;
  const int o0 = (tile / c.tiles_c) * 128, c0 = (tile % c.tiles_c) * c.nc_tile;
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_x) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_dy) : "memory");
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < c.stages; ++s) {
        mbar_init(full0 + 8 * s, 1);
        mbar_init(empty0 + 8 * s, 1);
      }
      mbar_init(done, 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  const bool do_db = c.with_db && (tile % c.tiles_c) == 0;
  if (do_db) {
    for (uint32_t i = threadIdx.x; i < (uint32_t)WM_GROUP_BYTES / 16u; i += blockDim.x)
      asm volatile("st.shared.v4.b32 [%0], {%1,%1,%1,%1};" ::"r"(ones + i * 16u), "r"(0x3f803f80u) : "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic writes -> visible to the tensor core's reads
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = warp_uniform(*tmem_slot_ptr);
  const uint32_t x_off = 2u * WM_GROUP_BYTES;  // slot = [dY group 0][dY group 1][X group 0 .. groups_c - 1]

  if (warp == 0) {
    if (lane == 0) {
      uint32_t it = 0;
      for (int b = split; b < c.blocks; b += c.S, ++it) {
        const uint32_t s = it % (uint32_t)c.stages, ph = (it / (uint32_t)c.stages) & 1u;
        const uint32_t slot = base + s * c.slot_bytes;
        mbar_wait(empty0 + 8 * s, ph ^ 1u);
        mbar_expect_tx(full0 + 8 * s, (uint32_t)(2 + c.groups_c) * WM_GROUP_BYTES);
        for (int g = 0; g < 2; ++g) tma_load_2d(slot + g * WM_GROUP_BYTES, &tma_dy, o0 + 64 * g, b * WM_KB, full0 + 8 * s);
        for (int g = 0; g < c.groups_c; ++g)
          tma_load_2d(slot + x_off + g * WM_GROUP_BYTES, &tma_x, c0 + 64 * g, b * WM_KB, full0 + 8 * s);
      }
    }
  } else if (warp == 1) {
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(c.nc_tile >> 3) << 17) |
                           ((128u >> 4) << 24);
    const uint64_t d0 = umma_desc_ex(0, WM_GROUP_BYTES, 1024, 2);  // 64-channel groups one tile apart
    const uint32_t idesc16 = (idesc & ~(0x3fu << 17)) | ((16u >> 3) << 17);
    uint32_t it = 0;
    for (int b = split; b < c.blocks; b += c.S, ++it) {
      const uint32_t s = it % (uint32_t)c.stages, ph = (it / (uint32_t)c.stages) & 1u;
      const uint32_t slot = base + s * c.slot_bytes;
      mbar_wait(full0 + 8 * s, ph);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      uint64_t da = d0 + (uint64_t)(slot >> 4), db = d0 + (uint64_t)((slot + x_off) >> 4);
      uint64_t d1 = d0 + (uint64_t)(ones >> 4);
#pragma unroll
      for (int ks = 0; ks < WM_KB / 16; ++ks) {
        if (do_db) {  // column sums of dY: accumulator columns 256..271
          umma_f16_elect(tmem_base + 256u, da, d1, idesc16, (uint32_t)((it | (uint32_t)ks) != 0u));
          d1 += 2048 >> 4;
        }
        umma_f16_elect(tmem_base, da, db, idesc, (uint32_t)((it | (uint32_t)ks) != 0u));
        da += 2048 >> 4;
        db += 2048 >> 4;
      }
      umma_commit_elect(empty0 + 8 * s);
    }
    umma_commit_elect(done);
  } else {
    const int q = warp & 3;
    mbar_wait_warp_sleepy(done, 0, 512);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const bool has_work = split < c.blocks;
    const int pitch = c.nc_tile + 4;  // floats
    const uint32_t stg = base + (uint32_t)(warp - 2) * (32u * (256 + 4) * 4u);
    for (int col = 0; col < c.nc_tile; col += 16) {
      uint32_t r[16];
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
          : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
            "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
          : "r"(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)col)
          : "memory");
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
      for (int v = 0; v < 4; ++v)
        asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(stg + (uint32_t)(lane * pitch + col + 4 * v) * 4u),
                     "r"(has_work ? r[4 * v] : 0u), "r"(has_work ? r[4 * v + 1] : 0u), "r"(has_work ? r[4 * v + 2] : 0u),
                     "r"(has_work ? r[4 * v + 3] : 0u)
                     : "memory");
    }
    if (do_db) {
      uint32_t r[16];
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
          : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
            "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
          : "r"(tmem_base + ((uint32_t)(q * 32) << 16) + 256u)
          : "memory");
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      const int o = o0 + q * 32 + lane;
      if (o < c.No) db_partial[(size_t)split * c.No + o] = has_work ? __uint_as_float(r[0]) : 0.f;
    }
    __syncwarp();
    // one row per trip: the warp writes the row's valid columns as consecutive float4s
    for (int rl = 0; rl < 32; ++rl) {
      const int o = o0 + q * 32 + rl;
      if (o >= c.No) break;
      float* drow = partial + ((size_t)split * c.No + o) * c.Nc + c0;
      for (int cc = lane * 4; cc < c.nc_tile && c0 + cc < c.Nc; cc += 128) {
        float4 v;
        asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                     : "r"(stg + (uint32_t)(rl * pitch + cc) * 4u));
        *reinterpret_cast<float4*>(drow + cc) = v;
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
  }
}

// ---------------------------------------------------------------------------------------------------------------
// CTA-pair version (cluster of two, tcgen05 cta_group::2) of the MN-major weight gradient.  A pair owns [256 o] x
// [nct c] accumulators: each CTA loads its own 128 output channels of dY and HALF of the X columns, the leader issues
// M = 256 instructions that read both halves out of the two CTAs' shared memory, and each CTA ends up with its 128
// output rows of every column in its tensor memory.  The single-CTA kernel reads X once per 128 output channels
// and dY once per 256 columns; the pair reads X once per 256 output channels and dY once - half the L2 -> SM
// traffic, which is what bounds the kernel (~95 B/clk/SM against a ~42 B/clk/SM ceiling on the Wide&Deep tower).
// One instruction covers at most 256 columns = 128 of each CTA's half, so a wider tile takes two per K step; the
// accumulator columns of instruction i are [256 i, 256 i + 2 w): CTA 0's w columns, then CTA 1's.
struct WgPairCfg {
  int M, No, Nc;
  int nct, half, groups_h;   // tile columns (multiple of 16), columns per CTA, 64-column groups loaded per CTA
  int tiles_c, tiles, S, stages, blocks;
  uint32_t slot_bytes;
  int with_db;
};
constexpr uint32_t WP_DBCOL = 496;  // accumulator columns of the bias gradient (with_db: nct <= 496)

__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&r)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr)
               : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

__global__ void __launch_bounds__(WG_THREADS, 1)
wgrad_mn_pair_kernel(const __grid_constant__ CUtensorMap tma_x, const __grid_constant__ CUtensorMap tma_dy,
                     float* __restrict__ partial, float* __restrict__ db_partial, const WgPairCfg c) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t ones = base + (uint32_t)c.stages * c.slot_bytes;
  const uint32_t bars = ones + (c.with_db ? (uint32_t)WM_GROUP_BYTES : 0u);
  const uint32_t full0 = bars, empty0 = bars + 8 * MAX_STAGES, done = empty0 + 8 * MAX_STAGES, tmem_slot = done + 8;
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - raw));
  const int warp = (int)warp_uniform(threadIdx.x >> 5), lane = threadIdx.x & 31;
  const uint32_t crank = warp_uniform(cluster_ctarank());
  const int unit = (int)(blockIdx.x >> 1);
  const int tile = unit % c.tiles, split = unit / c.tiles;
  const int o0 = (tile / c.tiles_c) * 256 + (int)crank * 128;  // this CTA's output channels
  const int c0 = (tile % c.tiles_c) * c.nct;                   // the pair's columns
  const int ch0 = c0 + (int)crank * c.half;                    // the half this CTA loads
  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_x) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_dy) : "memory");
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < c.stages; ++s) {
        mbar_init(full0 + 8 * s, 1);
        mbar_init(empty0 + 8 * s, 1);
      }
      mbar_init(done, 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  const bool do_db = c.with_db && (tile % c.tiles_c) == 0;
  if (do_db) {  // both CTAs: the instruction reads 8 of the 16 ones-columns from each
    for (uint32_t i = threadIdx.x; i < (uint32_t)WM_GROUP_BYTES / 16u; i += blockDim.x)
      asm volatile("st.shared.v4.b32 [%0], {%1,%1,%1,%1};" ::"r"(ones + i * 16u), "r"(0x3f803f80u) : "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  cluster_sync_all();  // the peer's barriers and ones tile exist before anything touches them
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = warp_uniform(*tmem_slot_ptr);
  const uint32_t x_off = 2u * WM_GROUP_BYTES;  // slot = [dY group 0][dY group 1][X group 0 .. groups_h - 1]
  const int n_inst = (c.half + 127) / 128;

  if (warp == 0) {
    if (lane == 0) {
      uint32_t it = 0;
      for (int b = split; b < c.blocks; b += c.S, ++it) {
        const uint32_t s = it % (uint32_t)c.stages, ph = (it / (uint32_t)c.stages) & 1u;
        const uint32_t slot = base + s * c.slot_bytes;
        mbar_wait(empty0 + 8 * s, ph ^ 1u);
        if (crank == 0) mbar_expect_tx(full0 + 8 * s, 2u * (uint32_t)(2 + c.groups_h) * WM_GROUP_BYTES);
        const uint32_t lbar = mapa_rank(full0 + 8 * s, 0);
        for (int g = 0; g < 2; ++g) tma_load_2d_pair(slot + g * WM_GROUP_BYTES, &tma_dy, o0 + 64 * g, b * WM_KB, lbar);
        for (int g = 0; g < c.groups_h; ++g)
          tma_load_2d_pair(slot + x_off + g * WM_GROUP_BYTES, &tma_x, ch0 + 64 * g, b * WM_KB, lbar);
      }
    }
  } else if (warp == 1) {
    if (crank == 0) {  // the leader issues for the pair (whole warp, one elected lane)
      const uint32_t idesc0 = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((256u >> 4) << 24);
      const uint64_t d0 = umma_desc_ex(0, WM_GROUP_BYTES, 1024, 2);
      uint32_t it = 0;
      for (int b = split; b < c.blocks; b += c.S, ++it) {
        const uint32_t s = it % (uint32_t)c.stages, ph = (it / (uint32_t)c.stages) & 1u;
        const uint32_t slot = base + s * c.slot_bytes;
        mbar_wait(full0 + 8 * s, ph);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        uint64_t da = d0 + (uint64_t)(slot >> 4), db = d0 + (uint64_t)((slot + x_off) >> 4);
        uint64_t d1 = d0 + (uint64_t)(ones >> 4);
#pragma unroll
        for (int ks = 0; ks < WM_KB / 16; ++ks) {
          const uint32_t acc = (uint32_t)((it | (uint32_t)ks) != 0u);
          if (do_db) {
            umma_f16_pair_elect(tmem_base + WP_DBCOL, da, d1, idesc0 | ((16u >> 3) << 17), acc);
            d1 += 2048 >> 4;
          }
          for (int i = 0; i < n_inst; ++i) {
            const uint32_t w = (uint32_t)((c.half - 128 * i) < 128 ? (c.half - 128 * i) : 128);
            umma_f16_pair_elect(tmem_base + 256u * (uint32_t)i, da, db + (uint64_t)((uint32_t)i * (2u * WM_GROUP_BYTES >> 4)),
                                idesc0 | (((2u * w) >> 3) << 17), acc);
          }
          da += 2048 >> 4;
          db += 2048 >> 4;
        }
        umma_commit_pair_elect(empty0 + 8 * s);
      }
      umma_commit_pair_elect(done);
    }
  } else {
    const int q = warp & 3;
    mbar_wait_warp_sleepy(done, 0, 512);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const bool has_work = split < c.blocks;
    const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
    const uint32_t stg = base + (uint32_t)(warp - 2) * (32u * (128 + 4) * 4u);
    const int o_first = o0 + q * 32;
    if (do_db) {
      uint32_t r[8];
      tmem_ld8(lane_base + WP_DBCOL, r);
      const int o = o_first + lane;
      if (o < c.No) db_partial[(size_t)split * c.No + o] = has_work ? __uint_as_float(r[0]) : 0.f;
    }
    for (int i = 0; i < n_inst; ++i) {
      const int w = (c.half - 128 * i) < 128 ? (c.half - 128 * i) : 128;
      const int pitch = w + 4;  // floats
      for (int h = 0; h < 2; ++h) {
        const uint32_t t0 = 256u * (uint32_t)i + (uint32_t)(h * w);   // accumulator columns of this segment
        const int gc = c0 + h * c.half + 128 * i;                     // where they belong in dW
        for (int col = 0; col < w; col += 8) {
          uint32_t r[8];
          tmem_ld8(lane_base + t0 + (uint32_t)col, r);
#pragma unroll
          for (int v = 0; v < 2; ++v)
            asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(stg + (uint32_t)(lane * pitch + col + 4 * v) * 4u),
                         "r"(has_work ? r[4 * v] : 0u), "r"(has_work ? r[4 * v + 1] : 0u), "r"(has_work ? r[4 * v + 2] : 0u),
                         "r"(has_work ? r[4 * v + 3] : 0u)
                         : "memory");
        }
        __syncwarp();
        for (int rl = 0; rl < 32; ++rl) {
          const int o = o_first + rl;
          if (o >= c.No) break;
          float* drow = partial + ((size_t)split * c.No + o) * c.Nc + gc;
          for (int cc = lane * 4; cc < w && gc + cc < c.Nc; cc += 128) {
            float4 v;
            asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];"
                         : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                         : "r"(stg + (uint32_t)(rl * pitch + cc) * 4u));
            *reinterpret_cast<float4*>(drow + cc) = v;
          }
        }
        __syncwarp();
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  cluster_sync_all();  // neither CTA leaves or frees tensor memory while the pair's instructions may still touch it
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
  }
}

int round_up(int a, int b) { return (a + b - 1) / b * b; }
constexpr bool PAIR_WGRAD_DEFAULT = false;  // see profiles/r2/wgrad_pair.md

// Chooses rows per block (R), stages and the split for one layer shape.  Returns 0 when the shape is not supported.
int make_cfg(int N, int H, int W, int Cin, int Cout, Wg3Cfg* out) {
  if (N < 1 || H < 1 || W < 1 || W + 2 > 256 || Cin % WG_NC != 0 || Cout % 64 != 0) return 0;
  Wg3Cfg c{};
  c.n_img = N; c.H = H; c.W = W; c.Cin = Cin; c.Cout = Cout;
  c.P = W + 2;
  c.mo = Cout % 128 == 0 ? 128 : 64;
  c.groups = c.mo / 64;
  const int budget = SMEM_LIMIT - 1024 - 512;
  const int want_stage = 56 * 1024;
  long long best_cost = -1;
  for (int R = 1; R <= H && R + 2 <= 256; ++R) {
    const int kpad = round_up(R * c.P, 16);
    const int xrows = kpad + 2 * c.P + 2;
    const int stage = c.groups * kpad * 128 + round_up(xrows * 64, 1024);
    if (stage > want_stage && R > 1) break;
    if (stage > budget / 2) break;
    const long long cost = (long long)((H + R - 1) / R) * kpad;  // padded reduction length per image
    if (best_cost < 0 || cost <= best_cost) {
      best_cost = cost;
      c.R = R;
      c.kpad = kpad;
    }
  }
  if (best_cost < 0) return 0;
  c.chunks = (H + c.R - 1) / c.R;
  c.dy_group_bytes = (uint32_t)c.kpad * 128u;
  c.dy_slot_bytes = (uint32_t)c.groups * c.dy_group_bytes;
  c.x_slot_bytes = (uint32_t)round_up((c.kpad + 2 * c.P + 2) * 64, 1024);
  c.x_box_bytes = (uint32_t)((c.R + 2) * c.P * 64);
  c.dy_box_bytes = (uint32_t)(c.R * c.P * 128);
  c.stages = budget / (int)(c.dy_slot_bytes + c.x_slot_bytes);
  if (c.stages > MAX_STAGES) c.stages = MAX_STAGES;
  if (c.stages < 2 || (size_t)c.stages * (c.dy_slot_bytes + c.x_slot_bytes) < 4u * 32u * (WG_COLS + 4) * 4u) return 0;  // epilogue staging
  c.tiles_c = Cin / WG_NC;
  c.tiles = (Cout / c.mo) * c.tiles_c;
  { const char* d = getenv("PDJ_WG_DEBUG"); c.debug = d ? atoi(d) : 0; }
  const int sms = pdj_tc_sm_count();
  const long long blocks = (long long)N * c.chunks;
  c.S = sms / c.tiles;
  if (c.S < 1) c.S = 1;
  if (c.S > blocks) c.S = (int)blocks;
  *out = c;
  return 1;
}

}  // namespace

extern "C" {

size_t pdj_conv3x3_wgrad_ws_bytes(int N, int H, int W, int Cin, int Cout) {
  Wg3Cfg c;
  if (!make_cfg(N, H, W, Cin, Cout, &c)) return 0;
  return (size_t)c.S * Cout * WG_TAPS * Cin * sizeof(float);
}

// dw[Cout][3][3][Cin] (bf16; the memory of a channels-last [Cout, Cin, 3, 3] weight) = gradient of the 3x3 / pad 1 /
// stride 1 convolution with respect to its weight, given x[N][H][W][Cin] and dy[N][H][W][Cout] (NHWC bf16).
// `ws`: pdj_conv3x3_wgrad_ws_bytes(...) bytes.  Cin % 32 == 0, Cout % 64 == 0, W <= 254.
int pdj_conv3x3_wgrad(const void* x, const void* dy, void* dw, void* ws, int N, int H, int W, int Cin, int Cout,
                      cudaStream_t st) {
  Wg3Cfg c;
  if (!make_cfg(N, H, W, Cin, Cout, &c) || ws == nullptr || ((uintptr_t)x & 15) || ((uintptr_t)dy & 15) || ((uintptr_t)dw & 7)) {
    pdj_set_error("pdj_conv3x3_wgrad: need Cin %% 32 == 0, Cout %% 64 == 0, W <= 254, 16-byte aligned tensors and a workspace "
                  "(N=%d H=%d W=%d Cin=%d Cout=%d)", N, H, W, Cin, Cout);
    return -1;
  }
  static bool attr = false;
  if (!attr) {
    PDJ_CUDA_CHECK(cudaFuncSetAttribute(wgrad3x3_tcgen05_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
    attr = true;
  }
  CUtensorMap mx, my;
  const uint64_t dx[4] = {(uint64_t)Cin, (uint64_t)W, (uint64_t)H, (uint64_t)N};
  const uint64_t sx[3] = {(uint64_t)Cin * 2, (uint64_t)W * Cin * 2, (uint64_t)H * W * Cin * 2};
  const uint32_t bx[4] = {(uint32_t)WG_NC, (uint32_t)c.P, (uint32_t)(c.R + 2), 1};
  if (int r = pdj_tc_make_map4(&mx, x, dx, sx, bx, 2)) return r;
  const uint64_t dd[4] = {(uint64_t)Cout, (uint64_t)W, (uint64_t)H, (uint64_t)N};
  const uint64_t sd[3] = {(uint64_t)Cout * 2, (uint64_t)W * Cout * 2, (uint64_t)H * W * Cout * 2};
  const uint32_t bd[4] = {64, (uint32_t)c.P, (uint32_t)c.R, 1};
  if (int r = pdj_tc_make_map4(&my, dy, dd, sd, bd, 3)) return r;
  const int smem = 1024 + c.stages * (int)(c.dy_slot_bytes + c.x_slot_bytes) + 512;
  wgrad3x3_tcgen05_kernel<<<c.tiles * c.S, WG_THREADS, smem, st>>>(mx, my, (float*)ws, c);
  PDJ_CUDA_CHECK(cudaGetLastError());
  const size_t n = (size_t)Cout * WG_TAPS * Cin;
  launch_reduce((const float*)ws, c.S, n, (__nv_bfloat16*)dw, st);
  PDJ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

static int make_mn_cfg(int M, int No, int Nc, WgMnCfg* out, int with_db = 0) {
  if (M < 1 || No < 1 || Nc < 8 || Nc % 8 != 0 || No % 8 != 0) return 0;
  WgMnCfg c{};
  c.M = M; c.No = No; c.Nc = Nc;
  c.with_db = with_db;
  const int parts = (Nc + 255) / 256;
  c.nc_tile = round_up((Nc + parts - 1) / parts, 64);
  c.groups_c = c.nc_tile / 64;
  c.tiles_c = (Nc + c.nc_tile - 1) / c.nc_tile;
  c.tiles = ((No + 127) / 128) * c.tiles_c;
  c.blocks = (M + WM_KB - 1) / WM_KB;
  c.slot_bytes = (uint32_t)(2 + c.groups_c) * WM_GROUP_BYTES;
  c.stages = (SMEM_LIMIT - 1024 - 512 - (with_db ? WM_GROUP_BYTES : 0)) / (int)c.slot_bytes;
  if (c.stages > MAX_STAGES) c.stages = MAX_STAGES;
  if (c.stages < 2) return 0;
  c.S = pdj_tc_sm_count() / c.tiles;
  if (c.S < 1) c.S = 1;
  if (c.S > c.blocks) c.S = c.blocks;
  *out = c;
  return 1;
}

// CTA pairs pay when there are more than 128 output channels to share an X tile between (PDJ_WGRAD_PAIR=0/1 overrides).
static int make_pair_cfg(int M, int No, int Nc, WgPairCfg* out, int with_db) {
  const char* e = getenv("PDJ_WGRAD_PAIR");
  const bool want = e ? atoi(e) != 0 : PAIR_WGRAD_DEFAULT;
  if (!want || M < 1 || No <= 128 || Nc < 16 || Nc % 8 != 0 || No % 8 != 0) return 0;
  WgPairCfg c{};
  c.M = M; c.No = No; c.Nc = Nc;
  c.with_db = with_db;
  const int maxw = 496;
  const int parts = (Nc + maxw - 1) / maxw;
  c.nct = round_up((Nc + parts - 1) / parts, 16);
  if (c.nct > maxw) return 0;
  c.half = c.nct / 2;
  c.groups_h = (c.half + 63) / 64;
  c.tiles_c = (Nc + c.nct - 1) / c.nct;
  c.tiles = ((No + 255) / 256) * c.tiles_c;
  c.blocks = (M + WM_KB - 1) / WM_KB;
  c.slot_bytes = (uint32_t)(2 + c.groups_h) * WM_GROUP_BYTES;
  c.stages = (SMEM_LIMIT - 1024 - 512 - (with_db ? WM_GROUP_BYTES : 0)) / (int)c.slot_bytes;
  if (c.stages > MAX_STAGES) c.stages = MAX_STAGES;
  if (c.stages < 2) return 0;
  const int units = pdj_tc_sm_count() / 2;
  if (units < c.tiles) return 0;
  c.S = units / c.tiles;
  if (c.S > c.blocks) c.S = c.blocks;
  *out = c;
  return 1;
}

size_t pdj_wgrad_mn_ws_bytes(int M, int No, int Nc) {
  WgMnCfg c;
  if (!make_mn_cfg(M, No, Nc, &c)) return 0;
  size_t n = (size_t)c.S * No * Nc * sizeof(float);
  WgPairCfg p;
  if (make_pair_cfg(M, No, Nc, &p, 0)) n = n > (size_t)p.S * No * Nc * sizeof(float) ? n : (size_t)p.S * No * Nc * sizeof(float);
  return n;
}

// dw[No][Nc] (bf16, row-major) = dy[M][No]^T . x[M][Nc]: the weight gradient of a linear layer or of a 1x1
// convolution over NHWC activations viewed as [pixels][channels].  Row pitches ldy / ldx in elements (multiples of
// 8); No and Nc multiples of 8.  `ws`: pdj_wgrad_mn_ws_bytes(M, No, Nc) bytes.
static int launch_wgrad_mn(const void* x, const void* dy, void* dw, void* db, void* ws, int M, int No, int Nc, int ldx, int ldy,
                           cudaStream_t st) {
  WgMnCfg c;
  if (!make_mn_cfg(M, No, Nc, &c, db != nullptr) || ws == nullptr || ldx % 8 != 0 || ldy % 8 != 0 || ldx < Nc || ldy < No ||
      ((uintptr_t)x & 15) || ((uintptr_t)dy & 15) || ((uintptr_t)dw & 7) || ((uintptr_t)db & 7)) {
    pdj_set_error("pdj_wgrad_mn: need No, Nc multiples of 8, row pitches multiples of 8 elements, 16-byte aligned tensors "
                  "and a workspace (M=%d No=%d Nc=%d ldx=%d ldy=%d)", M, No, Nc, ldx, ldy);
    return -1;
  }
  static bool attr = false;
  if (!attr) {
    PDJ_CUDA_CHECK(cudaFuncSetAttribute(wgrad_mn_tcgen05_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
    attr = true;
  }
  CUtensorMap mx, my;
  if (int r = pdj_tc_make_map(&mx, x, M, Nc, ldx, 64, WM_KB, true)) return r;
  if (int r = pdj_tc_make_map(&my, dy, M, No, ldy, 64, WM_KB, true)) return r;
  WgPairCfg pc;
  if (make_pair_cfg(M, No, Nc, &pc, db != nullptr)) {
    static bool pattr = false;
    if (!pattr) {
      PDJ_CUDA_CHECK(cudaFuncSetAttribute(wgrad_mn_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT));
      pattr = true;
    }
    const int psmem = 1024 + pc.stages * (int)pc.slot_bytes + (pc.with_db ? WM_GROUP_BYTES : 0) + 512;
    float* pdb = db != nullptr ? (float*)ws + (size_t)pc.S * No * Nc : nullptr;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * pc.tiles * pc.S);
    cfg.blockDim = dim3(WG_THREADS);
    cfg.dynamicSmemBytes = psmem;
    cfg.stream = st;
    cudaLaunchAttribute attr1[1];
    attr1[0].id = cudaLaunchAttributeClusterDimension;
    attr1[0].val.clusterDim.x = 2;
    attr1[0].val.clusterDim.y = 1;
    attr1[0].val.clusterDim.z = 1;
    cfg.attrs = attr1;
    cfg.numAttrs = 1;
    PDJ_CUDA_CHECK(cudaLaunchKernelEx(&cfg, wgrad_mn_pair_kernel, mx, my, (float*)ws, pdb, pc));
    launch_reduce((const float*)ws, pc.S, (size_t)No * Nc, (__nv_bfloat16*)dw, st);
    if (db != nullptr) launch_reduce(pdb, pc.S, (size_t)No, (__nv_bfloat16*)db, st);
    PDJ_CUDA_CHECK(cudaGetLastError());
    return 0;
  }
  const int smem = 1024 + c.stages * (int)c.slot_bytes + (c.with_db ? WM_GROUP_BYTES : 0) + 512;
  float* dbp = db != nullptr ? (float*)ws + (size_t)c.S * No * Nc : nullptr;
  wgrad_mn_tcgen05_kernel<<<c.tiles * c.S, WG_THREADS, smem, st>>>(mx, my, (float*)ws, dbp, c);
  PDJ_CUDA_CHECK(cudaGetLastError());
  launch_reduce((const float*)ws, c.S, (size_t)No * Nc, (__nv_bfloat16*)dw, st);
  if (db != nullptr) launch_reduce(dbp, c.S, (size_t)No, (__nv_bfloat16*)db, st);
  PDJ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int pdj_wgrad_mn(const void* x, const void* dy, void* dw, void* ws, int M, int No, int Nc, int ldx, int ldy,
                 cudaStream_t st) {
  return launch_wgrad_mn(x, dy, dw, nullptr, ws, M, No, Nc, ldx, ldy, st);
}

// The same plus db[No] (bf16) = the column sums of dy - the bias gradient - from the dY tiles the kernel already has in
// shared memory (one extra N = 16 tensor-core instruction per 16 rows against a tile of ones).
// `ws`: pdj_wgrad_mn_db_ws_bytes(M, No, Nc) bytes.
size_t pdj_wgrad_mn_db_ws_bytes(int M, int No, int Nc) {
  WgMnCfg c;
  if (!make_mn_cfg(M, No, Nc, &c, 1)) return 0;
  size_t n = (size_t)c.S * No * Nc * sizeof(float) + (size_t)c.S * No * sizeof(float);
  WgPairCfg p;
  if (make_pair_cfg(M, No, Nc, &p, 1)) {
    const size_t np = (size_t)p.S * No * Nc * sizeof(float) + (size_t)p.S * No * sizeof(float);
    n = n > np ? n : np;
  }
  return n;
}

int pdj_wgrad_mn_db(const void* x, const void* dy, void* dw, void* db, void* ws, int M, int No, int Nc, int ldx, int ldy,
                    cudaStream_t st) {
  if (db == nullptr) {
    pdj_set_error("pdj_wgrad_mn_db: db is NULL");
    return -1;
  }
  return launch_wgrad_mn(x, dy, dw, db, ws, M, No, Nc, ldx, ldy, st);
}

}  // extern "C"
