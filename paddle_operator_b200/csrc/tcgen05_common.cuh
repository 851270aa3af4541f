// Device-side building blocks shared by the tcgen05 kernels (gemm_tcgen05.cu, conv_tcgen05.cu): mbarrier /
// TMA / tcgen05 PTX wrappers, the shared-memory operand ring and the TMA-producer / MMA-issuer loops of the
// persistent GEMM pipeline, plus the host helpers of tma_util.cu.  sm_100a only.
#pragma once
#include <cuda.h>
#include <stdint.h>

#include "common.cuh"

namespace pdj_tc {

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
// Long waits (an epilogue warp waiting for a whole tile or a whole kernel's worth of MMAs): one lane polls with a
// back-off and the rest of the warp parks at __syncwarp.  128 threads spinning on try_wait are 128 shared-memory
// polls every few cycles, taken out of the 128 B/clk the tensor pipe needs for its operands.
__device__ __forceinline__ void mbar_wait_warp_sleepy(uint32_t bar, uint32_t parity, uint32_t ns) {
  if ((threadIdx.x & 31) == 0) {
    uint32_t done = 0;
    while (true) {
      asm volatile(
          "{\n\t"
          ".reg .pred p;\n\t"
          "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
          "selp.u32 %0, 1, 0, p;\n\t"
          "}"
          : "=r"(done)
          : "r"(bar), "r"(parity)
          : "memory");
      if (done) break;
      __nanosleep(ns);
    }
  }
  __syncwarp();
}
// True in exactly one lane of a fully active warp.  The MMA issuer runs its loops with the WHOLE warp and guards only
// the tcgen05 instructions with this: control flow and descriptor arithmetic then stay warp-uniform, the compiler
// keeps them in uniform registers, and a UTCHMMA issues without the R2UR "waterfall" that per-lane code needs
// (measured: ~120 -> ~56 cycles per N = 96 instruction).
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t"
      "}"
      : "=r"(pred));
  return pred != 0;
}
// A value the compiler can prove warp-uniform (lane 0's copy).
__device__ __forceinline__ uint32_t warp_uniform(uint32_t v) { return __shfl_sync(0xffffffffu, v, 0); }
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
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* map, int c0, int c1, int c2, int c3, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];" ::"r"(dst),
      "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(bar)
      : "memory");
}
// im2col-mode load over an NHWC tensor map (cuTensorMapEncodeIm2col): `pixels per column` output pixels starting at
// the one whose top-left filter tap sits at input coordinates (w, h) of image n, 64 channels from c; (off_w, off_h)
// select the filter tap.  Pixels are walked W -> H -> N inside the padded bounding box; padding reads as zero.
__device__ __forceinline__ void tma_load_im2col_4d(uint32_t dst, const CUtensorMap* map, int c, int w, int h, int n,
                                                   uint16_t off_w, uint16_t off_h, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.im2col.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6], {%7, %8};" ::
          "r"(dst), "l"(map), "r"(c), "r"(w), "r"(h), "r"(n), "r"(bar), "h"(off_w), "h"(off_h)
      : "memory");
}
// General shared-memory matrix descriptor (sm_100 version 1): start address, leading / stride byte offsets and the
// swizzle mode (2 = 128 B, 4 = 64 B, 6 = 32 B, 0 = none).  The base-offset field stays 0 even when `smem_addr` is not
// aligned to the swizzle pattern: the hardware applies the XOR to absolute shared-memory address bits, so a tile
// written by TMA can be read from any 16-byte-aligned row of it (profiles/r2/desc_shift_experiment.log).
__device__ __forceinline__ uint64_t umma_desc_ex(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t layout) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)(layout & 7) << 61;
  return d;
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

// Warp-wide forms: every lane of a converged warp executes the statement, one elected lane issues the instruction.
// No C++ branch surrounds it, so the operands stay in uniform registers (a branch on elect_one() makes the region
// divergent for the compiler and brings the per-lane R2UR loop back).
__device__ __forceinline__ void umma_f16_elect(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p, q;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "l"(da), "l"(db), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit_elect(uint32_t bar) {
  asm volatile(
      "{\n\t"
      ".reg .pred q;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t"
      "}" ::"r"(bar)
      : "memory");
}

// ---- CTA pairs (cluster of two, tcgen05 cta_group::2): see linear_tcgen05_kernel<*, PAIR2> in gemm_tcgen05.cu ----
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t mapa_rank(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// TMA load of a CTA pair: data into this CTA's shared memory, bytes counted on the LEADER's mbarrier.
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t leader_bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst),
      "l"(map), "r"(c0), "r"(c1), "r"(leader_bar)
      : "memory");
}
__device__ __forceinline__ void umma_f16_pair_elect(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p, q;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "@q tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d),
      "l"(da), "l"(db), "r"(idesc), "r"(accumulate)
      : "memory");
}
// Arrives on the barrier at this shared-memory offset in BOTH CTAs of the pair once the MMAs issued so far are done.
__device__ __forceinline__ void umma_commit_pair_elect(uint32_t bar) {
  asm volatile(
      "{\n\t"
      ".reg .pred q;\n\t"
      "elect.sync _|q, 0xffffffff;\n\t"
      "@q tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n\t"
      "}" ::"r"(bar),
      "h"((uint16_t)3)
      : "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_bar) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_bar) : "memory");
}

// Shared-memory ring + accumulator hand-off barriers of one CTA.
struct Pipe {
  uint32_t a_smem, b_smem, b_stage_bytes;
  uint32_t full0, empty0;      // per ring slot: TMA -> MMA, MMA -> TMA
  uint32_t tfull0, tempty0;    // per accumulator buffer: MMA -> epilogue, epilogue -> MMA
  int stages;
  int b_mn;                    // B is given as [K][N] (N contiguous, e.g. a weight used transposed): MN-major operand;
                               // 1: one 2-D box per 64-column group, 2: one 4-D box {64, BK, groups, 1} (N % 64 == 0)
};
constexpr int B_MN_GROUP_BYTES = BK * 128;  // MN-major B stage: 64-column groups of [BK k-rows][128 B], 8 KiB each
__device__ __host__ __forceinline__ uint32_t b_stage_bytes_for(int block_n, int b_mn) {
  return b_mn ? (uint32_t)((block_n + 63) / 64) * B_MN_GROUP_BYTES : (uint32_t)block_n * BK * 2;
}

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
      if (p.b_mn == 2) {  // all 64-column groups of the stage in one request
        tma_load_4d(p.b_smem + s * p.b_stage_bytes, tma_b, 0, kb * BK, n0 >> 6, 0, p.full0 + 8 * s);
      } else if (p.b_mn) {  // [k rows][64 columns] boxes of the [K][N] matrix, one per 64-column group
        for (int g = 0; g * 64 < block_n; ++g)
          tma_load_2d(p.b_smem + s * p.b_stage_bytes + g * B_MN_GROUP_BYTES, tma_b, n0 + 64 * g, kb * BK, p.full0 + 8 * s);
      } else {
        tma_load_2d(p.b_smem + s * p.b_stage_bytes, tma_b, kb * BK, n0, p.full0 + 8 * s);
      }
    }
  }
}

// MMA issuer: called by the WHOLE warp (converged); one elected lane issues.  Tile i accumulates into TMEM buffer
// i & 1.  Descriptors are built once and advanced by adding to their address field, all in uniform registers.
__device__ __forceinline__ void mma_loop(const Pipe& p, uint32_t tmem_base, int total, int block_n, int num_kb) {
  // instruction descriptor: D = f32, A = B = bf16, both K-major, N >> 3, M >> 4
  // (bit 16: B is MN-major - its 64-column groups sit B_MN_GROUP_BYTES apart, 8-row k groups 1024 B apart, and one
  // K = 16 step is sixteen 128-byte rows)
  const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (p.b_mn ? (1u << 16) : 0u) | ((uint32_t)(block_n >> 3) << 17) |
                         ((uint32_t)(BM >> 4) << 24);
  const uint64_t d0 = umma_desc(0);
  const uint64_t d0b = p.b_mn ? umma_desc_ex(0, B_MN_GROUP_BYTES, 1024, 2) : d0;
  const uint32_t kstep_b = p.b_mn ? (16u * 128u) >> 4 : (uint32_t)(UK * 2) >> 4;
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
      const uint64_t da = d0 + (uint64_t)((p.a_smem + s * A_STAGE_BYTES) >> 4);
      const uint64_t db = d0b + (uint64_t)((p.b_smem + s * p.b_stage_bytes) >> 4);
#pragma unroll
      for (int k = 0; k < BK / UK; ++k)
        umma_f16_elect(tmem_d, da + (uint64_t)(k * UK * 2 >> 4), db + (uint64_t)((uint32_t)k * kstep_b), idesc, (uint32_t)((kb | k) != 0));
      umma_commit_elect(p.empty0 + 8 * s);  // slot is free once these MMAs have read it
    }
    umma_commit_elect(p.tfull0 + 8 * buf);  // accumulator complete
  }
}

}  // namespace pdj_tc

// ---- host helpers (tma_util.cu) ---------------------------------------------------------------
// Tensor-map encoders (cuTensorMapEncodeTiled through the runtime's driver entry point; they bind the
// primary context first when the calling thread has none) and the cached SM count.
int pdj_tc_make_map(CUtensorMap* map, const void* ptr, int rows, int cols, int ld_elems, int box_cols, int box_rows,
                    bool swizzle);
int pdj_tc_make_map3(CUtensorMap* map, const void* ptr, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t s1, uint64_t s2,
                     int b0, int b1);
// Map over a row-major bf16 matrix b[K][N] (N % 64 == 0) as {64, K, N / 64, 1}: one box {64, box_k, groups, 1} lands as
// `groups` consecutive [box_k][64] tiles, 128-byte swizzled - an MN-major tcgen05 B stage in one request.
int pdj_tc_make_map_kn(CUtensorMap* map, const void* ptr, int K, int N, int ld_elems, int box_k, int groups);
// 4-D tiled map over an NHWC bf16 tensor {C, W, H, N} (strides in bytes for W, H, N); swizzle: 0 none, 1 32 B,
// 2 64 B, 3 128 B.  Out-of-range box elements (negative or past-the-end coordinates) read as zero.
int pdj_tc_make_map4(CUtensorMap* map, const void* ptr, const uint64_t dims[4], const uint64_t strides[3], const uint32_t box[4],
                     int swizzle);
// im2col map over an NHWC bf16 tensor for a ksize x ksize window with symmetric padding `pad` and traversal stride
// `stride`: `pixels` output pixels x `channels` (<= 64) channels per load, 128-byte swizzle.
int pdj_tc_make_im2col4(CUtensorMap* map, const void* ptr, int C, int W, int H, int N, int ksize, int pad, int stride,
                        int channels, int pixels);
int pdj_tc_sm_count();
