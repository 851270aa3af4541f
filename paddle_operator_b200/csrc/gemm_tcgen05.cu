// Fused Linear (+bias)(+ReLU)(+mask) on the 5th-generation tensor cores:
//   out[M,N] = epilogue( A[M,K] . B[N,K]^T )       bf16 in, fp32 accumulate in TMEM, bf16 out
//
// Blackwell-native structure (sm_100a only):
//   * operands arrive by TMA (cp.async.bulk.tensor.2d, 128-byte swizzle) into a
//     4-stage shared-memory ring guarded by mbarriers;
//   * one elected thread issues tcgen05.mma (cta_group::1, kind::f16, M=128,
//     N=block_n, K=16) with the accumulator in tensor memory; tcgen05.commit
//     releases ring slots and finally signals the epilogue;
//   * four epilogue warps read the accumulator with tcgen05.ld (32 lanes x 32 bit
//     x 16 columns), add the bias, apply ReLU and/or a ReLU-derivative mask, pack
//     to bf16 and store.
// Out-of-range rows / columns / K tail are handled by TMA zero fill and guarded
// stores, so M, N, K need no padding (only 16-byte aligned row pitches).
//
// Used by the Wide&Deep MLP (models/wide_deep.py) forward and input-gradient
// GEMMs.  Reference parity: none (the reference has no compute code); workload of
// deploy/examples/wide_and_deep.yaml.
#include <cuda.h>

#include "common.cuh"
#include "../../include/pdjob_b200.h"

namespace {

constexpr int BM = 128;          // rows of A per CTA = UMMA M
constexpr int BK = 64;           // K per stage: 64 bf16 = one 128-byte swizzle row
constexpr int UK = 16;           // K per tcgen05.mma for 16-bit inputs
constexpr int STAGES = 4;
constexpr int MAX_BN = 256;
constexpr int A_STAGE_BYTES = BM * BK * 2;        // 16 KiB
constexpr int B_STAGE_BYTES = MAX_BN * BK * 2;    // 32 KiB
constexpr int TMEM_COLS = 256;
constexpr int SMEM_BYTES = STAGES * (A_STAGE_BYTES + B_STAGE_BYTES) + 256 + 1024;  // + barriers + alignment slack
constexpr int THREADS = 192;     // warp 0: TMA, warp 1: MMA + TMEM owner, warps 2-5: epilogue

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
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

__global__ void __launch_bounds__(THREADS, 1)
linear_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                      __nv_bfloat16* __restrict__ out, const float* __restrict__ bias,
                      const __nv_bfloat16* __restrict__ mask, int M, int N, int K, int ldo, int block_n, int relu) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;  // SWIZZLE_128B wants 1024-byte alignment
  const uint32_t a_smem = base;
  const uint32_t b_smem = base + STAGES * A_STAGE_BYTES;
  const uint32_t bars = b_smem + STAGES * B_STAGE_BYTES;
  const uint32_t full0 = bars, empty0 = bars + 8 * STAGES, tmem_full = bars + 16 * STAGES, tmem_slot = tmem_full + 8;
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * BM, n0 = blockIdx.y * block_n;
  const int num_kb = (K + BK - 1) / BK;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tma_b) : "memory");
  }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < STAGES; ++s) {
        mbar_init(full0 + 8 * s, 1);
        mbar_init(empty0 + 8 * s, 1);
      }
      mbar_init(tmem_full, 1);
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

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      const uint32_t stage_bytes = (uint32_t)(BM + block_n) * BK * 2;
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % STAGES;
        const uint32_t ph = (uint32_t)(kb / STAGES) & 1u;
        mbar_wait(empty0 + 8 * s, ph ^ 1u);
        mbar_expect_tx(full0 + 8 * s, stage_bytes);
        tma_load_2d(a_smem + s * A_STAGE_BYTES, &tma_a, kb * BK, m0, full0 + 8 * s);
        tma_load_2d(b_smem + s * B_STAGE_BYTES, &tma_b, kb * BK, n0, full0 + 8 * s);
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (lane == 0) {
      // instruction descriptor: D = f32, A = B = bf16, both K-major, N >> 3, M >> 4
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(block_n >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % STAGES;
        const uint32_t ph = (uint32_t)(kb / STAGES) & 1u;
        mbar_wait(full0 + 8 * s, ph);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
        for (int k = 0; k < BK / UK; ++k) {
          const uint64_t da = umma_desc(a_smem + s * A_STAGE_BYTES + k * UK * 2);
          const uint64_t db = umma_desc(b_smem + s * B_STAGE_BYTES + k * UK * 2);
          umma_f16(tmem_base, da, db, idesc, (uint32_t)((kb | k) != 0));
        }
        umma_commit(empty0 + 8 * s);  // slot is free once these MMAs have read it
      }
      umma_commit(tmem_full);  // accumulator complete
    }
  } else {
    // ===== epilogue: warps 2..5 own TMEM lane quarters (warp % 4) =====
    mbar_wait(tmem_full, 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const int q = warp & 3;
    const int row = m0 + q * 32 + lane;
    const uint32_t taddr_row = tmem_base + ((uint32_t)(q * 32) << 16);
    for (int c0 = 0; c0 < block_n; c0 += 16) {
      uint32_t r[16];
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
          : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
            "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
          : "r"(taddr_row + (uint32_t)c0)
          : "memory");
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      const int col = n0 + c0;
      if (row < M && col < N) {
        float v[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          v[j] = __uint_as_float(r[j]);
          if (bias != nullptr && col + j < N) v[j] += __ldg(bias + col + j);
          if (relu) v[j] = fmaxf(v[j], 0.f);
        }
        __nv_bfloat16* o = out + (size_t)row * ldo + col;
        if (mask != nullptr) {
          const __nv_bfloat16* mk = mask + (size_t)row * ldo + col;
#pragma unroll
          for (int j = 0; j < 16; ++j)
            if (col + j < N && !(__bfloat162float(mk[j]) > 0.f)) v[j] = 0.f;
        }
        if (col + 16 <= N && (((size_t)row * ldo + col) & 7) == 0) {
          uint4 p0, p1;
          p0.x = pdj_f2_to_bf2(v[0], v[1]); p0.y = pdj_f2_to_bf2(v[2], v[3]);
          p0.z = pdj_f2_to_bf2(v[4], v[5]); p0.w = pdj_f2_to_bf2(v[6], v[7]);
          p1.x = pdj_f2_to_bf2(v[8], v[9]); p1.y = pdj_f2_to_bf2(v[10], v[11]);
          p1.z = pdj_f2_to_bf2(v[12], v[13]); p1.w = pdj_f2_to_bf2(v[14], v[15]);
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
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn g_encode = nullptr;

int make_map(CUtensorMap* map, const void* ptr, int rows, int cols, int ld_elems, int box_rows) {
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
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};       // innermost first
  const cuuint64_t strides[1] = {(cuuint64_t)ld_elems * 2};              // bytes, for dim 1
  const cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    pdj_set_error("cuTensorMapEncodeTiled failed with CUresult %d (rows=%d cols=%d ld=%d box_rows=%d ptr=%p)", (int)r, rows,
                  cols, ld_elems, box_rows, ptr);
    return (int)r;
  }
  return 0;
}

bool g_attr_set = false;

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

int pdj_linear_bf16(const void* a, const void* b, void* out, const float* bias, const void* mask, int M, int N, int K,
                    int lda, int ldb, int ldo, int relu, cudaStream_t st) {
  if (M < 1 || N < 1 || K < 1 || lda % 8 != 0 || ldb % 8 != 0 || ((uintptr_t)a & 15) || ((uintptr_t)b & 15)) {
    pdj_set_error("pdj_linear_bf16: need M,N,K >= 1, 16-byte aligned operands and row pitches that are multiples of 8 elements "
                  "(lda=%d ldb=%d)", lda, ldb);
    return -1;
  }
  const int bn = pdj_linear_block_n(N);
  CUtensorMap ma, mb;
  if (int r = make_map(&ma, a, M, K, lda, BM)) return r;
  if (int r = make_map(&mb, b, N, K, ldb, bn)) return r;
  if (!g_attr_set) {
    PDJ_CUDA_CHECK(cudaFuncSetAttribute(linear_tcgen05_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    g_attr_set = true;
  }
  dim3 grid((M + BM - 1) / BM, (N + bn - 1) / bn);
  linear_tcgen05_kernel<<<grid, THREADS, SMEM_BYTES, st>>>(ma, mb, (__nv_bfloat16*)out, bias, (const __nv_bfloat16*)mask, M,
                                                            N, K, ldo, bn, relu);
  PDJ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // extern "C"
