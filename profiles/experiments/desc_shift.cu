// Experiment (not product code): which shared-memory descriptor tricks does tcgen05.mma accept on sm_100a?
//   E1  K-major A operand whose start address is shifted by d rows of 128 B (d not a multiple of 8),
//       with the descriptor's base-offset field 0 or (addr >> 7) & 7.
//   E2  MN-major A operand (M = 64) shifted by d reduction rows.
//   E3  MN-major A operand with M = 128 = two 64-element column groups LBO bytes apart, shifted by d rows.
//   E4  cp.async.bulk.tensor im2col loads (3x3 window, pad 1, stride 1 and 2) against a host im2col.
// Build:  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I include -I paddle_operator_b200/csrc \
//              profiles/experiments/desc_shift.cu paddle_operator_b200/csrc/tma_util.cu -o gpurun_out/desc_shift -lcuda
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>

#include <vector>

#include "tcgen05_common.cuh"

void pdj_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vfprintf(stderr, fmt, ap);
  va_end(ap);
  fprintf(stderr, "\n");
}

using namespace pdj_tc;

__device__ __forceinline__ uint64_t mk_desc(uint32_t addr, uint32_t lbo, uint32_t sbo, uint32_t base_off) {
  uint64_t d = 0;
  d |= (uint64_t)((addr & 0x3FFFF) >> 4);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)(base_off & 7) << 49;
  d |= (uint64_t)2 << 61;
  return d;
}

// mode 1: K-major A (M=128) x K-major B (N=64), K=64
// mode 2: MN-major A (M=64)  x MN-major B (N=64), K=128
// mode 3: MN-major A (M=128, two groups) x MN-major B (N=64), K=128
__global__ void __launch_bounds__(128, 1)
mma_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b, float* out, int mode,
           int delta, int variant) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t a_smem = base;                 // up to 64 KiB
  const uint32_t b_smem = base + 65536;         // 16 KiB
  const uint32_t bar = b_smem + 16384, bar2 = bar + 8, slot = bar + 16;
  uint32_t* slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (slot - raw));
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    mbar_init(bar2, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot), "r"(128) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *slot_ptr;
  if (threadIdx.x == 0) {
    if (mode == 1) {
      mbar_expect_tx(bar, 2 * 16384 + 8192);
      tma_load_2d(a_smem, &map_a, 0, 0, bar);
      tma_load_2d(a_smem + 16384, &map_a, 0, 128, bar);
      tma_load_2d(b_smem, &map_b, 0, 0, bar);
    } else if (mode == 2) {
      mbar_expect_tx(bar, 2 * 16384 + 16384);
      tma_load_2d(a_smem, &map_a, 0, 0, bar);
      tma_load_2d(a_smem + 16384, &map_a, 0, 128, bar);
      tma_load_2d(b_smem, &map_b, 0, 0, bar);
    } else {
      mbar_expect_tx(bar, 4 * 16384 + 16384);
      tma_load_2d(a_smem, &map_a, 0, 0, bar);              // group 0, rows 0..127
      tma_load_2d(a_smem + 16384, &map_a, 0, 128, bar);    // group 0, rows 128..255
      tma_load_2d(a_smem + 32768, &map_a, 64, 0, bar);     // group 1
      tma_load_2d(a_smem + 49152, &map_a, 64, 128, bar);
      tma_load_2d(b_smem, &map_b, 0, 0, bar);
    }
    mbar_wait(bar, 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if (mode == 1) {
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((64u >> 3) << 17) | ((128u >> 4) << 24);
      for (int k = 0; k < 4; ++k) {
        const uint32_t aa = a_smem + (uint32_t)delta * 128u + k * 32;
        const uint64_t da = mk_desc(aa, 16, 1024, variant ? (aa >> 7) & 7 : 0);
        const uint64_t db = mk_desc(b_smem + k * 32, 16, 1024, 0);
        umma_f16(tmem, da, db, idesc, (uint32_t)(k != 0));
      }
    } else {
      const uint32_t m = mode == 2 ? 64u : 128u;
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((64u >> 3) << 17) | ((m >> 4) << 24);
      for (int k = 0; k < 8; ++k) {
        const uint32_t aa = a_smem + (uint32_t)delta * 128u + k * 2048;
        const uint64_t da = mk_desc(aa, mode == 3 ? 32768 : 16, 1024, variant ? (aa >> 7) & 7 : 0);
        const uint64_t db = mk_desc(b_smem + k * 2048, 16, 1024, 0);
        umma_f16(tmem, da, db, idesc, (uint32_t)(k != 0));
      }
    }
    umma_commit(bar2);
  }
  mbar_wait(bar2, 0);
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  for (int c0 = 0; c0 < 64; c0 += 16) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    int row = warp * 32 + lane;
    bool ok = true;
    if (mode == 2) { row = warp * 16 + lane; ok = lane < 16; }
    if (ok)
      for (int j = 0; j < 16; ++j) out[row * 64 + c0 + j] = __uint_as_float(r[j]);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(128) : "memory");
}

// E4: one im2col load of `pixels` x 64 channels, dumped unswizzled.
__global__ void im2col_kernel(const __grid_constant__ CUtensorMap map, __nv_bfloat16* out, int pixels, int c0, int w0, int h0,
                              int n0, int offw, int offh) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t bar = base + 32768;
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    mbar_expect_tx(bar, (uint32_t)pixels * 128u);
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.im2col.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6], {%7, %8};" ::
            "r"(base), "l"(&map), "r"(c0), "r"(w0), "r"(h0), "r"(n0), "r"(bar), "h"((uint16_t)offw), "h"((uint16_t)offh)
        : "memory");
    mbar_wait(bar, 0);
  }
  __syncthreads();
  const __nv_bfloat16* s = reinterpret_cast<const __nv_bfloat16*>(smem_raw + (base - raw));
  for (int i = threadIdx.x; i < pixels * 64; i += blockDim.x) out[i] = s[i];
}

static float bf(__nv_bfloat16 v) { return __bfloat162float(v); }

int main() {
  cudaSetDevice(0);
  cudaFree(0);
  srand(7);
  // ---- E1..E3 ----------------------------------------------------------------------------------
  const int AR = 256;
  std::vector<__nv_bfloat16> hA(AR * 64), hA3(AR * 128), hB(64 * 64), hB2(128 * 64);
  for (auto& v : hA) v = __float2bfloat16((rand() % 17 - 8) / 8.f);
  for (auto& v : hA3) v = __float2bfloat16((rand() % 17 - 8) / 8.f);
  for (auto& v : hB) v = __float2bfloat16((rand() % 17 - 8) / 8.f);
  for (auto& v : hB2) v = __float2bfloat16((rand() % 17 - 8) / 8.f);
  __nv_bfloat16 *dA, *dA3, *dB, *dB2;
  float* dOut;
  cudaMalloc(&dA, hA.size() * 2); cudaMalloc(&dA3, hA3.size() * 2); cudaMalloc(&dB, hB.size() * 2); cudaMalloc(&dB2, hB2.size() * 2);
  cudaMalloc(&dOut, 128 * 64 * 4);
  cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dA3, hA3.data(), hA3.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB2, hB2.data(), hB2.size() * 2, cudaMemcpyHostToDevice);
  CUtensorMap mA, mA3, mB, mB2;
  if (pdj_tc_make_map(&mA, dA, AR, 64, 64, 64, 128, true)) return 1;
  if (pdj_tc_make_map(&mA3, dA3, AR, 128, 128, 64, 128, true)) return 1;
  if (pdj_tc_make_map(&mB, dB, 64, 64, 64, 64, 64, true)) return 1;
  if (pdj_tc_make_map(&mB2, dB2, 128, 64, 64, 64, 128, true)) return 1;
  const int smem = 1024 + 65536 + 16384 + 64;
  cudaFuncSetAttribute(mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  std::vector<float> hOut(128 * 64);
  const int deltas[] = {0, 1, 2, 3, 5, 8, 9, 58};
  for (int mode = 1; mode <= 3; ++mode)
    for (int variant = 0; variant < 2; ++variant)
      for (int delta : deltas) {
        cudaMemset(dOut, 0, 128 * 64 * 4);
        mma_kernel<<<1, 128, smem>>>(mode == 3 ? mA3 : mA, mode == 1 ? mB : mB2, dOut, mode, delta, variant);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("E%d variant %d delta %d: CUDA error %s\n", mode, variant, delta, cudaGetErrorString(e)); return 2; }
        cudaMemcpy(hOut.data(), dOut, 128 * 64 * 4, cudaMemcpyDeviceToHost);
        double worst = 0;
        const int M = mode == 2 ? 64 : 128;
        for (int m = 0; m < M; ++m)
          for (int n = 0; n < 64; ++n) {
            double ref = 0;
            if (mode == 1) for (int k = 0; k < 64; ++k) ref += bf(hA[(m + delta) * 64 + k]) * bf(hB[n * 64 + k]);
            else if (mode == 2) for (int k = 0; k < 128; ++k) ref += bf(hA[(k + delta) * 64 + m]) * bf(hB2[k * 64 + n]);
            else for (int k = 0; k < 128; ++k) ref += bf(hA3[(k + delta) * 128 + m]) * bf(hB2[k * 64 + n]);
            worst = fmax(worst, fabs(ref - hOut[m * 64 + n]));
          }
        printf("E%d base_offset=%s delta=%2d  max_abs_err=%.4f %s\n", mode, variant ? "addr>>7&7" : "0", delta, worst,
               worst < 1e-2 ? "OK" : "WRONG");
      }
  // ---- E4: im2col ------------------------------------------------------------------------------
  typedef CUresult (*EncIm2col)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                const int*, const int*, cuuint32_t, cuuint32_t, const cuuint32_t*, CUtensorMapInterleave,
                                CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
  void* fp = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &fp, cudaEnableDefault, &q) != cudaSuccess || !fp) { printf("no im2col encoder\n"); return 3; }
  EncIm2col enc = (EncIm2col)fp;
  const int N = 5, H = 6, W = 6, C = 128;
  std::vector<__nv_bfloat16> hX((size_t)N * H * W * C);
  for (size_t i = 0; i < hX.size(); ++i) hX[i] = __float2bfloat16((float)((i * 7919) % 251) - 125.f);
  __nv_bfloat16 *dX, *dO;
  cudaMalloc(&dX, hX.size() * 2);
  cudaMalloc(&dO, 256 * 64 * 2);
  cudaMemcpy(dX, hX.data(), hX.size() * 2, cudaMemcpyHostToDevice);
  cudaFuncSetAttribute(im2col_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 1024 + 32768 + 64);
  for (int stride = 1; stride <= 2; ++stride) {
    const cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
    const cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
    const int lower[2] = {-1, -1}, upper[2] = {-1, -1};
    const cuuint32_t estr[4] = {1, (cuuint32_t)stride, (cuuint32_t)stride, 1};
    const int pixels = 128;
    CUtensorMap mX;
    CUresult r = enc(&mX, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, dX, dims, strides, lower, upper, 64, pixels, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("E4 stride %d: encode failed %d\n", stride, (int)r); continue; }
    const int Q = (W + 2 - 3) / stride + 1, P = (H + 2 - 3) / stride + 1;
    std::vector<__nv_bfloat16> hO(256 * 64);
    // start at output pixel m0 (not a row start), channel block 1, every tap
    const int m0 = stride == 1 ? 9 : 4;
    for (int rr = 0; rr < 3; ++rr)
      for (int ss = 0; ss < 3; ++ss) {
        const int n0 = m0 / (P * Q), p0 = (m0 % (P * Q)) / Q, q0 = m0 % Q;
        cudaMemset(dO, 0xff, 256 * 64 * 2);
        im2col_kernel<<<1, 128, 1024 + 32768 + 64>>>(mX, dO, pixels, 64, q0 * stride - 1, p0 * stride - 1, n0, ss, rr);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("E4 stride %d tap (%d,%d): CUDA error %s\n", stride, rr, ss, cudaGetErrorString(e)); return 4; }
        cudaMemcpy(hO.data(), dO, 256 * 64 * 2, cudaMemcpyDeviceToHost);
        int bad = 0, checked = 0;
        for (int i = 0; i < pixels; ++i) {
          const int m = m0 + i;
          const int n = m / (P * Q), p = (m % (P * Q)) / Q, qq = m % Q;
          const int y = p * stride - 1 + rr, x = qq * stride - 1 + ss;
          for (int c = 0; c < 64; ++c) {
            float ref = 0.f;
            if (n < N && y >= 0 && y < H && x >= 0 && x < W) ref = bf(hX[(((size_t)n * H + y) * W + x) * C + 64 + c]);
            if (n < N) { ++checked; if (bf(hO[i * 64 + c]) != ref) ++bad; }
          }
        }
        printf("E4 im2col stride=%d tap=(%d,%d): %d of %d elements wrong %s\n", stride, rr, ss, bad, checked, bad ? "WRONG" : "OK");
      }
  }
  return 0;
}
