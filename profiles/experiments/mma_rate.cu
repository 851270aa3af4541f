// Experiment (not product code): cycles per tcgen05.mma (cta_group::1, kind::f16, bf16, K = 16 per instruction) on
// one SM as a function of M, N and the operand majors, operands in shared memory (SS mode), 128-byte swizzle.
// One thread issues `iters` instructions back to back, then commits and waits.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I include -I paddle_operator_b200/csrc \
//             profiles/experiments/mma_rate.cu paddle_operator_b200/csrc/tma_util.cu -o profiles/experiments/bin/mma_rate -lcuda
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>

#include "tcgen05_common.cuh"

void pdj_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vfprintf(stderr, fmt, ap);
  va_end(ap);
  fprintf(stderr, "\n");
}
using namespace pdj_tc;

// a_mn / b_mn: operand is MN-major; rot: number of accumulator column groups the instructions rotate over;
// b_sw: swizzle code of B (2 = 128 B, 4 = 64 B); adv: advance the descriptors every instruction (k-slices of a tile)
__global__ void __launch_bounds__(128, 1)
rate_kernel(long long* out, int M, int N, int a_mn, int b_mn, int rot, int iters, int b_sw, int adv, int aoff = 0, int boff = 0) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t a_smem = base, b_smem = base + 65536, bar = base + 131072, slot = bar + 16;
  uint32_t* slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (slot - raw));
  for (uint32_t i = threadIdx.x; i < 131072 / 16; i += 128)
    asm volatile("st.shared.v4.b32 [%0], {%1,%1,%1,%1};" ::"r"(base + i * 16u), "r"(0u) : "memory");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *slot_ptr;
  if (threadIdx.x == 0) {
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) |
                           ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
    // descriptors are loop-invariant except for a k-slice advance; eight instructions per trip, two accumulators
    const uint32_t a_smem2 = a_smem + (uint32_t)aoff, b_smem2 = b_smem + (uint32_t)boff;
    const uint64_t da0 = a_mn ? umma_desc_ex(a_smem2, 16384, 1024, 2) : umma_desc_ex(a_smem2, 16, 1024, 2);
    const uint64_t db0 = b_mn ? (b_sw == 4 ? umma_desc_ex(b_smem2, 64, 512, 4) : umma_desc_ex(b_smem2, 16384, 1024, 2))
                              : umma_desc_ex(b_smem2, 16, 1024, 2);
    const uint64_t sa = adv ? (a_mn ? 128u : 2u) : 0u, sb = adv ? (b_mn ? (b_sw == 4 ? 64u : 128u) : 2u) : 0u;
    const uint32_t t1addr = tmem + (rot > 1 ? (uint32_t)N : 0u);
    const long long t0 = clock64();
    for (int i = 0; i < iters; i += 8) {
      umma_f16(tmem, da0, db0, idesc, 1u);
      umma_f16(t1addr, da0 + sa, db0 + sb, idesc, 1u);
      umma_f16(tmem, da0 + 2 * sa, db0 + 2 * sb, idesc, 1u);
      umma_f16(t1addr, da0 + 3 * sa, db0 + 3 * sb, idesc, 1u);
      umma_f16(tmem, da0, db0, idesc, 1u);
      umma_f16(t1addr, da0 + sa, db0 + sb, idesc, 1u);
      umma_f16(tmem, da0 + 2 * sa, db0 + 2 * sb, idesc, 1u);
      umma_f16(t1addr, da0 + 3 * sa, db0 + 3 * sb, idesc, 1u);
    }
    umma_commit(bar);
    mbar_wait(bar, 0);
    const long long t1 = clock64();
    out[0] = t1 - t0;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
}

// The wgrad inner loop: per k-step three N = 96 instructions sharing one A slice, B moved by 0 / P / 2P rows, three
// accumulators; `ksteps` k-steps per block, a commit per block, blocks cycling over `stages` slot bases.
__global__ void __launch_bounds__(128, 1)
wgrad_loop_kernel(long long* out, int blocks, int ksteps, int stages, int P, int commit_every) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t bar = base + 131072 + 65536, slot = bar + 64;
  uint32_t* slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (slot - raw));
  for (uint32_t i = threadIdx.x; i < (131072 + 65536) / 16; i += 128)
    asm volatile("st.shared.v4.b32 [%0], {%1,%1,%1,%1};" ::"r"(base + i * 16u), "r"(0u) : "memory");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (threadIdx.x == 0) {
    for (int i = 0; i < 8; ++i) mbar_init(bar + 8 * i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (threadIdx.x < 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(slot), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = warp_uniform(*slot_ptr);
  if (warp_uniform(threadIdx.x >> 5) == 0) {
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((96u >> 3) << 17) | ((128u >> 4) << 24);
    const uint64_t da0 = umma_desc_ex(0, 16384, 1024, 2), db0 = umma_desc_ex(0, 64, 512, 4);
    const uint32_t row_step = (uint32_t)P * 4u;
    const uint32_t slot_bytes = 32768 + 13312;
    const long long t0 = clock64();
    for (int b = 0; b < blocks; ++b) {
      const uint32_t s = (uint32_t)b % (uint32_t)stages;
      uint64_t da = da0 + (uint64_t)((base + s * slot_bytes) >> 4);
      uint64_t db = db0 + (uint64_t)((base + s * slot_bytes + 32768) >> 4);
      for (int ks = 0; ks < ksteps; ++ks) {
        if (elect_one()) {
          umma_f16(tmem, da, db, idesc, 1u);
          umma_f16(tmem + 96, da, db + row_step, idesc, 1u);
          umma_f16(tmem + 192, da, db + 2 * row_step, idesc, 1u);
        }
        da += 128;
        db += 64;
      }
      if (commit_every && (b % commit_every) == commit_every - 1 && elect_one()) umma_commit(bar + 8 * (s & 3));
      __syncwarp();
    }
    if (elect_one()) umma_commit(bar + 56);
    __syncwarp();
    mbar_wait(bar + 56, 0);
    if (threadIdx.x == 0) out[0] = clock64() - t0;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
}

int main() {
  cudaSetDevice(0);
  long long* d;
  cudaMalloc(&d, 8);
  const int smem = 1024 + 131072 + 64;
  cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const int iters = 4096;
  printf("%4s %4s %5s %5s %4s %4s %4s  cycles/mma  (floor max(M,128)*N/256)\n", "M", "N", "A", "B", "rot", "bsw", "adv");
  for (int M : {128, 64})
    for (int N : {32, 64, 96, 128, 256})
      for (int maj = 0; maj < 4; ++maj)
        for (int rot : {1, 2}) {
          const int a_mn = maj & 1, b_mn = maj >> 1;
          if (rot * N > 512) continue;
          for (int bsw : {2, 4}) {
            if (bsw == 4 && !(b_mn && (N == 32 || N == 96))) continue;
            for (int adv : {0, 1}) {
              rate_kernel<<<1, 128, smem>>>(d, M, N, a_mn, b_mn, rot, 64, bsw, adv);  // warm-up
              rate_kernel<<<1, 128, smem>>>(d, M, N, a_mn, b_mn, rot, iters, bsw, adv);
              cudaError_t e = cudaDeviceSynchronize();
              if (e != cudaSuccess) { printf("CUDA error %s\n", cudaGetErrorString(e)); return 1; }
              long long cyc;
              cudaMemcpy(&cyc, d, 8, cudaMemcpyDeviceToHost);
              printf("%4d %4d %5s %5s %4d %4d %4d  %8.1f   (%d)\n", M, N, a_mn ? "MN" : "K", b_mn ? "MN" : "K", rot, bsw, adv,
                     (double)cyc / iters, (M > 128 ? M : 128) * N / 256);
            }
          }
        }
  printf("row-shifted operands (M=128, N=96, A MN-major SW128, B MN-major SW64 with LBO = 64 B):\n");
  for (int aoff : {0, 128, 384})
    for (int boff : {0, 64, 384, 1920}) {
      rate_kernel<<<1, 128, smem>>>(d, 128, 96, 1, 1, 2, 64, 4, 1, aoff, boff);
      rate_kernel<<<1, 128, smem>>>(d, 128, 96, 1, 1, 2, iters, 4, 1, aoff, boff);
      cudaDeviceSynchronize();
      long long cyc;
      cudaMemcpy(&cyc, d, 8, cudaMemcpyDeviceToHost);
      printf("  A offset %4d B, B offset %4d B: %6.1f cycles/mma\n", aoff, boff, (double)cyc / iters);
    }
  printf("the wgrad inner loop (3 x N=96 per k-step, 8 k-steps per block, 48 blocks):\n");
  {
    const int smem2 = 1024 + 131072 + 65536 + 256;
    cudaFuncSetAttribute(wgrad_loop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem2);
    for (int ce : {0, 1, 4})
      for (int st : {1, 4}) {
        wgrad_loop_kernel<<<1, 128, smem2>>>(d, 4, 8, st, 30, ce);
        wgrad_loop_kernel<<<1, 128, smem2>>>(d, 48, 8, st, 30, ce);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("CUDA error %s\n", cudaGetErrorString(e)); return 1; }
        long long cyc;
        cudaMemcpy(&cyc, d, 8, cudaMemcpyDeviceToHost);
        printf("  commit every %d blocks, %d slot bases: %7.1f cycles/mma (%lld total)\n", ce, st, (double)cyc / (48 * 8 * 3), cyc);
      }
  }
  return 0;
}
