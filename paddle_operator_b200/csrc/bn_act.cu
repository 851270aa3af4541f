// Fused training-mode BatchNorm (+residual add)(+ReLU) for NHWC bf16 tensors,
// forward and backward.  ResNet-50 at bf16 on B200 is HBM-bound on exactly
// these passes (the convolutions are tensor-core work in cuDNN); fusing the
// normalise / add / activation passes and their backward removes about a third
// of the activation traffic of a step.
//
// Reference parity: none - the reference launches a user image that runs
// Paddle's ResNet (deploy/examples/resnet.yaml:13-19); this is the compute
// layer BASELINE.json asks the runtime to own.
//
// x is viewed as [M, C] (M = N*H*W rows, C channels contiguous), C % 8 == 0,
// C <= 2048 so that 256 % min(C/8,32) == 0.
#include <cuda.h>
#include <stdlib.h>

#include "common.cuh"
#include "../../include/pdjob_b200.h"

namespace {

constexpr int kThreads = 256;

struct Tile {
  int cvt;        // channel-vectors (of 8) per block: min(C/8, cvt_max)
  int rows_par;   // rows processed in parallel by a block: 256 / cvt
  int cv;         // this thread's channel-vector index within the tensor
  int row0;       // this thread's first row
  int row_step;   // row stride between iterations
};

__device__ __forceinline__ Tile make_tile(int C, int cvt_max) {
  Tile t;
  const int c8 = C >> 3;
  t.cvt = c8 < cvt_max ? c8 : cvt_max;
  t.rows_par = kThreads / t.cvt;
  t.cv = blockIdx.y * t.cvt + (threadIdx.x % t.cvt);
  t.row0 = blockIdx.x * t.rows_par + threadIdx.x / t.cvt;
  t.row_step = gridDim.x * t.rows_par;
  return t;
}

__device__ __forceinline__ void unpack8(const uint4& u, float* f) {
  float2 a = pdj_bf2_to_f2(u.x), b = pdj_bf2_to_f2(u.y), c = pdj_bf2_to_f2(u.z), d = pdj_bf2_to_f2(u.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
__device__ __forceinline__ uint4 pack8(const float* f) {
  uint4 o;
  o.x = pdj_f2_to_bf2(f[0], f[1]); o.y = pdj_f2_to_bf2(f[2], f[3]);
  o.z = pdj_f2_to_bf2(f[4], f[5]); o.w = pdj_f2_to_bf2(f[6], f[7]);
  return o;
}
__device__ __forceinline__ uint4 ldg16(const void* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }
// Same load as an ordered (volatile) asm statement: the compiler keeps a batch of these in program
// order ahead of the arithmetic that consumes them, which is what puts several loads in flight per
// thread in the reduction kernels (with plain __ldg the loads were sunk next to their uses).
__device__ __forceinline__ uint4 ldg16_ordered(const void* p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p));
  return v;
}

// Block-level reduction of 16 per-thread accumulators over the threads that
// share a channel-vector, then one partial row per block:
//   partial[blockIdx.x][k][C] for k in {0,1}, channel = cv*8 + j.
// Returns true in the last block (per blockIdx.y) to arrive, after which all
// partials of this channel tile are visible to it.
__device__ __forceinline__ bool reduce_and_publish(const Tile& t, int C, const float* acc /*16*/,
                                                   float* __restrict__ partial, unsigned int* __restrict__ counter) {
  __shared__ float sm[16][kThreads + 1];
  __shared__ bool is_last;
#pragma unroll
  for (int k = 0; k < 16; ++k) sm[k][threadIdx.x] = acc[k];
  __syncthreads();
  const int nout = t.cvt * 16;
  for (int o = threadIdx.x; o < nout; o += kThreads) {
    const int k = o / t.cvt, lcv = o % t.cvt;
    float s = 0.f;
    for (int j = 0; j < t.rows_par; ++j) s += sm[k][lcv + j * t.cvt];
    const int ch = (blockIdx.y * t.cvt + lcv) * 8 + (k & 7);
    partial[((size_t)blockIdx.x * 2 + (k >> 3)) * C + ch] = s;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int ticket = atomicAdd(&counter[blockIdx.y], 1u);
    is_last = (ticket == gridDim.x - 1);
    if (is_last) counter[blockIdx.y] = 0;  // self-reset for the next launch
  }
  __syncthreads();
  if (is_last) __threadfence();
  return is_last;
}

// Last-block finalisation: sums partial[b][k][ch0 + c] over the nb row blocks
// with all 256 threads (256 / nch threads per channel, then a shared-memory
// fold).  Returns true in the threads that own a channel (threadIdx.x < nch),
// with the two sums in s and q.
__device__ __forceinline__ bool sum_partials(const float* partial, int C, int ch0, int nch, unsigned int nb, float& s,
                                             float& q) {
  __shared__ float red[2][kThreads];
  const int lanes = kThreads / nch;
  const int c = threadIdx.x % nch;
  const unsigned int sub = threadIdx.x / nch;
  float a = 0.f, b2 = 0.f;
#pragma unroll 4
  for (unsigned int b = sub; b < nb; b += lanes) {
    a += partial[((size_t)b * 2 + 0) * C + ch0 + c];
    b2 += partial[((size_t)b * 2 + 1) * C + ch0 + c];
  }
  red[0][threadIdx.x] = a;
  red[1][threadIdx.x] = b2;
  __syncthreads();
  if ((int)threadIdx.x >= nch) return false;
  for (int j = 1; j < lanes; ++j) {
    a += red[0][threadIdx.x + j * nch];
    b2 += red[1][threadIdx.x + j * nch];
  }
  s = a;
  q = b2;
  return true;
}

// Per-channel epilogue of the statistics pass: mean / rstd, the fused scale and shift of the
// apply pass, and the running statistics (unbiased variance, like torch).
__device__ __forceinline__ void finalize_channel(int ch, float s, float q, int M, int C, const float* __restrict__ gamma,
                                                 const float* __restrict__ beta, float eps, float momentum,
                                                 float* __restrict__ running_mean, float* __restrict__ running_var,
                                                 float* __restrict__ mean_rstd, float* __restrict__ scale_shift) {
  const float inv_m = 1.f / (float)M;
  const float mean = s * inv_m;
  float var = q * inv_m - mean * mean;
  var = var < 0.f ? 0.f : var;
  const float rstd = rsqrtf(var + eps);
  mean_rstd[ch] = mean;
  mean_rstd[C + ch] = rstd;
  const float sc = gamma[ch] * rstd;
  scale_shift[ch] = sc;
  scale_shift[C + ch] = beta[ch] - mean * sc;
  if (running_mean != nullptr) {
    const float unbiased = M > 1 ? var * ((float)M / (float)(M - 1)) : var;
    running_mean[ch] = (1.f - momentum) * running_mean[ch] + momentum * mean;
    running_var[ch] = (1.f - momentum) * running_var[ch] + momentum * unbiased;
  }
}

// ---- forward --------------------------------------------------------------
template <int U>
__global__ void __launch_bounds__(kThreads)
bn_stats_kernel(const __nv_bfloat16* __restrict__ x, int M, int C, const float* __restrict__ gamma,
                const float* __restrict__ beta, float eps, float momentum, float* __restrict__ running_mean,
                float* __restrict__ running_var, float* __restrict__ mean_rstd /*[2][C]*/,
                float* __restrict__ scale_shift /*[2][C]*/, float* __restrict__ partial,
                unsigned int* __restrict__ counter, int cvt_max) {
  const Tile t = make_tile(C, cvt_max);
  float acc[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) acc[k] = 0.f;
  const char* xp = reinterpret_cast<const char*>(x) + (size_t)t.cv * 16;
  const size_t row_bytes = (size_t)C * 2;
  int r = t.row0;
  const int rs = t.row_step;
  // U independent rows per iteration: U x 16 B in flight per thread.
  for (; r + (U - 1) * rs < M; r += U * rs) {
    uint4 u[U];
#pragma unroll
    for (int k = 0; k < U; ++k) u[k] = ldg16_ordered(xp + (size_t)(r + k * rs) * row_bytes);
#pragma unroll
    for (int k = 0; k < U; ++k) {
      float f[8];
      unpack8(u[k], f);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        acc[j] += f[j];
        acc[8 + j] = fmaf(f[j], f[j], acc[8 + j]);
      }
    }
  }
  for (; r < M; r += rs) {
    float f[8];
    unpack8(ldg16(xp + (size_t)r * row_bytes), f);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      acc[j] += f[j];
      acc[8 + j] = fmaf(f[j], f[j], acc[8 + j]);
    }
  }
  if (!reduce_and_publish(t, C, acc, partial, counter)) return;
  // Last block of this channel tile: finalise its channels.
  const int nch = t.cvt * 8;
  float s, q;
  if (sum_partials(partial, C, blockIdx.y * nch, nch, gridDim.x, s, q)) {
    finalize_channel(blockIdx.y * nch + threadIdx.x, s, q, M, C, gamma, beta, eps, momentum, running_mean, running_var,
                     mean_rstd, scale_shift);
  }
}

// Statistics already reduced to `nb` partial rows by the producer of x (the 1x1-convolution GEMM
// of conv_tcgen05.cu writes them from its epilogue): only the finalisation is left.  8 channels per
// block, 32 threads per channel.
__global__ void __launch_bounds__(kThreads)
bn_finalize_kernel(const float* __restrict__ partial, unsigned int nb, int M, int C, const float* __restrict__ gamma,
                   const float* __restrict__ beta, float eps, float momentum, float* __restrict__ running_mean,
                   float* __restrict__ running_var, float* __restrict__ mean_rstd, float* __restrict__ scale_shift) {
  float s, q;
  if (sum_partials(partial, C, blockIdx.x * 8, 8, nb, s, q))
    finalize_channel(blockIdx.x * 8 + threadIdx.x, s, q, M, C, gamma, beta, eps, momentum, running_mean, running_var,
                     mean_rstd, scale_shift);
}

template <bool RELU, bool RES>
__global__ void __launch_bounds__(kThreads)
bn_apply_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ res,
                __nv_bfloat16* __restrict__ y, size_t n_vec, int C, const float* __restrict__ scale_shift,
                uint8_t* __restrict__ gate) {
  // gate (optional, RELU only): one byte per 8-element vector, bit j = (y_j > 0).  The backward
  // kernels read it instead of y: 1/16 of the bytes.
  const int c8 = C >> 3;
  const int cv = threadIdx.x % c8;  // 256 % c8 == 0 and the grid stride is a multiple of 256
  float sc[8], sh[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    sc[j] = scale_shift[cv * 8 + j];
    sh[j] = scale_shift[C + cv * 8 + j];
  }
  const size_t stride = (size_t)gridDim.x * kThreads;
  const uint4* xv = reinterpret_cast<const uint4*>(x);
  const uint4* rv = reinterpret_cast<const uint4*>(res);
  uint4* yv = reinterpret_cast<uint4*>(y);
  size_t i = (size_t)blockIdx.x * kThreads + threadIdx.x;
  for (; i + stride < n_vec; i += 2 * stride) {
    const uint4 a0 = ldg16(xv + i), a1 = ldg16(xv + i + stride);
    uint4 b0, b1;
    if (RES) { b0 = ldg16(rv + i); b1 = ldg16(rv + i + stride); }
    float f[8], g[8], p[8], q[8];
    unpack8(a0, f); unpack8(a1, g);
    if (RES) { unpack8(b0, p); unpack8(b1, q); }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      f[j] = fmaf(f[j], sc[j], sh[j]);
      g[j] = fmaf(g[j], sc[j], sh[j]);
      if (RES) { f[j] += p[j]; g[j] += q[j]; }
      if (RELU) { f[j] = fmaxf(f[j], 0.f); g[j] = fmaxf(g[j], 0.f); }
    }
    yv[i] = pack8(f);
    yv[i + stride] = pack8(g);
    if (RELU && gate != nullptr) {
      uint32_t b0 = 0, b1 = 0;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        b0 |= (f[j] > 0.f ? 1u : 0u) << j;
        b1 |= (g[j] > 0.f ? 1u : 0u) << j;
      }
      gate[i] = (uint8_t)b0;
      gate[i + stride] = (uint8_t)b1;
    }
  }
  if (i < n_vec) {
    float f[8], p[8];
    unpack8(ldg16(xv + i), f);
    if (RES) unpack8(ldg16(rv + i), p);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      f[j] = fmaf(f[j], sc[j], sh[j]);
      if (RES) f[j] += p[j];
      if (RELU) f[j] = fmaxf(f[j], 0.f);
    }
    yv[i] = pack8(f);
    if (RELU && gate != nullptr) {
      uint32_t b0 = 0;
#pragma unroll
      for (int j = 0; j < 8; ++j) b0 |= (f[j] > 0.f ? 1u : 0u) << j;
      gate[i] = (uint8_t)b0;
    }
  }
}

// ---- backward -------------------------------------------------------------
// dz = dy * (y > 0) when RELU.  Pass 1: sum(dz), sum(dz * xhat) per channel;
// the last block turns them into dgamma/dbeta and the three coefficients of
//   dx = A*dz + B*x + Cc.
template <bool RELU, int U>
__global__ void __launch_bounds__(kThreads)
bn_bwd_reduce_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ y,
                     const __nv_bfloat16* __restrict__ x, int M, int C, const float* __restrict__ gamma,
                     const float* __restrict__ mean_rstd, float* __restrict__ dgamma_dbeta /*[2][C] fp32*/,
                     __nv_bfloat16* __restrict__ dgb_bf16 /*dgamma [C] or null*/, __nv_bfloat16* __restrict__ db_bf16 /*dbeta [C]*/,
                     float* __restrict__ coef /*[3][C]*/, float* __restrict__ partial,
                     unsigned int* __restrict__ counter, int cvt_max, const uint8_t* __restrict__ gate) {
  const Tile t = make_tile(C, cvt_max);
  float mu[8], rs[8], acc[16];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    mu[j] = mean_rstd[t.cv * 8 + j];
    rs[j] = mean_rstd[C + t.cv * 8 + j];
    acc[j] = 0.f;
    acc[8 + j] = 0.f;
  }
  const size_t row_bytes = (size_t)C * 2;
  const int c8g = C >> 3;
  const size_t coff = (size_t)t.cv * 16;
  const char* dyp = reinterpret_cast<const char*>(dy) + coff;
  const char* yp = reinterpret_cast<const char*>(y) + coff;
  const char* xp = reinterpret_cast<const char*>(x) + coff;
  // sum(dz * xhat) = rstd * (sum(dz * x) - mean * sum(dz)): accumulate the raw
  // sums here and fold mean / rstd in once per channel at the end.
  int r = t.row0;
  const int rstep = t.row_step;
  for (; U == 2 && r + rstep < M; r += 2 * rstep) {
    const size_t o0 = (size_t)r * row_bytes, o1 = (size_t)(r + rstep) * row_bytes;
    const uint4 ud0 = ldg16_ordered(dyp + o0), ux0 = ldg16_ordered(xp + o0), ud1 = ldg16_ordered(dyp + o1),
                ux1 = ldg16_ordered(xp + o1);
    uint4 uy0, uy1;
    uint32_t gb0 = 0, gb1 = 0;
    if (RELU) {
      if (gate != nullptr) {
        gb0 = __ldg(gate + (size_t)r * c8g + t.cv);
        gb1 = __ldg(gate + (size_t)(r + rstep) * c8g + t.cv);
      } else {
        uy0 = ldg16(yp + o0);
        uy1 = ldg16(yp + o1);
      }
    }
    float d[8], xx[8], yy[8];
    unpack8(ud0, d); unpack8(ux0, xx);
    if (RELU) {
      if (gate != nullptr) {
#pragma unroll
        for (int j = 0; j < 8; ++j) yy[j] = (float)((gb0 >> j) & 1u);
      } else {
        unpack8(uy0, yy);
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float dz = RELU ? (yy[j] > 0.f ? d[j] : 0.f) : d[j];
      acc[j] += dz;
      acc[8 + j] = fmaf(dz, xx[j] - mu[j], acc[8 + j]);
    }
    unpack8(ud1, d); unpack8(ux1, xx);
    if (RELU) {
      if (gate != nullptr) {
#pragma unroll
        for (int j = 0; j < 8; ++j) yy[j] = (float)((gb1 >> j) & 1u);
      } else {
        unpack8(uy1, yy);
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float dz = RELU ? (yy[j] > 0.f ? d[j] : 0.f) : d[j];
      acc[j] += dz;
      acc[8 + j] = fmaf(dz, xx[j] - mu[j], acc[8 + j]);
    }
  }
  for (; r < M; r += rstep) {
    const size_t o = (size_t)r * row_bytes;
    const uint4 ud = ldg16(dyp + o), ux = ldg16(xp + o);
    uint4 uy;
    uint32_t gb = 0;
    if (RELU) {
      if (gate != nullptr) gb = __ldg(gate + (size_t)r * c8g + t.cv);
      else uy = ldg16(yp + o);
    }
    float d[8], xx[8], yy[8];
    unpack8(ud, d);
    unpack8(ux, xx);
    if (RELU) {
      if (gate != nullptr) {
#pragma unroll
        for (int j = 0; j < 8; ++j) yy[j] = (float)((gb >> j) & 1u);
      } else {
        unpack8(uy, yy);
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float dz = RELU ? (yy[j] > 0.f ? d[j] : 0.f) : d[j];
      acc[j] += dz;
      acc[8 + j] = fmaf(dz, xx[j] - mu[j], acc[8 + j]);
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[8 + j] *= rs[j];
  if (!reduce_and_publish(t, C, acc, partial, counter)) return;
  const int nch = t.cvt * 8;
  const float inv_m = 1.f / (float)M;
  float sdz, sdzx;
  if (sum_partials(partial, C, blockIdx.y * nch, nch, gridDim.x, sdz, sdzx)) {
    const int ch = blockIdx.y * nch + threadIdx.x;
    dgamma_dbeta[ch] = sdzx;
    dgamma_dbeta[C + ch] = sdz;
    if (dgb_bf16 != nullptr) {
      dgb_bf16[ch] = __float2bfloat16(sdzx);
      db_bf16[ch] = __float2bfloat16(sdz);
    }
    const float r = mean_rstd[C + ch], m = mean_rstd[ch];
    const float A = gamma[ch] * r;
    const float B = -A * r * sdzx * inv_m;
    coef[ch] = A;
    coef[C + ch] = B;
    coef[2 * C + ch] = -B * m - A * sdz * inv_m;
  }
}

template <bool RELU, bool WRITE_DZ>
__global__ void __launch_bounds__(kThreads)
bn_bwd_apply_kernel(const __nv_bfloat16* __restrict__ dy, const __nv_bfloat16* __restrict__ y,
                    const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ dx,
                    __nv_bfloat16* __restrict__ dres, size_t n_vec, int C, const float* __restrict__ coef,
                    const uint8_t* __restrict__ gate) {
  const int c8 = C >> 3;
  const int cv = threadIdx.x % c8;
  float A[8], B[8], Cc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    A[j] = coef[cv * 8 + j];
    B[j] = coef[C + cv * 8 + j];
    Cc[j] = coef[2 * C + cv * 8 + j];
  }
  const size_t stride = (size_t)gridDim.x * kThreads;
  const uint4* dyv = reinterpret_cast<const uint4*>(dy);
  const uint4* yv = reinterpret_cast<const uint4*>(y);
  const uint4* xv = reinterpret_cast<const uint4*>(x);
  uint4* dxv = reinterpret_cast<uint4*>(dx);
  uint4* drv = reinterpret_cast<uint4*>(dres);
  for (size_t i = (size_t)blockIdx.x * kThreads + threadIdx.x; i < n_vec; i += stride) {
    const uint4 ud = ldg16(dyv + i), ux = ldg16(xv + i);
    uint4 uy;
    uint32_t gb = 0;
    if (RELU) {
      if (gate != nullptr) gb = __ldg(gate + i);
      else uy = ldg16(yv + i);
    }
    float d[8], xx[8], yy[8], o[8];
    unpack8(ud, d);
    unpack8(ux, xx);
    if (RELU) {
      if (gate != nullptr) {
#pragma unroll
        for (int j = 0; j < 8; ++j) yy[j] = (float)((gb >> j) & 1u);
      } else {
        unpack8(uy, yy);
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (RELU) d[j] = yy[j] > 0.f ? d[j] : 0.f;
      o[j] = fmaf(A[j], d[j], fmaf(B[j], xx[j], Cc[j]));
    }
    dxv[i] = pack8(o);
    if (WRITE_DZ) drv[i] = pack8(d);
  }
}

// ---- TMA-streamed reductions ----------------------------------------------------------------------
// The two reduction passes (statistics; backward sums) are pure streaming reads.  With ordinary loads
// their bandwidth is capped by how many 16-byte loads the compiler keeps in flight per thread (about
// 70 % of the HBM rate measured by ncu).  Here one thread streams [128 rows x 64 channels] boxes into a
// shared-memory ring with cp.async.bulk.tensor (TMA), so ~100 KB per CTA are in flight regardless of
// instruction scheduling, and the 256 threads reduce each box from shared memory.  Same partials /
// last-block finalisation as above, so results are bit-identical across runs.
constexpr int TR = 128;  // rows per box
constexpr int TC = 64;   // channels per box (128 bytes: a warp reads four full rows, conflict-free)
constexpr int TBOX_BYTES = TR * TC * 2;

__device__ __forceinline__ uint32_t bn_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void bn_mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void bn_mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bn_mbar_wait(uint32_t bar, uint32_t parity) {
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
__device__ __forceinline__ void bn_tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(dst),
      "l"(map), "r"(c0), "r"(c1), "r"(bar)
      : "memory");
}
__device__ __forceinline__ uint4 bn_lds16(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}

// The block-level fold of reduce_and_publish, with its scratch in caller-provided shared memory (the
// streaming ring, idle by then) instead of a static array.
__device__ __forceinline__ bool reduce_and_publish_dyn(const Tile& t, int C, const float* acc /*16*/,
                                                       float* __restrict__ partial, unsigned int* __restrict__ counter,
                                                       float* sm /*[16][kThreads + 1]*/) {
  __shared__ bool is_last;
#pragma unroll
  for (int k = 0; k < 16; ++k) sm[k * (kThreads + 1) + threadIdx.x] = acc[k];
  __syncthreads();
  const int nout = t.cvt * 16;
  for (int o = threadIdx.x; o < nout; o += kThreads) {
    const int k = o / t.cvt, lcv = o % t.cvt;
    float s = 0.f;
    for (int j = 0; j < t.rows_par; ++j) s += sm[k * (kThreads + 1) + lcv + j * t.cvt];
    const int ch = (blockIdx.y * t.cvt + lcv) * 8 + (k & 7);
    partial[((size_t)blockIdx.x * 2 + (k >> 3)) * C + ch] = s;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int ticket = atomicAdd(&counter[blockIdx.y], 1u);
    is_last = (ticket == gridDim.x - 1);
    if (is_last) counter[blockIdx.y] = 0;
  }
  __syncthreads();
  if (is_last) __threadfence();
  return is_last;
}

__device__ __forceinline__ Tile make_tile_tma() {
  Tile t;
  t.cvt = TC / 8;                       // 8 channel-vectors per block
  t.rows_par = kThreads / t.cvt;        // 32 rows per pass
  t.cv = blockIdx.y * t.cvt + (threadIdx.x % t.cvt);
  t.row0 = threadIdx.x / t.cvt;
  t.row_step = t.rows_par;
  return t;
}

// NT tensors per stage (1: statistics, 2: backward sums).  STAGES boxes per tensor in the ring.
template <int NT, int STAGES>
struct TmaRing {
  uint32_t base, bars;
  __device__ __forceinline__ void init(uint8_t* raw) {
    const uint32_t r = bn_smem_u32(raw);
    base = (r + 1023u) & ~1023u;
    bars = base + STAGES * NT * TBOX_BYTES;
    if (threadIdx.x == 0) {
      for (int s = 0; s < STAGES; ++s) bn_mbar_init(bars + 8 * s, 1);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
  }
  __device__ __forceinline__ uint32_t slot(int s, int tensor) const { return base + (uint32_t)(s * NT + tensor) * TBOX_BYTES; }
  __device__ __forceinline__ uint32_t bar(int s) const { return bars + 8 * s; }
};

template <int STAGES>
__global__ void __launch_bounds__(kThreads, 2)
bn_stats_tma_kernel(const __grid_constant__ CUtensorMap tma_x, int M, int C, const float* __restrict__ gamma,
                    const float* __restrict__ beta, float eps, float momentum, float* __restrict__ running_mean,
                    float* __restrict__ running_var, float* __restrict__ mean_rstd, float* __restrict__ scale_shift,
                    float* __restrict__ partial, unsigned int* __restrict__ counter) {
  extern __shared__ uint8_t bn_dyn_smem[];
  TmaRing<1, STAGES> ring;
  ring.init(bn_dyn_smem);
  const Tile t = make_tile_tma();
  const int ntiles = (M + TR - 1) / TR;
  const int c0 = blockIdx.y * TC;
  float acc[16];
#pragma unroll
  for (int k = 0; k < 16; ++k) acc[k] = 0.f;
  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      const int rt = blockIdx.x + s * gridDim.x;
      if (rt < ntiles) {
        bn_mbar_expect_tx(ring.bar(s), TBOX_BYTES);
        bn_tma_load_2d(ring.slot(s, 0), &tma_x, c0, rt * TR, ring.bar(s));
      }
    }
  }
  const uint32_t toff = (uint32_t)t.row0 * (TC * 2) + (uint32_t)(threadIdx.x % t.cvt) * 16u;
  int it = 0;
  for (int rt = blockIdx.x; rt < ntiles; rt += gridDim.x, ++it) {
    const int s = it % STAGES;
    bn_mbar_wait(ring.bar(s), (uint32_t)(it / STAGES) & 1u);
    const uint32_t p = ring.slot(s, 0) + toff;
#pragma unroll
    for (int k = 0; k < TR / 32; ++k) {
      float f[8];
      unpack8(bn_lds16(p + (uint32_t)k * 32u * (TC * 2)), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        acc[j] += f[j];
        acc[8 + j] = fmaf(f[j], f[j], acc[8 + j]);
      }
    }
    __syncthreads();  // everyone is done with this slot
    if (threadIdx.x == 0) {
      const int nrt = rt + STAGES * gridDim.x;
      if (nrt < ntiles) {
        bn_mbar_expect_tx(ring.bar(s), TBOX_BYTES);
        bn_tma_load_2d(ring.slot(s, 0), &tma_x, c0, nrt * TR, ring.bar(s));
      }
    }
  }
  __syncthreads();
  float* scratch = reinterpret_cast<float*>(bn_dyn_smem + (ring.base - bn_smem_u32(bn_dyn_smem)));
  if (!reduce_and_publish_dyn(t, C, acc, partial, counter, scratch)) return;
  const int nch = t.cvt * 8;
  float sum, sq;
  if (sum_partials(partial, C, blockIdx.y * nch, nch, gridDim.x, sum, sq))
    finalize_channel(blockIdx.y * nch + threadIdx.x, sum, sq, M, C, gamma, beta, eps, momentum, running_mean, running_var,
                     mean_rstd, scale_shift);
}

template <bool RELU, int STAGES>
__global__ void __launch_bounds__(kThreads, 2)
bn_bwd_reduce_tma_kernel(const __grid_constant__ CUtensorMap tma_dy, const __grid_constant__ CUtensorMap tma_x, int M, int C,
                         const float* __restrict__ gamma, const float* __restrict__ mean_rstd,
                         float* __restrict__ dgamma_dbeta, __nv_bfloat16* __restrict__ dgb_bf16, __nv_bfloat16* __restrict__ db_bf16,
                         float* __restrict__ coef,
                         float* __restrict__ partial, unsigned int* __restrict__ counter, const uint8_t* __restrict__ gate) {
  extern __shared__ uint8_t bn_dyn_smem[];
  TmaRing<2, STAGES> ring;
  ring.init(bn_dyn_smem);
  const Tile t = make_tile_tma();
  const int ntiles = (M + TR - 1) / TR;
  const int c0 = blockIdx.y * TC;
  const int c8g = C >> 3;
  float mu[8], rs[8], acc[16];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    mu[j] = mean_rstd[t.cv * 8 + j];
    rs[j] = mean_rstd[C + t.cv * 8 + j];
    acc[j] = 0.f;
    acc[8 + j] = 0.f;
  }
  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      const int rt = blockIdx.x + s * gridDim.x;
      if (rt < ntiles) {
        bn_mbar_expect_tx(ring.bar(s), 2 * TBOX_BYTES);
        bn_tma_load_2d(ring.slot(s, 0), &tma_dy, c0, rt * TR, ring.bar(s));
        bn_tma_load_2d(ring.slot(s, 1), &tma_x, c0, rt * TR, ring.bar(s));
      }
    }
  }
  const uint32_t toff = (uint32_t)t.row0 * (TC * 2) + (uint32_t)(threadIdx.x % t.cvt) * 16u;
  // gate bits of this thread's four rows of a box, fetched one box ahead
  uint32_t gb[TR / 32], gnext[TR / 32];
  auto load_gates = [&](int rt, uint32_t* g) {
#pragma unroll
    for (int k = 0; k < TR / 32; ++k) {
      const int row = rt * TR + t.row0 + 32 * k;
      g[k] = (RELU && rt < ntiles && row < M) ? (uint32_t)__ldg(gate + (size_t)row * c8g + t.cv) : 0u;
    }
  };
  if (RELU) load_gates(blockIdx.x, gnext);
  int it = 0;
  for (int rt = blockIdx.x; rt < ntiles; rt += gridDim.x, ++it) {
    const int s = it % STAGES;
    if (RELU) {
#pragma unroll
      for (int k = 0; k < TR / 32; ++k) gb[k] = gnext[k];
      load_gates(rt + gridDim.x, gnext);
    }
    bn_mbar_wait(ring.bar(s), (uint32_t)(it / STAGES) & 1u);
    const uint32_t pd = ring.slot(s, 0) + toff, px = ring.slot(s, 1) + toff;
#pragma unroll
    for (int k = 0; k < TR / 32; ++k) {
      float d[8], xx[8];
      unpack8(bn_lds16(pd + (uint32_t)k * 32u * (TC * 2)), d);
      unpack8(bn_lds16(px + (uint32_t)k * 32u * (TC * 2)), xx);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float dz = RELU ? (((gb[k] >> j) & 1u) ? d[j] : 0.f) : d[j];
        acc[j] += dz;
        acc[8 + j] = fmaf(dz, xx[j] - mu[j], acc[8 + j]);
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      const int nrt = rt + STAGES * gridDim.x;
      if (nrt < ntiles) {
        bn_mbar_expect_tx(ring.bar(s), 2 * TBOX_BYTES);
        bn_tma_load_2d(ring.slot(s, 0), &tma_dy, c0, nrt * TR, ring.bar(s));
        bn_tma_load_2d(ring.slot(s, 1), &tma_x, c0, nrt * TR, ring.bar(s));
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[8 + j] *= rs[j];
  __syncthreads();
  float* scratch = reinterpret_cast<float*>(bn_dyn_smem + (ring.base - bn_smem_u32(bn_dyn_smem)));
  if (!reduce_and_publish_dyn(t, C, acc, partial, counter, scratch)) return;
  const int nch = t.cvt * 8;
  const float inv_m = 1.f / (float)M;
  float sdz, sdzx;
  if (sum_partials(partial, C, blockIdx.y * nch, nch, gridDim.x, sdz, sdzx)) {
    const int ch = blockIdx.y * nch + threadIdx.x;
    dgamma_dbeta[ch] = sdzx;
    dgamma_dbeta[C + ch] = sdz;
    if (dgb_bf16 != nullptr) {
      dgb_bf16[ch] = __float2bfloat16(sdzx);
      db_bf16[ch] = __float2bfloat16(sdz);
    }
    const float r = mean_rstd[C + ch], m = mean_rstd[ch];
    const float A = gamma[ch] * r;
    const float B = -A * r * sdzx * inv_m;
    coef[ch] = A;
    coef[C + ch] = B;
    coef[2 * C + ch] = -B * m - A * sdz * inv_m;
  }
}

int env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return (v && *v) ? atoi(v) : dflt;
}
// Tuning knobs (read once): row blocks per SM for the two reduction kernels,
// minimum rows per thread, rows in flight per thread in the statistics pass.
int knob_blocks_per_sm() { static int v = env_int("PDJ_BN_BLOCKS_PER_SM", 4); return v; }
int knob_min_rows() { static int v = env_int("PDJ_BN_MIN_ROWS", 4); return v; }
int knob_stats_unroll() { static int v = env_int("PDJ_BN_STATS_UNROLL", 1); return v; }
int knob_bwd_unroll() { static int v = env_int("PDJ_BN_BWD_UNROLL", 1); return v; }
int knob_cvt() {
  static int v = env_int("PDJ_BN_CVT", 4);
  return (v == 1 || v == 2 || v == 4 || v == 8 || v == 16 || v == 32) ? v : 4;
}

// Channel-vectors per block actually used for width C.  The "last block" tickets are one counter per
// channel tile (blockIdx.y) and only 64 of them exist (the first 256 bytes of the workspace, and of the
// built-in array of pdj_bn_fwd_noalloc), so narrow tiles chosen with PDJ_BN_CVT=1|2 are widened for
// C > 512 / 1024 until the channel-tile count is at most 64.  Used for the grid AND for the kernels'
// cvt_max argument, which must agree.
constexpr int kMaxChannelTiles = 64;
int eff_cvt(int C) {
  const int c8 = C / 8;
  int cvt = knob_cvt();
  while (c8 / cvt > kMaxChannelTiles) cvt *= 2;
  return c8 < cvt ? c8 : cvt;
}

// TMA-streamed reductions: on by default for C % 64 == 0 (PDJ_BN_TMA=0 restores the plain-load kernels).
int knob_tma() { static int v = env_int("PDJ_BN_TMA", 1); return v; }
constexpr int kStatsStages = 4, kBwdStages = 3;
constexpr int kStatsSmem = 1024 + kStatsStages * TBOX_BYTES + 64;
constexpr int kBwdSmem = 1024 + kBwdStages * 2 * TBOX_BYTES + 64;

bool tma_path(int M, int C) { return knob_tma() != 0 && C % TC == 0 && M >= TR; }

// Row-block count of the streamed kernels: two resident CTAs per SM over all channel tiles.
int tma_grid_rows(int M, int C, int sms) {
  const int gy = C / TC;
  const int ntiles = (M + TR - 1) / TR;
  int gx = (2 * sms + gy - 1) / gy;
  if (gx > ntiles) gx = ntiles;
  if (gx < 1) gx = 1;
  if (gx > PDJ_BN_MAX_ROW_BLOCKS) gx = PDJ_BN_MAX_ROW_BLOCKS;
  return gx;
}

int ensure_tma_attrs() {
  static bool done = false;
  if (done) return 0;
  PDJ_CUDA_CHECK(cudaFuncSetAttribute(bn_stats_tma_kernel<kStatsStages>, cudaFuncAttributeMaxDynamicSharedMemorySize, kStatsSmem));
  PDJ_CUDA_CHECK(cudaFuncSetAttribute(bn_bwd_reduce_tma_kernel<true, kBwdStages>, cudaFuncAttributeMaxDynamicSharedMemorySize, kBwdSmem));
  PDJ_CUDA_CHECK(cudaFuncSetAttribute(bn_bwd_reduce_tma_kernel<false, kBwdStages>, cudaFuncAttributeMaxDynamicSharedMemorySize, kBwdSmem));
  done = true;
  return 0;
}

template <typename... A>
void launch_stats(dim3 g, cudaStream_t st, A... a) {
  switch (knob_stats_unroll()) {
    case 1: bn_stats_kernel<1><<<g, kThreads, 0, st>>>(a...); break;
    case 4: bn_stats_kernel<4><<<g, kThreads, 0, st>>>(a...); break;
    default: bn_stats_kernel<2><<<g, kThreads, 0, st>>>(a...); break;
  }
}

int grid_rows(int M, int C, int sm_count) {
  const int c8 = C / 8;
  const int cvt = eff_cvt(C);
  const int gy = c8 / cvt;
  const int rows_par = kThreads / cvt;
  int gx = (knob_blocks_per_sm() * sm_count + gy - 1) / gy;
  const int mr = knob_min_rows();
  const int max_gx = (M + rows_par * mr - 1) / (rows_par * mr);  // at least `mr` rows per thread
  if (gx > max_gx) gx = max_gx;
  if (gx < 1) gx = 1;
  if (gx > PDJ_BN_MAX_ROW_BLOCKS) gx = PDJ_BN_MAX_ROW_BLOCKS;
  return gx;
}

int check_shape(int M, int C) {
  const int c8 = C / 8;
  if (C % 8 != 0 || c8 < 1 || c8 > 256 || (c8 & (c8 - 1)) != 0 || M < 1) {
    pdj_set_error("bn: unsupported shape M=%d C=%d (C must be 8..2048, C/8 a power of two)", M, C);
    return -1;
  }
  return 0;
}

int g_sm_count = 0;
int sm_count() {
  if (g_sm_count == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_sm_count, cudaDevAttrMultiProcessorCount, dev);
    if (g_sm_count <= 0) g_sm_count = 148;
  }
  return g_sm_count;
}

int flat_grid(size_t n_vec) {
  size_t want = (n_vec + (size_t)kThreads * 4 - 1) / ((size_t)kThreads * 4);
  const size_t cap = (size_t)sm_count() * 8;
  if (want > cap) want = cap;
  if (want < 1) want = 1;
  return (int)want;
}

int launch_apply(const void* x, const void* res, void* y, uint8_t* gate_out, int M, int C, const float* scale_shift,
                 int relu, cudaStream_t st) {
  const size_t n_vec = (size_t)M * (C / 8);
  const int fg = flat_grid(n_vec);
  const __nv_bfloat16* xb = (const __nv_bfloat16*)x;
  const __nv_bfloat16* rb = (const __nv_bfloat16*)res;
  __nv_bfloat16* yb = (__nv_bfloat16*)y;
  if (relu && res) bn_apply_kernel<true, true><<<fg, kThreads, 0, st>>>(xb, rb, yb, n_vec, C, scale_shift, gate_out);
  else if (relu) bn_apply_kernel<true, false><<<fg, kThreads, 0, st>>>(xb, rb, yb, n_vec, C, scale_shift, gate_out);
  else if (res) bn_apply_kernel<false, true><<<fg, kThreads, 0, st>>>(xb, rb, yb, n_vec, C, scale_shift, gate_out);
  else bn_apply_kernel<false, false><<<fg, kThreads, 0, st>>>(xb, rb, yb, n_vec, C, scale_shift, gate_out);
  PDJ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // namespace

extern "C" {

size_t pdj_bn_workspace_bytes() {
  // partials [PDJ_BN_MAX_ROW_BLOCKS][2][2048] fp32 + 64 counters
  return (size_t)PDJ_BN_MAX_ROW_BLOCKS * 2 * 2048 * sizeof(float) + 256;
}

int pdj_bn_fwd_gate(const void* x, const void* res, void* y, void* gate, int M, int C, const float* gamma,
                    const float* beta, float eps, float momentum, float* running_mean, float* running_var,
                    float* mean_rstd, float* scale_shift, int relu, void* workspace, cudaStream_t st);

int pdj_bn_fwd(const void* x, const void* res, void* y, int M, int C, const float* gamma, const float* beta,
               float eps, float momentum, float* running_mean, float* running_var, float* mean_rstd,
               float* scale_shift, int relu, void* workspace, cudaStream_t st) {
  return pdj_bn_fwd_gate(x, res, y, nullptr, M, C, gamma, beta, eps, momentum, running_mean, running_var, mean_rstd,
                         scale_shift, relu, workspace, st);
}

int pdj_bn_fwd_gate(const void* x, const void* res, void* y, void* gate, int M, int C, const float* gamma,
                    const float* beta, float eps, float momentum, float* running_mean, float* running_var,
                    float* mean_rstd, float* scale_shift, int relu, void* workspace, cudaStream_t st) {
  if (check_shape(M, C)) return -1;
  uint8_t* gate_out = reinterpret_cast<uint8_t*>(gate);
  float* partial = reinterpret_cast<float*>(reinterpret_cast<char*>(workspace) + 256);
  unsigned int* counter = reinterpret_cast<unsigned int*>(workspace);
  if (tma_path(M, C)) {
    if (int r = ensure_tma_attrs()) return r;
    CUtensorMap mx;
    if (int r = pdj_tma_map_2d_bf16(&mx, x, M, C, C, TC, TR, 0)) return r;
    dim3 g(tma_grid_rows(M, C, sm_count()), C / TC);
    bn_stats_tma_kernel<kStatsStages><<<g, kThreads, kStatsSmem, st>>>(mx, M, C, gamma, beta, eps, momentum, running_mean,
                                                                      running_var, mean_rstd, scale_shift, partial, counter);
  } else {
    const int c8 = C / 8, cvt = eff_cvt(C);
    dim3 g(grid_rows(M, C, sm_count()), c8 / cvt);
    launch_stats(g, st, (const __nv_bfloat16*)x, M, C, gamma, beta, eps, momentum, running_mean, running_var, mean_rstd,
                 scale_shift, partial, counter, cvt);
  }
  PDJ_CUDA_CHECK(cudaGetLastError());
  return launch_apply(x, res, y, gate_out, M, C, scale_shift, relu, st);
}

// Same as pdj_bn_fwd_gate, but the per-channel sums come from `partial` ([nb][2][C] fp32: sum, sum
// of squares; written by pdj_conv1x1_bf16) instead of a pass over x.
int pdj_bn_fwd_pre(const void* x, const void* res, void* y, void* gate, int M, int C, const float* gamma,
                   const float* beta, float eps, float momentum, float* running_mean, float* running_var,
                   float* mean_rstd, float* scale_shift, int relu, const float* partial, int nb, cudaStream_t st) {
  if (check_shape(M, C)) return -1;
  if (partial == nullptr || nb < 1) {
    pdj_set_error("pdj_bn_fwd_pre: no partial sums (nb=%d)", nb);
    return -1;
  }
  bn_finalize_kernel<<<C / 8, kThreads, 0, st>>>(partial, (unsigned int)nb, M, C, gamma, beta, eps, momentum, running_mean,
                                                running_var, mean_rstd, scale_shift);
  PDJ_CUDA_CHECK(cudaGetLastError());
  return launch_apply(x, res, y, reinterpret_cast<uint8_t*>(gate), M, C, scale_shift, relu, st);
}

int pdj_bn_bwd_gate(const void* dy, const void* y, const void* gate, const void* x, void* dx, void* dres, int M, int C,
                    const float* gamma, const float* mean_rstd, float* dgamma_dbeta, void* dgb_bf16, float* coef,
                    int relu, void* workspace, cudaStream_t st);

int pdj_bn_bwd(const void* dy, const void* y, const void* x, void* dx, void* dres, int M, int C, const float* gamma,
               const float* mean_rstd, float* dgamma_dbeta, void* dgb_bf16, float* coef, int relu,
               void* workspace, cudaStream_t st) {
  return pdj_bn_bwd_gate(dy, y, nullptr, x, dx, dres, M, C, gamma, mean_rstd, dgamma_dbeta, dgb_bf16, coef, relu,
                         workspace, st);
}

// `gate` (from pdj_bn_fwd_gate) replaces y as the source of the ReLU derivative when given.
int pdj_bn_bwd_gate(const void* dy, const void* y, const void* gate, const void* x, void* dx, void* dres, int M, int C,
                    const float* gamma, const float* mean_rstd, float* dgamma_dbeta, void* dgb_bf16, float* coef,
                    int relu, void* workspace, cudaStream_t st) {
  return pdj_bn_bwd_gate2(dy, y, gate, x, dx, dres, M, C, gamma, mean_rstd, dgamma_dbeta, dgb_bf16,
                          dgb_bf16 ? (void*)((__nv_bfloat16*)dgb_bf16 + C) : nullptr, coef, relu, workspace, st);
}

// The same with the two bf16 gradients at separate addresses (a trainer's flat gradient buffer).
int pdj_bn_bwd_gate2(const void* dy, const void* y, const void* gate, const void* x, void* dx, void* dres, int M, int C,
                     const float* gamma, const float* mean_rstd, float* dgamma_dbeta, void* dgb_bf16, void* dbeta_bf16,
                     float* coef, int relu, void* workspace, cudaStream_t st) {
  if (check_shape(M, C)) return -1;
  const uint8_t* gate_in = reinterpret_cast<const uint8_t*>(gate);
  float* partial = reinterpret_cast<float*>(reinterpret_cast<char*>(workspace) + 256);
  unsigned int* counter = reinterpret_cast<unsigned int*>(workspace);
  const int c8 = C / 8, cvt = eff_cvt(C);
  dim3 g(grid_rows(M, C, sm_count()), c8 / cvt);
  const __nv_bfloat16* dyb = (const __nv_bfloat16*)dy;
  const __nv_bfloat16* yb = (const __nv_bfloat16*)y;
  const __nv_bfloat16* xb = (const __nv_bfloat16*)x;
  __nv_bfloat16* dgb = (__nv_bfloat16*)dgb_bf16;
  __nv_bfloat16* dbb = (__nv_bfloat16*)dbeta_bf16;
  const int cm = cvt;
  if (tma_path(M, C) && (!relu || gate_in != nullptr)) {
    if (int r = ensure_tma_attrs()) return r;
    CUtensorMap md, mx;
    if (int r = pdj_tma_map_2d_bf16(&md, dy, M, C, C, TC, TR, 0)) return r;
    if (int r = pdj_tma_map_2d_bf16(&mx, x, M, C, C, TC, TR, 0)) return r;
    dim3 gt(tma_grid_rows(M, C, sm_count()), C / TC);
    if (relu)
      bn_bwd_reduce_tma_kernel<true, kBwdStages><<<gt, kThreads, kBwdSmem, st>>>(md, mx, M, C, gamma, mean_rstd, dgamma_dbeta,
                                                                                dgb, dbb, coef, partial, counter, gate_in);
    else
      bn_bwd_reduce_tma_kernel<false, kBwdStages><<<gt, kThreads, kBwdSmem, st>>>(md, mx, M, C, gamma, mean_rstd, dgamma_dbeta,
                                                                                 dgb, dbb, coef, partial, counter, gate_in);
  } else if (knob_bwd_unroll() == 2) {
    if (relu)
      bn_bwd_reduce_kernel<true, 2><<<g, kThreads, 0, st>>>(dyb, yb, xb, M, C, gamma, mean_rstd, dgamma_dbeta, dgb, dbb, coef,
                                                            partial, counter, cm, gate_in);
    else
      bn_bwd_reduce_kernel<false, 2><<<g, kThreads, 0, st>>>(dyb, yb, xb, M, C, gamma, mean_rstd, dgamma_dbeta, dgb, dbb,
                                                             coef, partial, counter, cm, gate_in);
  } else {
    if (relu)
      bn_bwd_reduce_kernel<true, 1><<<g, kThreads, 0, st>>>(dyb, yb, xb, M, C, gamma, mean_rstd, dgamma_dbeta, dgb, dbb, coef,
                                                            partial, counter, cm, gate_in);
    else
      bn_bwd_reduce_kernel<false, 1><<<g, kThreads, 0, st>>>(dyb, yb, xb, M, C, gamma, mean_rstd, dgamma_dbeta, dgb, dbb,
                                                             coef, partial, counter, cm, gate_in);
  }
  PDJ_CUDA_CHECK(cudaGetLastError());
  const size_t n_vec = (size_t)M * c8;
  const int fg = flat_grid(n_vec);
  __nv_bfloat16* dxb = (__nv_bfloat16*)dx;
  __nv_bfloat16* drb = (__nv_bfloat16*)dres;
  if (relu && dres) bn_bwd_apply_kernel<true, true><<<fg, kThreads, 0, st>>>(dyb, yb, xb, dxb, drb, n_vec, C, coef, gate_in);
  else if (relu) bn_bwd_apply_kernel<true, false><<<fg, kThreads, 0, st>>>(dyb, yb, xb, dxb, drb, n_vec, C, coef, gate_in);
  else if (dres) bn_bwd_apply_kernel<false, true><<<fg, kThreads, 0, st>>>(dyb, yb, xb, dxb, drb, n_vec, C, coef, gate_in);
  else bn_bwd_apply_kernel<false, false><<<fg, kThreads, 0, st>>>(dyb, yb, xb, dxb, drb, n_vec, C, coef, gate_in);
  PDJ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // extern "C"

// ---------------------------------------------------------------------------
// pdj_bn_fwd_noalloc: the forward BatchNorm(+add)(+ReLU) for callers that must
// not allocate device memory.
//
// Who that is: (1) plain C / FFI callers that drive this library without
// PyTorch and therefore have no allocator to get a workspace from; (2) rank
// loops that run after the job's HBM budget has been frozen (arena + framework
// pool sized at start-up), where a late cudaMalloc can fail outright and, worse,
// cudaMalloc/cudaFree synchronise the whole device - which stalls this rank while
// its peers spin inside the fused collectives waiting for it.
//
// The block partial sums and the "last block" tickets that pdj_bn_fwd keeps in a
// caller-provided workspace live here in one file-scope device array instead.
// Every GPU that loads this library carries those PDJ_BN_NOALLOC_SCRATCH_BYTES
// (2 MiB + 256 B) until the process exits.
//
// Limits: C a power-of-two multiple of 8 up to PDJ_BN_NOALLOC_MAX_C (2048) and
// 1 <= M <= PDJ_BN_NOALLOC_MAX_ROWS (2^24 rows); anything else is rejected with
// an error message and nothing is launched.
//
// Set-up: the first call on each device resolves the array's device address
// (cudaGetSymbolAddress, which may also load the module) and caches it. Make
// that first call during start-up, before any stream capture begins.
//
// Sharing: there is ONE array per device. All calls use it, and it keeps
// whatever the previous call left behind; the ticket counters are therefore
// cleared (cudaMemsetAsync on the caller's stream) at the start of every call,
// the partial sums are fully rewritten before they are read. Two calls that can
// overlap on the GPU (different streams, or graphs replayed concurrently) race
// on the array and produce garbage - serialise them.
//
// Calls no cudaMalloc / cudaFree / allocating library routine.
#include <mutex>

namespace {

constexpr size_t kNoAllocCounterBytes = 256;
constexpr size_t kNoAllocPartialBytes = PDJ_BN_NOALLOC_SCRATCH_BYTES - kNoAllocCounterBytes;

// In host code this name is only a handle for cudaGetSymbolAddress - never pass
// it to a kernel directly.
static __device__ __align__(256) unsigned char g_bn_noalloc_scratch[PDJ_BN_NOALLOC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
void* g_noalloc_ptr[kMaxDevices] = {nullptr};
int g_noalloc_sms[kMaxDevices] = {0};
std::mutex g_noalloc_mu;

// Once per device, with that device current.
int noalloc_setup(int dev, void** ptr, int* sms) {
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (g_noalloc_ptr[dev] == nullptr) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_bn_noalloc_scratch);
    if (e != cudaSuccess || p == nullptr) {
      pdj_set_error("pdj_bn_fwd_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
      return e != cudaSuccess ? (int)e : -1;
    }
    int n = 0;
    e = cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess || n <= 0) {
      pdj_set_error("pdj_bn_fwd_noalloc: cannot read the SM count of device %d: %s", dev, cudaGetErrorString(e));
      return e != cudaSuccess ? (int)e : -1;
    }
    g_noalloc_sms[dev] = n;
    g_noalloc_ptr[dev] = p;
  }
  *ptr = g_noalloc_ptr[dev];
  *sms = g_noalloc_sms[dev];
  return 0;
}

}  // namespace

extern "C" int pdj_bn_fwd_noalloc(const void* x, const void* res, void* y, int M, int C, const float* gamma,
                                  const float* beta, float eps, float momentum, float* running_mean,
                                  float* running_var, float* mean_rstd, float* scale_shift, int relu,
                                  cudaStream_t st) {
  const int c8 = C / 8;
  if (C % 8 != 0 || c8 < 1 || (c8 & (c8 - 1)) != 0 || C > PDJ_BN_NOALLOC_MAX_C) {
    pdj_set_error("pdj_bn_fwd_noalloc: C=%d is not supported (power-of-two multiple of 8, at most %d)", C,
                  PDJ_BN_NOALLOC_MAX_C);
    return -1;
  }
  if (M < 1 || M > PDJ_BN_NOALLOC_MAX_ROWS) {
    pdj_set_error("pdj_bn_fwd_noalloc: M=%d rows is outside 1..%d, the range the built-in scratch is sized for", M,
                  PDJ_BN_NOALLOC_MAX_ROWS);
    return -1;
  }
  int dev = 0;
  PDJ_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) {
    pdj_set_error("pdj_bn_fwd_noalloc: device index %d is out of range", dev);
    return -1;
  }
  void* scratch = nullptr;
  int sms = 0;
  if (int rc = noalloc_setup(dev, &scratch, &sms)) return rc;

  unsigned int* counter = reinterpret_cast<unsigned int*>(scratch);
  float* partial = reinterpret_cast<float*>(reinterpret_cast<char*>(scratch) + kNoAllocCounterBytes);
  uint8_t* gate_out = nullptr;
  // Tickets must start at zero whatever the previous user of the array did.
  PDJ_CUDA_CHECK(cudaMemsetAsync(counter, 0, kNoAllocCounterBytes, st));

  const int cvt = eff_cvt(C);  // at most 64 channel tiles: the ticket region is 64 counters
  int gx = grid_rows(M, C, sms);
  const int cap = (int)(kNoAllocPartialBytes / ((size_t)2 * C * sizeof(float)));  // row blocks that fit
  if (gx > cap) gx = cap;
  dim3 g(gx, c8 / cvt);
  launch_stats(g, st, (const __nv_bfloat16*)x, M, C, gamma, beta, eps, momentum, running_mean, running_var, mean_rstd,
               scale_shift, partial, counter, cvt);
  PDJ_CUDA_CHECK(cudaGetLastError());
  const size_t n_vec = (size_t)M * c8;
  size_t fg = (n_vec + (size_t)kThreads * 4 - 1) / ((size_t)kThreads * 4);
  if (fg > (size_t)sms * 8) fg = (size_t)sms * 8;
  if (fg < 1) fg = 1;
  const __nv_bfloat16* xb = (const __nv_bfloat16*)x;
  const __nv_bfloat16* rb = (const __nv_bfloat16*)res;
  __nv_bfloat16* yb = (__nv_bfloat16*)y;
  if (relu && res) bn_apply_kernel<true, true><<<(int)fg, kThreads, 0, st>>>(xb, rb, yb, n_vec, C, scale_shift, gate_out);
  else if (relu) bn_apply_kernel<true, false><<<(int)fg, kThreads, 0, st>>>(xb, rb, yb, n_vec, C, scale_shift, gate_out);
  else if (res) bn_apply_kernel<false, true><<<(int)fg, kThreads, 0, st>>>(xb, rb, yb, n_vec, C, scale_shift, gate_out);
  else bn_apply_kernel<false, false><<<(int)fg, kThreads, 0, st>>>(xb, rb, yb, n_vec, C, scale_shift, gate_out);
  PDJ_CUDA_CHECK(cudaGetLastError());
  return 0;
}
