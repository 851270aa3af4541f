// Max pooling for NHWC bf16 with a saved arg-max, forward and backward.
//
// The stock NHWC max-pool backward is an atomics-based scatter that ran at
// ~0.5 TB/s in the ResNet-50 step profile (profiles/); this pair streams at HBM
// speed: forward stores the window position of each maximum (one byte per
// element), backward is a gather over the <= ceil(k/s)^2 windows that cover an
// input pixel, so it is deterministic and atomic-free.
//
// Reference parity: none (the reference has no model code); part of the
// ResNet workload of deploy/examples/resnet.yaml:13-19.
#include "common.cuh"
#include "../../include/pdjob_b200.h"

namespace {

__device__ __forceinline__ void unpack8p(const uint4& u, float* f) {
  float2 a = pdj_bf2_to_f2(u.x), b = pdj_bf2_to_f2(u.y), c = pdj_bf2_to_f2(u.z), d = pdj_bf2_to_f2(u.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}

__global__ void __launch_bounds__(256)
maxpool_fwd_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, uint8_t* __restrict__ idx,
                   int N, int H, int W, int C, int Ho, int Wo, int k, int s, int p) {
  const int c8 = C >> 3;
  // 32-bit index arithmetic (the host rejects tensors with >= 2^32 vectors):
  // 64-bit div/mod per element dominated the first version of this kernel.
  const uint32_t total = (uint32_t)N * Ho * Wo * c8;
  const uint32_t stride = gridDim.x * blockDim.x;
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
    const int cv = (int)(i % (uint32_t)c8);
    uint32_t t = i / (uint32_t)c8;
    const int wo = (int)(t % (uint32_t)Wo); t /= (uint32_t)Wo;
    const int ho = (int)(t % (uint32_t)Ho);
    const int n = (int)(t / (uint32_t)Ho);
    float best[8];
    uint32_t arg[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { best[j] = -INFINITY; arg[j] = 0; }
    for (int kh = 0; kh < k; ++kh) {
      const int h = ho * s - p + kh;
      if (h < 0 || h >= H) continue;
      for (int kw = 0; kw < k; ++kw) {
        const int w = wo * s - p + kw;
        if (w < 0 || w >= W) continue;
        const uint4 u = __ldg(reinterpret_cast<const uint4*>(x + (((size_t)n * H + h) * W + w) * C) + cv);
        float f[8];
        unpack8p(u, f);
        const uint32_t code = (uint32_t)(kh * k + kw);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          if (f[j] > best[j] || f[j] != f[j]) { best[j] = f[j]; arg[j] = code; }  // first max wins; NaN propagates
        }
      }
    }
    uint4 o;
    o.x = pdj_f2_to_bf2(best[0], best[1]); o.y = pdj_f2_to_bf2(best[2], best[3]);
    o.z = pdj_f2_to_bf2(best[4], best[5]); o.w = pdj_f2_to_bf2(best[6], best[7]);
    reinterpret_cast<uint4*>(y)[i] = o;
    uint2 a;
    a.x = arg[0] | (arg[1] << 8) | (arg[2] << 16) | (arg[3] << 24);
    a.y = arg[4] | (arg[5] << 8) | (arg[6] << 16) | (arg[7] << 24);
    reinterpret_cast<uint2*>(idx)[i] = a;
  }
}

// One block per input row (n, h): the window rows that cover h are the same for the whole block, and the
// only per-element index arithmetic left is w / cv from the thread index.  K, S, P > 0 are compile-time for
// the ResNet stem pool (3, 2, 1), which turns the window bounds into shifts; K = 0 is the generic version.
template <int K, int S, int P>
__global__ void __launch_bounds__(256)
maxpool_bwd_kernel(const __nv_bfloat16* __restrict__ dy, const uint8_t* __restrict__ idx,
                   __nv_bfloat16* __restrict__ dx, int H, int W, int C, int Ho, int Wo, int k_rt, int s_rt, int p_rt) {
  const int k = K > 0 ? K : k_rt, s = K > 0 ? S : s_rt, p = K > 0 ? P : p_rt;
  const int c8 = C >> 3;
  const int h = (int)(blockIdx.x % (uint32_t)H);
  const int n = (int)(blockIdx.x / (uint32_t)H);
  // windows ho with ho*s - p <= h <= ho*s - p + k - 1
  int ho_lo = h + p - k + 1;
  ho_lo = ho_lo <= 0 ? 0 : (ho_lo + s - 1) / s;
  int ho_hi = (h + p) / s;
  if (ho_hi > Ho - 1) ho_hi = Ho - 1;
  const int row_items = W * c8;
  const uint4* dyv = reinterpret_cast<const uint4*>(dy);
  const uint2* idv = reinterpret_cast<const uint2*>(idx);
  uint4* dxv = reinterpret_cast<uint4*>(dx) + ((size_t)n * H + h) * row_items;
  for (int it = threadIdx.x; it < row_items; it += 256) {
    const int w = it / c8, cv = it - w * c8;
    int wo_lo = w + p - k + 1;
    wo_lo = wo_lo <= 0 ? 0 : (wo_lo + s - 1) / s;
    int wo_hi = (w + p) / s;
    if (wo_hi > Wo - 1) wo_hi = Wo - 1;
    float g[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) g[j] = 0.f;
    for (int ho = ho_lo; ho <= ho_hi; ++ho) {
      const int kh = h - (ho * s - p);
      const size_t orow = ((size_t)n * Ho + ho) * Wo;
      for (int wo = wo_lo; wo <= wo_hi; ++wo) {
        const uint32_t code = (uint32_t)(kh * k + (w - (wo * s - p)));
        const size_t o = (orow + wo) * c8 + cv;
        const uint2 a = __ldg(idv + o);
        const uint4 u = __ldg(dyv + o);
        float f[8];
        unpack8p(u, f);
        // byte j of the arg-max word equals `code` <=> this input element won window (ho, wo)
        const uint32_t pat = code * 0x01010101u;
        const uint32_t x0 = a.x ^ pat, x1 = a.y ^ pat;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (((x0 >> (8 * j)) & 0xffu) == 0u) g[j] += f[j];
          if (((x1 >> (8 * j)) & 0xffu) == 0u) g[4 + j] += f[4 + j];
        }
      }
    }
    uint4 out;
    out.x = pdj_f2_to_bf2(g[0], g[1]); out.y = pdj_f2_to_bf2(g[2], g[3]);
    out.z = pdj_f2_to_bf2(g[4], g[5]); out.w = pdj_f2_to_bf2(g[6], g[7]);
    dxv[it] = out;
  }
}

int pool_grid(size_t total) {
  size_t g = (total + 255) / 256;
  if (g > 148 * 16) g = 148 * 16;
  if (g < 1) g = 1;
  return (int)g;
}

}  // namespace

extern "C" {

int pdj_maxpool_fwd(const void* x, void* y, void* idx, int N, int H, int W, int C, int k, int s, int p,
                    cudaStream_t st) {
  if (C % 8 != 0 || k < 1 || k > 15 || s < 1 || p < 0 || 2 * p > k) {
    pdj_set_error("pdj_maxpool_fwd: need C %% 8 == 0, 1 <= k <= 15, s >= 1, 0 <= 2p <= k");
    return -1;
  }
  const int Ho = (H + 2 * p - k) / s + 1, Wo = (W + 2 * p - k) / s + 1;
  if ((size_t)N * H * W * (C / 8) >= 0xffffffffull) {
    pdj_set_error("pdj_maxpool_fwd: tensor too large for 32-bit indexing");
    return -1;
  }
  maxpool_fwd_kernel<<<pool_grid((size_t)N * Ho * Wo * (C / 8)), 256, 0, st>>>(
      (const __nv_bfloat16*)x, (__nv_bfloat16*)y, (uint8_t*)idx, N, H, W, C, Ho, Wo, k, s, p);
  PDJ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int pdj_maxpool_bwd(const void* dy, const void* idx, void* dx, int N, int H, int W, int C, int k, int s, int p,
                    cudaStream_t st) {
  if (C % 8 != 0 || k < 1 || k > 15 || s < 1 || p < 0 || 2 * p > k) {
    pdj_set_error("pdj_maxpool_bwd: need C %% 8 == 0, 1 <= k <= 15, s >= 1, 0 <= 2p <= k");
    return -1;
  }
  const int Ho = (H + 2 * p - k) / s + 1, Wo = (W + 2 * p - k) / s + 1;
  if ((size_t)N * H * W * (C / 8) >= 0xffffffffull) {
    pdj_set_error("pdj_maxpool_bwd: tensor too large for 32-bit indexing");
    return -1;
  }
  const unsigned int rows = (unsigned int)N * (unsigned int)H;
  if (k == 3 && s == 2 && p == 1)
    maxpool_bwd_kernel<3, 2, 1><<<rows, 256, 0, st>>>((const __nv_bfloat16*)dy, (const uint8_t*)idx, (__nv_bfloat16*)dx, H, W,
                                                      C, Ho, Wo, k, s, p);
  else
    maxpool_bwd_kernel<0, 0, 0><<<rows, 256, 0, st>>>((const __nv_bfloat16*)dy, (const uint8_t*)idx, (__nv_bfloat16*)dx, H, W,
                                                      C, Ho, Wo, k, s, p);
  PDJ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // extern "C"
