// Flat-buffer optimiser, input pre-processing and the classification loss.
//
// Reference parity: none in the reference (it has no training code); these are
// the per-step kernels of the Collective-mode ResNet-50 workload that
// deploy/examples/resnet.yaml launches, owned here by the runtime.
#include "common.cuh"
#include "../../include/pdjob_b200.h"

namespace {

// hyper = {lr, momentum, weight_decay, grad_scale}; read from device memory so
// the launch can sit inside a CUDA graph while the schedule changes.
__global__ void __launch_bounds__(256)
sgd_momentum_kernel(float* __restrict__ w, float* __restrict__ mom, const __nv_bfloat16* __restrict__ grad,
                    __nv_bfloat16* __restrict__ param, size_t n_vec, size_t wd_split_vec,
                    const float* __restrict__ hyper) {
  const float lr = hyper[0], mu = hyper[1], wd = hyper[2], gs = hyper[3];
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_vec; i += stride) {
    const uint4 g4 = __ldg(reinterpret_cast<const uint4*>(grad) + i);
    float4 w0 = reinterpret_cast<float4*>(w)[2 * i], w1 = reinterpret_cast<float4*>(w)[2 * i + 1];
    float4 m0 = reinterpret_cast<float4*>(mom)[2 * i], m1 = reinterpret_cast<float4*>(mom)[2 * i + 1];
    float g[8];
    {
      float2 a = pdj_bf2_to_f2(g4.x), b = pdj_bf2_to_f2(g4.y), c = pdj_bf2_to_f2(g4.z), d = pdj_bf2_to_f2(g4.w);
      g[0] = a.x; g[1] = a.y; g[2] = b.x; g[3] = b.y; g[4] = c.x; g[5] = c.y; g[6] = d.x; g[7] = d.y;
    }
    float ww[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
    float mm[8] = {m0.x, m0.y, m0.z, m0.w, m1.x, m1.y, m1.z, m1.w};
    const float dec = i < wd_split_vec ? wd : 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float gj = fmaf(dec, ww[j], g[j] * gs);
      mm[j] = fmaf(mu, mm[j], gj);
      ww[j] = fmaf(-lr, mm[j], ww[j]);
    }
    reinterpret_cast<float4*>(w)[2 * i] = make_float4(ww[0], ww[1], ww[2], ww[3]);
    reinterpret_cast<float4*>(w)[2 * i + 1] = make_float4(ww[4], ww[5], ww[6], ww[7]);
    reinterpret_cast<float4*>(mom)[2 * i] = make_float4(mm[0], mm[1], mm[2], mm[3]);
    reinterpret_cast<float4*>(mom)[2 * i + 1] = make_float4(mm[4], mm[5], mm[6], mm[7]);
    uint4 o;
    o.x = pdj_f2_to_bf2(ww[0], ww[1]); o.y = pdj_f2_to_bf2(ww[2], ww[3]);
    o.z = pdj_f2_to_bf2(ww[4], ww[5]); o.w = pdj_f2_to_bf2(ww[6], ww[7]);
    reinterpret_cast<uint4*>(param)[i] = o;
  }
}

// uint8 NHWC (C = 3) -> bf16 NHWC, (x/255 - mean) / std per channel.
__global__ void __launch_bounds__(256)
preprocess_kernel(const uint8_t* __restrict__ in, __nv_bfloat16* __restrict__ out, size_t n_px, float m0, float m1,
                  float m2, float s0, float s1, float s2) {
  // One thread per 4 pixels = 12 input bytes (3 x u32) -> 12 bf16 (24 bytes).
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  const size_t n4 = n_px / 4;
  const float mean[3] = {m0, m1, m2}, inv[3] = {1.f / (255.f * s0), 1.f / (255.f * s1), 1.f / (255.f * s2)};
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    const uint32_t* p = reinterpret_cast<const uint32_t*>(in) + i * 3;
    const uint32_t u[3] = {__ldg(p), __ldg(p + 1), __ldg(p + 2)};
    float f[12];
#pragma unroll
    for (int k = 0; k < 12; ++k) {
      const float v = (float)((u[k >> 2] >> ((k & 3) * 8)) & 0xffu);
      const int ch = k % 3;
      f[k] = v * inv[ch] - mean[ch] / (ch == 0 ? s0 : (ch == 1 ? s1 : s2));
    }
    uint32_t* o = reinterpret_cast<uint32_t*>(out) + i * 6;
#pragma unroll
    for (int k = 0; k < 6; ++k) o[k] = pdj_f2_to_bf2(f[2 * k], f[2 * k + 1]);
  }
  // tail pixels (n_px % 4), handled by the first threads
  const size_t tail0 = n4 * 4;
  for (size_t px = tail0 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; px < n_px; px += stride) {
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
      const float v = (float)in[px * 3 + ch];
      const float s = ch == 0 ? s0 : (ch == 1 ? s1 : s2);
      out[px * 3 + ch] = __float2bfloat16(v / (255.f * s) - mean[ch] / s);
    }
  }
}

// uint8 NHWC (C = 3) -> normalised bf16 in the space-to-depth layout of the ResNet stem:
// out[n, by + 2, bx + 2, (dy*2 + dx)*3 + c] = norm(in[n, 2*by + dy, 2*bx + dx, c]), 16 channels per
// pixel (12 used, 4 zero), spatial size (H/2 + 3, W/2 + 3) whose border (2 before, 1 after) stays
// zero.  A 7x7 / stride-2 / pad-3 convolution over the image equals a 4x4 / stride-1 / pad-0
// convolution over this tensor (weights re-arranged, see models/resnet.py), which runs on the
// tensor cores at K = 256 instead of K = 147 with 3-channel gathers.
__global__ void __launch_bounds__(256)
preprocess_s2d_kernel(const uint8_t* __restrict__ in, __nv_bfloat16* __restrict__ out, int N, int H, int W, float m0,
                      float m1, float m2, float s0, float s1, float s2) {
  const int Hb = H >> 1, Wb = W >> 1, Ho = Hb + 3, Wo = Wb + 3;
  const uint32_t total = (uint32_t)N * Hb * Wb;
  const float inv[3] = {1.f / (255.f * s0), 1.f / (255.f * s1), 1.f / (255.f * s2)};
  const float off[3] = {m0 / s0, m1 / s1, m2 / s2};
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int bx = (int)(i % (uint32_t)Wb);
    uint32_t t = i / (uint32_t)Wb;
    const int by = (int)(t % (uint32_t)Hb);
    const int n = (int)(t / (uint32_t)Hb);
    float f[16];
#pragma unroll
    for (int dy = 0; dy < 2; ++dy) {
      const uint16_t* p = reinterpret_cast<const uint16_t*>(in + (((size_t)n * H + 2 * by + dy) * W + 2 * bx) * 3);
      const uint32_t a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);  // 6 bytes: two RGB pixels
      const uint32_t px[6] = {a & 0xffu, a >> 8, b & 0xffu, b >> 8, c & 0xffu, c >> 8};
#pragma unroll
      for (int k = 0; k < 6; ++k) f[dy * 6 + k] = (float)px[k] * inv[k % 3] - off[k % 3];
    }
    f[12] = f[13] = f[14] = f[15] = 0.f;
    uint4 o0, o1;
    o0.x = pdj_f2_to_bf2(f[0], f[1]); o0.y = pdj_f2_to_bf2(f[2], f[3]);
    o0.z = pdj_f2_to_bf2(f[4], f[5]); o0.w = pdj_f2_to_bf2(f[6], f[7]);
    o1.x = pdj_f2_to_bf2(f[8], f[9]); o1.y = pdj_f2_to_bf2(f[10], f[11]);
    o1.z = 0u; o1.w = 0u;
    uint4* dst = reinterpret_cast<uint4*>(out + (((size_t)n * Ho + by + 2) * Wo + bx + 2) * 16);
    dst[0] = o0;
    dst[1] = o1;
  }
}

// Softmax cross-entropy, forward and backward in one pass over the logits.
// One block per row; writes dlogits = (softmax - onehot) * grad_scale (bf16),
// and atomically accumulates loss_sum += -log p[label] (caller zeroes it).
__global__ void __launch_bounds__(256)
softmax_xent_kernel(const __nv_bfloat16* __restrict__ logits, const int64_t* __restrict__ labels,
                    __nv_bfloat16* __restrict__ dlogits, float* __restrict__ loss_sum, int ncls, float grad_scale) {
  __shared__ float red[32];
  const int row = blockIdx.x;
  const __nv_bfloat16* lr = logits + (size_t)row * ncls;
  float mx = -INFINITY;
  for (int j = threadIdx.x; j < ncls; j += blockDim.x) mx = fmaxf(mx, __bfloat162float(lr[j]));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = mx;
  __syncthreads();
  mx = red[0];
  for (int w = 1; w < (int)(blockDim.x >> 5); ++w) mx = fmaxf(mx, red[w]);
  __syncthreads();
  float sum = 0.f;
  for (int j = threadIdx.x; j < ncls; j += blockDim.x) sum += __expf(__bfloat162float(lr[j]) - mx);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = sum;
  __syncthreads();
  sum = 0.f;
  for (int w = 0; w < (int)(blockDim.x >> 5); ++w) sum += red[w];
  const float inv = 1.f / sum;
  const int lab = (int)labels[row];
  __nv_bfloat16* dr = dlogits + (size_t)row * ncls;
  for (int j = threadIdx.x; j < ncls; j += blockDim.x) {
    const float p = __expf(__bfloat162float(lr[j]) - mx) * inv;
    dr[j] = __float2bfloat16((p - (j == lab ? 1.f : 0.f)) * grad_scale);
  }
  if (threadIdx.x == 0) {
    const float lp = __bfloat162float(lr[lab]) - mx - __logf(sum);
    atomicAdd(loss_sum, -lp);
  }
}

// Column sums of a row-major bf16 matrix: out[c] += sum_r x[r, c] (fp32, caller zeroes out).
// The bias gradient of a linear layer.  One block takes ROWS rows; a thread owns two columns.
constexpr int kColsumRows = 128;
__global__ void __launch_bounds__(256)
colsum_kernel(const __nv_bfloat16* __restrict__ x, float* __restrict__ out, int M, int N, int ld) {
  const int r0 = blockIdx.x * kColsumRows;
  const int r1 = r0 + kColsumRows < M ? r0 + kColsumRows : M;
  for (int c = 2 * threadIdx.x; c < N; c += 2 * blockDim.x) {
    float a = 0.f, b = 0.f;
    if (c + 1 < N && (ld & 1) == 0) {
      for (int r = r0; r < r1; ++r) {
        const uint32_t u = __ldg(reinterpret_cast<const uint32_t*>(x + (size_t)r * ld + c));
        const float2 f = pdj_bf2_to_f2(u);
        a += f.x;
        b += f.y;
      }
      atomicAdd(out + c, a);
      atomicAdd(out + c + 1, b);
    } else {
      for (int r = r0; r < r1; ++r) {
        a += __bfloat162float(x[(size_t)r * ld + c]);
        if (c + 1 < N) b += __bfloat162float(x[(size_t)r * ld + c + 1]);
      }
      atomicAdd(out + c, a);
      if (c + 1 < N) atomicAdd(out + c + 1, b);
    }
  }
}

// ---- Weighted column sums, deterministic, one launch ---------------------------------------------------
//   res[c]          = sum_r w[r] * x[r, c]        c < N        (x bf16, N % 8 == 0)
//   res[N + j]      = sum_r w[r] * xe[r, j]       j < Ne       (xe fp32, optional)
//   res[N + Ne]     = sum_r w[r]                               (optional "ones" column)
// with w[r] = wgt[r] * gscale[0] (either may be absent = 1).  Bias gradients of linear layers (w = 1) and every
// parameter gradient of the Wide&Deep head (x = top activation, xe = dense features, w = dlogit).
// Block b reduces a fixed row range: a thread owns 8 adjacent columns of every RG-th row, the row groups are
// folded in shared memory in a fixed order, the block writes one row of `partial`; the block that draws the last
// ticket adds the partial rows in block order, converts and scatters the result to up to four destinations
// (a trainer's flat gradient buffer) and resets the ticket.  Same bits every run, no memset, no conversion pass.
constexpr int kCsThreads = 1024;
constexpr int kCsMaxSlots = 128;  // 8-column slots: N / 8 + ceil((Ne + ones) / 8)
struct CsSeg { void* p; int start, n, bf16; };
struct CsOut { CsSeg seg[4]; };

__global__ void __launch_bounds__(kCsThreads)
colsum_w_kernel(const __nv_bfloat16* __restrict__ x, int ld, int M, int N, const float* __restrict__ wgt,
                const float* __restrict__ gscale, const float* __restrict__ xe, int lde, int Ne, int ones,
                float* __restrict__ partial, unsigned int* __restrict__ counter, CsOut out) {
  extern __shared__ float cs_smem[];  // [RG][slots * 8]
  const int nv = N >> 3, ne = Ne + ones, slots = nv + ((ne + 7) >> 3), ntot = N + ne;
  const int RG = kCsThreads / slots;
  const int slot = threadIdx.x % slots, rg = threadIdx.x / slots;
  const int rows_per = (M + gridDim.x - 1) / gridDim.x;
  const int r0 = blockIdx.x * rows_per, r1 = min(M, r0 + rows_per);
  const float gs = gscale ? __ldg(gscale) : 1.f;
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  if (rg < RG) {
    if (slot < nv) {
      const __nv_bfloat16* xp = x + (size_t)slot * 8;
      auto fold = [&](const uint4 v, const float w) {
        const float2 a = pdj_bf2_to_f2(v.x), b = pdj_bf2_to_f2(v.y), c = pdj_bf2_to_f2(v.z), d = pdj_bf2_to_f2(v.w);
        acc[0] = fmaf(w, a.x, acc[0]); acc[1] = fmaf(w, a.y, acc[1]);
        acc[2] = fmaf(w, b.x, acc[2]); acc[3] = fmaf(w, b.y, acc[3]);
        acc[4] = fmaf(w, c.x, acc[4]); acc[5] = fmaf(w, c.y, acc[5]);
        acc[6] = fmaf(w, d.x, acc[6]); acc[7] = fmaf(w, d.y, acc[7]);
      };
      // eight rows per round, the last round predicated: all loads of a round are issued before the first use (one
      // block per SM, so the memory-level parallelism has to come from inside the thread) and no serial tail remains
      for (int r = r0 + rg; r < r1; r += 8 * RG) {
        uint4 v[8];
        float w[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int ru = r + u * RG;
          const bool in = ru < r1;
          v[u] = in ? __ldg(reinterpret_cast<const uint4*>(xp + (size_t)ru * ld)) : make_uint4(0u, 0u, 0u, 0u);
          w[u] = in ? (wgt ? __ldg(wgt + ru) : 1.f) : 0.f;
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) fold(v[u], w[u] * gs);
      }
    } else {
      const int e0 = (slot - nv) * 8;
      for (int r = r0 + rg; r < r1; r += RG) {
        const float w = wgt ? __ldg(wgt + r) * gs : gs;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int e = e0 + j;
          if (e < Ne) acc[j] = fmaf(w, __ldg(xe + (size_t)r * lde + e), acc[j]);
          else if (e < ne) acc[j] += w;
        }
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) cs_smem[(rg * slots + slot) * 8 + j] = acc[j];
  }
  __syncthreads();
  const int pitch = (ntot + 3) & ~3;  // partial rows are whole float4s
  float* prow = partial + (size_t)blockIdx.x * pitch;
  for (int c = threadIdx.x; c < pitch; c += kCsThreads) {
    float t = 0.f;
    if (c < ntot)
      for (int g = 0; g < RG; ++g) t += cs_smem[g * slots * 8 + c];
    prow[c] = t;
  }
  __shared__ unsigned int ticket;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) ticket = atomicAdd(counter, 1u);
  __syncthreads();
  if (ticket != gridDim.x - 1) return;
  __threadfence();
  // last block: P2 row groups x float4 columns in parallel (independent loads), then a fixed-order fold
  const int nvec = pitch >> 2, P2 = kCsThreads / nvec;
  const int v = threadIdx.x % nvec, g2 = threadIdx.x / nvec;
  if (g2 < P2) {
    float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 8
    for (int b = g2; b < (int)gridDim.x; b += P2) {
      const float4 q = __ldcg(reinterpret_cast<const float4*>(partial + (size_t)b * pitch) + v);
      t.x += q.x; t.y += q.y; t.z += q.z; t.w += q.w;
    }
    reinterpret_cast<float4*>(cs_smem)[g2 * nvec + v] = t;
  }
  __syncthreads();
  for (int c = threadIdx.x; c < ntot; c += kCsThreads) {
    float t = 0.f;
    for (int g = 0; g < P2; ++g) t += cs_smem[g * pitch + c];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const CsSeg sg = out.seg[k];
      if (sg.p != nullptr && c >= sg.start && c < sg.start + sg.n) {
        if (sg.bf16) reinterpret_cast<__nv_bfloat16*>(sg.p)[c - sg.start] = __float2bfloat16(t);
        else reinterpret_cast<float*>(sg.p)[c - sg.start] = t;
      }
    }
  }
  if (threadIdx.x == 0) *counter = 0u;
}

// ---- Wide&Deep head: output layer + wide part + binary cross-entropy in one pass -------------------
// One warp per sample:
//   z = h[b,:] . w_out + b_out + wide_sum[b] + dense[b,:] . w_d + b_d
//   loss[b] = max(z,0) - z*y + log1p(exp(-|z|)),   dlogit[b] = (sigmoid(z) - y) * grad_scale
// replacing the gemv, bias adds, sigmoid / log-sigmoid / mean chain and their backward (about fifteen
// launches of a few microseconds each in a 1 ms step).
__global__ void __launch_bounds__(256)
wd_head_kernel(const __nv_bfloat16* __restrict__ h, int ld, const __nv_bfloat16* __restrict__ w_out,
               const __nv_bfloat16* __restrict__ b_out, const float* __restrict__ wide_sum,
               const float* __restrict__ dense, const __nv_bfloat16* __restrict__ w_d,
               const __nv_bfloat16* __restrict__ b_d, const float* __restrict__ labels, float* __restrict__ loss,
               float* __restrict__ dlogit, int B, int N, int Dd, float grad_scale) {
  const int lane = threadIdx.x & 31;
  const int warps = (gridDim.x * blockDim.x) >> 5;
  const float bias = (b_out ? __bfloat162float(b_out[0]) : 0.f) + (b_d ? __bfloat162float(b_d[0]) : 0.f);
  for (int b = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; b < B; b += warps) {
    const __nv_bfloat16* hr = h + (size_t)b * ld;
    float acc = 0.f;
    for (int c = 2 * lane; c + 1 < N; c += 64) {  // N is even (checked on the host)
      const float2 hv = pdj_bf2_to_f2(__ldg(reinterpret_cast<const uint32_t*>(hr + c)));
      const float2 wv = pdj_bf2_to_f2(__ldg(reinterpret_cast<const uint32_t*>(w_out + c)));
      acc = fmaf(hv.x, wv.x, acc);
      acc = fmaf(hv.y, wv.y, acc);
    }
    if (lane < Dd) acc = fmaf(dense[(size_t)b * Dd + lane], __bfloat162float(w_d[lane]), acc);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) {
      const float z = acc + bias + (wide_sum ? wide_sum[b] : 0.f);
      const float y = labels[b];
      const float e = __expf(-fabsf(z));
      loss[b] = fmaxf(z, 0.f) - z * y + log1pf(e);
      const float sig = z >= 0.f ? 1.f / (1.f + e) : e / (1.f + e);
      dlogit[b] = (sig - y) * grad_scale;
    }
  }
}

// dz[b,j] = (h[b,j] > 0) ? dlogit[b] * w_out[j] : 0  - the gradient entering the top ReLU layer, mask applied.
__global__ void __launch_bounds__(256)
wd_head_dz_kernel(const float* __restrict__ dlogit, const __nv_bfloat16* __restrict__ w_out,
                  const __nv_bfloat16* __restrict__ h, int ldh, __nv_bfloat16* __restrict__ dz, int ldz, int B, int N8) {
  const size_t total = (size_t)B * N8;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int b = (int)(i / N8), v = (int)(i % N8);
    const uint4 hv = __ldg(reinterpret_cast<const uint4*>(h + (size_t)b * ldh) + v);
    const uint4 wv = __ldg(reinterpret_cast<const uint4*>(w_out) + v);
    const float g = dlogit[b];
    const uint32_t hh[4] = {hv.x, hv.y, hv.z, hv.w}, ww[4] = {wv.x, wv.y, wv.z, wv.w};
    uint32_t o[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float2 hf = pdj_bf2_to_f2(hh[k]), wf = pdj_bf2_to_f2(ww[k]);
      o[k] = pdj_f2_to_bf2(hf.x > 0.f ? g * wf.x : 0.f, hf.y > 0.f ? g * wf.y : 0.f);
    }
    reinterpret_cast<uint4*>(dz + (size_t)b * ldz)[v] = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

int grid_for(size_t n, int per_block) {
  size_t g = (n + per_block - 1) / per_block;
  if (g > 148 * 8) g = 148 * 8;
  if (g < 1) g = 1;
  return (int)g;
}

}  // namespace

extern "C" {

int pdj_sgd_momentum(float* w, float* mom, const void* grad, void* param, size_t n, size_t wd_split,
                     const float* hyper, cudaStream_t st) {
  if (n % 8 != 0 || wd_split % 8 != 0) {
    pdj_set_error("pdj_sgd_momentum: n and wd_split must be multiples of 8");
    return -1;
  }
  sgd_momentum_kernel<<<grid_for(n / 8, 256 * 2), 256, 0, st>>>(w, mom, (const __nv_bfloat16*)grad,
                                                                (__nv_bfloat16*)param, n / 8, wd_split / 8, hyper);
  PDJ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int pdj_preprocess_u8(const void* in, void* out, size_t n_px, const float* mean3, const float* std3,
                      cudaStream_t st) {
  preprocess_kernel<<<grid_for(n_px / 4 + 1, 256 * 4), 256, 0, st>>>((const uint8_t*)in, (__nv_bfloat16*)out, n_px,
                                                                     mean3[0], mean3[1], mean3[2], std3[0], std3[1],
                                                                     std3[2]);
  PDJ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int pdj_preprocess_u8_s2d(const void* in, void* out, int N, int H, int W, const float* mean3, const float* std3,
                          cudaStream_t st) {
  if (H % 2 != 0 || W % 2 != 0 || N < 1) {
    pdj_set_error("pdj_preprocess_u8_s2d: H and W must be even");
    return -1;
  }
  const size_t total = (size_t)N * (H / 2) * (W / 2);
  if (total >= 0xffffffffull) {
    pdj_set_error("pdj_preprocess_u8_s2d: tensor too large for 32-bit indexing");
    return -1;
  }
  preprocess_s2d_kernel<<<grid_for(total, 256), 256, 0, st>>>((const uint8_t*)in, (__nv_bfloat16*)out, N, H, W, mean3[0],
                                                              mean3[1], mean3[2], std3[0], std3[1], std3[2]);
  PDJ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int pdj_colsum_bf16(const void* x, float* out, int M, int N, int ld, cudaStream_t st) {
  if (M < 1 || N < 1) return 0;
  PDJ_CUDA_CHECK(cudaMemsetAsync(out, 0, (size_t)N * sizeof(float), st));
  colsum_kernel<<<(M + kColsumRows - 1) / kColsumRows, 256, 0, st>>>((const __nv_bfloat16*)x, out, M, N, ld);
  PDJ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int pdj_colsum_w_ws_bytes(int M, int N, int Ne, int ones) {
  (void)M;
  return 256 + 148 * ((N + Ne + ones + 3) & ~3) * (int)sizeof(float);
}

// See colsum_w_kernel.  `ws`: pdj_colsum_w_ws_bytes bytes whose first 4 are zero before the first call (the kernel
// leaves them zero); dst[k] receives result columns [dst_start[k], dst_start[k] + dst_n[k]) as bf16 or fp32.
int pdj_colsum_w(const void* x, int ld, int M, int N, const float* wgt, const float* gscale, const float* xe, int lde,
                 int Ne, int ones, void* ws, void* const* dst, const int* dst_start, const int* dst_n,
                 const int* dst_bf16, int ndst, cudaStream_t st) {
  const int ne = Ne + (ones ? 1 : 0), slots = N / 8 + (ne + 7) / 8;
  if (M < 1 || N < 0 || N % 8 != 0 || ld % 8 != 0 || ((uintptr_t)x & 15) || slots < 1 || slots > kCsMaxSlots || ndst < 1 ||
      ndst > 4 || (Ne > 0 && xe == nullptr)) {
    pdj_set_error("pdj_colsum_w: need N %% 8 == 0, ld %% 8 == 0, 16-byte aligned x, at most %d column slots, 1..4 destinations "
                  "(N=%d ld=%d Ne=%d ndst=%d)", kCsMaxSlots, N, ld, Ne, ndst);
    return -1;
  }
  CsOut out{};
  for (int k = 0; k < ndst; ++k) out.seg[k] = CsSeg{dst[k], dst_start[k], dst_n[k], dst_bf16[k]};
  int grid = (M + 63) / 64;
  if (grid > 148) grid = 148;
  const int RG = kCsThreads / slots;
  const size_t smem = (size_t)RG * slots * 8 * sizeof(float);
  unsigned int* counter = reinterpret_cast<unsigned int*>(ws);
  float* partial = reinterpret_cast<float*>(reinterpret_cast<char*>(ws) + 256);
  colsum_w_kernel<<<grid, kCsThreads, smem, st>>>((const __nv_bfloat16*)x, ld, M, N, wgt, gscale, xe, lde, Ne, ones ? 1 : 0,
                                                  partial, counter, out);
  PDJ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int pdj_wd_head(const void* h, int ld, const void* w_out, const void* b_out, const float* wide_sum, const float* dense,
                const void* w_d, const void* b_d, const float* labels, float* loss, float* dlogit, int B, int N, int Dd,
                float grad_scale, cudaStream_t st) {
  if (B < 1 || N < 2 || N % 2 != 0 || ld % 2 != 0 || Dd < 0 || Dd > 32 || ((uintptr_t)h & 3) || ((uintptr_t)w_out & 3)) {
    pdj_set_error("pdj_wd_head: need even N and ld, at most 32 dense features, 4-byte aligned rows (N=%d ld=%d Dd=%d)", N, ld, Dd);
    return -1;
  }
  wd_head_kernel<<<grid_for((size_t)B * 32, 256), 256, 0, st>>>(
      (const __nv_bfloat16*)h, ld, (const __nv_bfloat16*)w_out, (const __nv_bfloat16*)b_out, wide_sum, dense,
      (const __nv_bfloat16*)w_d, (const __nv_bfloat16*)b_d, labels, loss, dlogit, B, N, Dd, grad_scale);
  PDJ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int pdj_wd_head_dz(const float* dlogit, const void* w_out, const void* h, int ldh, void* dz, int ldz, int B, int N,
                   cudaStream_t st) {
  if (B < 1 || N < 8 || N % 8 != 0 || ldh % 8 != 0 || ldz % 8 != 0 || ((uintptr_t)h & 15) || ((uintptr_t)dz & 15) ||
      ((uintptr_t)w_out & 15)) {
    pdj_set_error("pdj_wd_head_dz: need N, ldh, ldz multiples of 8 and 16-byte aligned tensors (N=%d)", N);
    return -1;
  }
  wd_head_dz_kernel<<<grid_for((size_t)B * (N / 8), 256), 256, 0, st>>>(dlogit, (const __nv_bfloat16*)w_out,
                                                                     (const __nv_bfloat16*)h, ldh, (__nv_bfloat16*)dz, ldz,
                                                                     B, N / 8);
  PDJ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

int pdj_softmax_xent(const void* logits, const int64_t* labels, void* dlogits, float* loss_sum, int rows, int ncls,
                     float grad_scale, cudaStream_t st) {
  softmax_xent_kernel<<<rows, 256, 0, st>>>((const __nv_bfloat16*)logits, labels, (__nv_bfloat16*)dlogits, loss_sum,
                                            ncls, grad_scale);
  PDJ_CUDA_CHECK(cudaGetLastError());
  return 0;
}

}  // extern "C"
