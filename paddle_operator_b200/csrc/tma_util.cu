// Host-side TMA utilities: tensor-map encoding for the tcgen05 / TMA kernels.
#include <cuda.h>
#include <cuda_bf16.h>

#include "common.cuh"
#include "tcgen05_common.cuh"

namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn g_encode = nullptr;

// cuTensorMapEncodeTiled is a driver call and needs a current context.  A thread that has only made
// context-free runtime calls so far (an autograd worker entering backward) has none: bind the primary
// context of its current device.
int ensure_current_context() {
  typedef CUresult (*GetCur)(CUcontext*);
  typedef CUresult (*SetCur)(CUcontext);
  typedef CUresult (*Retain)(CUcontext*, CUdevice);
  typedef CUresult (*DevGet)(CUdevice*, int);
  static GetCur get_cur = nullptr;
  static SetCur set_cur = nullptr;
  static Retain retain = nullptr;
  static DevGet dev_get = nullptr;
  if (get_cur == nullptr) {
    void *a = nullptr, *b = nullptr, *c = nullptr, *d = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuCtxGetCurrent", &a, cudaEnableDefault, &q) != cudaSuccess || a == nullptr ||
        cudaGetDriverEntryPoint("cuCtxSetCurrent", &b, cudaEnableDefault, &q) != cudaSuccess || b == nullptr ||
        cudaGetDriverEntryPoint("cuDevicePrimaryCtxRetain", &c, cudaEnableDefault, &q) != cudaSuccess || c == nullptr ||
        cudaGetDriverEntryPoint("cuDeviceGet", &d, cudaEnableDefault, &q) != cudaSuccess || d == nullptr) {
      pdj_set_error("driver context entry points are unavailable");
      return -1;
    }
    set_cur = reinterpret_cast<SetCur>(b);
    retain = reinterpret_cast<Retain>(c);
    dev_get = reinterpret_cast<DevGet>(d);
    get_cur = reinterpret_cast<GetCur>(a);
  }
  CUcontext ctx = nullptr;
  if (get_cur(&ctx) == CUDA_SUCCESS && ctx != nullptr) return 0;
  int dev = 0;
  PDJ_CUDA_CHECK(cudaGetDevice(&dev));
  CUdevice cudev;
  CUresult r = dev_get(&cudev, dev);
  if (r == CUDA_SUCCESS) r = retain(&ctx, cudev);
  if (r == CUDA_SUCCESS) r = set_cur(ctx);
  if (r != CUDA_SUCCESS) {
    pdj_set_error("could not bind the primary context of device %d to this thread (CUresult %d)", dev, (int)r);
    return (int)r;
  }
  return 0;
}

int resolve_encode() {
  if (g_encode != nullptr) return 0;
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
  if (e != cudaSuccess || p == nullptr || q != cudaDriverEntryPointSuccess) {
    pdj_set_error("cuTensorMapEncodeTiled is unavailable: %s", cudaGetErrorString(e));
    return -1;
  }
  g_encode = reinterpret_cast<EncodeTiledFn>(p);
  return 0;
}

int make_map(CUtensorMap* map, const void* ptr, int rows, int cols, int ld_elems, int box_cols, int box_rows, bool swizzle) {
  if (int r = resolve_encode()) return r;
  if (int r = ensure_current_context()) return r;
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};       // innermost first
  const cuuint64_t strides[1] = {(cuuint64_t)ld_elems * 2};              // bytes, for dim 1
  const cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    pdj_set_error("cuTensorMapEncodeTiled failed with CUresult %d (rows=%d cols=%d ld=%d box=%dx%d ptr=%p)", (int)r, rows, cols,
                  ld_elems, box_rows, box_cols, ptr);
    return (int)r;
  }
  return 0;
}

// 3-D map {d0 contiguous elements, d1 with stride s1 bytes, d2 with stride s2 bytes}, 128-byte swizzle.
int make_map3(CUtensorMap* map, const void* ptr, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t s1, uint64_t s2, int b0,
              int b1) {
  if (int r = resolve_encode()) return r;
  if (int r = ensure_current_context()) return r;
  const cuuint64_t dims[3] = {d0, d1, d2};
  const cuuint64_t strides[2] = {s1, s2};
  const cuuint32_t box[3] = {(cuuint32_t)b0, (cuuint32_t)b1, 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    pdj_set_error("cuTensorMapEncodeTiled (3-D) failed with CUresult %d (dims %llu x %llu x %llu, strides %llu, %llu)", (int)r,
                  (unsigned long long)d0, (unsigned long long)d1, (unsigned long long)d2, (unsigned long long)s1,
                  (unsigned long long)s2);
    return (int)r;
  }
  return 0;
}

typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const int*, const int*, cuuint32_t, cuuint32_t, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeIm2colFn g_encode_im2col = nullptr;

int resolve_encode_im2col() {
  if (g_encode_im2col != nullptr) return 0;
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &p, cudaEnableDefault, &q);
  if (e != cudaSuccess || p == nullptr || q != cudaDriverEntryPointSuccess) {
    pdj_set_error("cuTensorMapEncodeIm2col is unavailable: %s", cudaGetErrorString(e));
    return -1;
  }
  g_encode_im2col = reinterpret_cast<EncodeIm2colFn>(p);
  return 0;
}

int g_sms = 0;

}  // namespace

int pdj_tc_make_map4(CUtensorMap* map, const void* ptr, const uint64_t dims[4], const uint64_t strides[3], const uint32_t box[4],
                     int swizzle) {
  if (int r = resolve_encode()) return r;
  if (int r = ensure_current_context()) return r;
  const cuuint64_t d[4] = {dims[0], dims[1], dims[2], dims[3]};
  const cuuint64_t s[3] = {strides[0], strides[1], strides[2]};
  const cuuint32_t b[4] = {box[0], box[1], box[2], box[3]};
  const cuuint32_t estr[4] = {1, 1, 1, 1};
  static const CUtensorMapSwizzle modes[4] = {CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_SWIZZLE_32B, CU_TENSOR_MAP_SWIZZLE_64B,
                                              CU_TENSOR_MAP_SWIZZLE_128B};
  CUresult r = g_encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(ptr), d, s, b, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        modes[swizzle & 3], CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    pdj_set_error("cuTensorMapEncodeTiled (4-D) failed with CUresult %d (dims %llu x %llu x %llu x %llu, box %u x %u x %u x %u)", (int)r,
                  (unsigned long long)d[0], (unsigned long long)d[1], (unsigned long long)d[2], (unsigned long long)d[3], b[0], b[1],
                  b[2], b[3]);
    return (int)r;
  }
  return 0;
}

int pdj_tc_make_im2col4(CUtensorMap* map, const void* ptr, int C, int W, int H, int N, int ksize, int pad, int stride, int channels,
                        int pixels) {
  if (int r = resolve_encode_im2col()) return r;
  if (int r = ensure_current_context()) return r;
  const cuuint64_t d[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  const cuuint64_t s[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
  // bounding box of the positions the top-left filter tap may take: [-pad, dim + pad - ksize] in W and H
  const int lower[2] = {-pad, -pad};
  const int upper[2] = {pad - (ksize - 1), pad - (ksize - 1)};
  const cuuint32_t estr[4] = {1, (cuuint32_t)stride, (cuuint32_t)stride, 1};
  CUresult r = g_encode_im2col(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(ptr), d, s, lower, upper,
                               (cuuint32_t)channels, (cuuint32_t)pixels, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                               CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    pdj_set_error("cuTensorMapEncodeIm2col failed with CUresult %d (C=%d W=%d H=%d N=%d k=%d pad=%d stride=%d)", (int)r, C, W, H, N,
                  ksize, pad, stride);
    return (int)r;
  }
  return 0;
}

int pdj_tc_make_map(CUtensorMap* map, const void* ptr, int rows, int cols, int ld_elems, int box_cols, int box_rows,
                    bool swizzle) {
  return make_map(map, ptr, rows, cols, ld_elems, box_cols, box_rows, swizzle);
}

int pdj_tc_make_map3(CUtensorMap* map, const void* ptr, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t s1, uint64_t s2,
                     int b0, int b1) {
  return make_map3(map, ptr, d0, d1, d2, s1, s2, b0, b1);
}

int pdj_tc_make_map_kn(CUtensorMap* map, const void* ptr, int K, int N, int ld_elems, int box_k, int groups) {
  const uint64_t dims[4] = {64, (uint64_t)K, (uint64_t)(N / 64), 1};
  const uint64_t strides[3] = {(uint64_t)ld_elems * 2, 128, (uint64_t)ld_elems * 2 * (uint64_t)K};
  const uint32_t box[4] = {64, (uint32_t)box_k, (uint32_t)groups, 1};
  return pdj_tc_make_map4(map, ptr, dims, strides, box, 3);
}

int pdj_tc_sm_count() {
  if (g_sms == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&g_sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || g_sms <= 0)
      g_sms = 148;
  }
  return g_sms;
}

int pdj_tma_map_2d_bf16(void* map_out, const void* ptr, int rows, int cols, int ld_elems, int box_cols, int box_rows,
                        int swizzle128) {
  return make_map(reinterpret_cast<CUtensorMap*>(map_out), ptr, rows, cols, ld_elems, box_cols, box_rows, swizzle128 != 0);
}
