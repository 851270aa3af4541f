// Shared device helpers for the pdjob_b200 kernels (sm_100a only).
//
// The reference (lfeng-nl/paddle-operator) has no device code at all; it only
// wires PADDLE_TRAINER_ENDPOINTS etc. (controllers/paddlejob_helper.go:258-268)
// so that Paddle can bootstrap its own collectives.  Everything below is the
// data plane this runtime supplies in its place: one process per GPU, peers
// mapped into each other's address space through NVSwitch (cuMem VMM handles),
// system-scope flags for device-side barriers.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#define PDJ_MAX_RANKS 8
// The signal pad has PDJ_MAX_BLOCKS barrier slots, split into channels of
// PDJ_CHANNEL_BLOCKS.  Kernels that may run concurrently on different streams
// (a bucket all-reduce next to an embedding push) must use different channels:
// a slot's epoch counter belongs to one kernel at a time.
#define PDJ_MAX_BLOCKS 1024
#define PDJ_CHANNEL_BLOCKS 256
// Bytes reserved at the head of every rank's arena for the signal pad.
#define PDJ_PAD_BYTES 65536
// Spin-wait bound (ns).  A peer that never shows up turns into a trapped
// kernel (a CUDA error on sync), never into a hung GPU.
#define PDJ_SPIN_TIMEOUT_NS 8000000000ULL

struct PdjSignalPad {
  uint32_t flag[PDJ_MAX_BLOCKS][PDJ_MAX_RANKS];  // flag[block][src_rank], written by src_rank
  uint32_t epoch[PDJ_MAX_BLOCKS];                // local: last epoch consumed by block b
  uint32_t error;                                // set before a timeout trap
  uint32_t reserved[63];
};
static_assert(sizeof(PdjSignalPad) <= PDJ_PAD_BYTES, "signal pad too large");

// Passed by value to every communicating kernel.  base[i] is rank i's arena as
// mapped in THIS process.  With gridDim.y == 1 the kernel acts as `rank`; the
// single-GPU emulation used by the tests launches gridDim.y == nranks
// cooperatively over nranks arenas that all live on one device, and each
// y-slice acts as rank + blockIdx.y.
struct PdjComm {
  int nranks;
  int rank;
  int pad_base;  // first barrier slot of this communicator view (channel * PDJ_CHANNEL_BLOCKS)
  int spin_timeout_ms;  // 0 = PDJ_SPIN_TIMEOUT_NS; elastic jobs lower it to detect a lost peer sooner
  char* base[PDJ_MAX_RANKS];
  char* mc_base;  // NVLS multicast mapping of the arena (nullptr when unavailable)
};

__device__ __forceinline__ void pdj_st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t pdj_ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long pdj_globaltimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// 16-byte loads/stores that are safe on peer-mapped memory written by another
// GPU during this kernel's lifetime (so no .nc), and that do not pollute L1.
__device__ __forceinline__ uint4 pdj_ld16(const void* p) {
  uint4 v;
  asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void pdj_st16(void* p, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

__device__ __forceinline__ PdjSignalPad* pdj_pad(const PdjComm& c, int r) {
  return reinterpret_cast<PdjSignalPad*>(c.base[r]);
}

// Barrier between block blockIdx.x of every rank.  `epoch` must be the same
// on all ranks and strictly increasing per (block index) across kernels; the
// kernels read it from pad->epoch[] so that CUDA-graph replays stay in step.
__device__ __forceinline__ int pdj_slot(const PdjComm& c) { return c.pad_base + (int)blockIdx.x; }

__device__ __forceinline__ void pdj_block_barrier(const PdjComm& c, int rank, uint32_t epoch) {
  __syncthreads();
  const int t = threadIdx.x;
  const int slot = pdj_slot(c);
  if (t < c.nranks) {
    __threadfence_system();
    pdj_st_release_sys(&pdj_pad(c, t)->flag[slot][rank], epoch);
    PdjSignalPad* me = pdj_pad(c, rank);
    const uint32_t* f = &me->flag[slot][t];
    const unsigned long long t0 = pdj_globaltimer();
    const unsigned long long limit =
        c.spin_timeout_ms > 0 ? (unsigned long long)c.spin_timeout_ms * 1000000ULL : PDJ_SPIN_TIMEOUT_NS;
    uint32_t spins = 0;
    while ((int32_t)(pdj_ld_acquire_sys(f) - epoch) < 0) {
      if (((++spins) & 0x3ff) == 0 && pdj_globaltimer() - t0 > limit) {
        me->error = 1u + (uint32_t)t;
        __threadfence_system();
        __trap();
      }
    }
  }
  __syncthreads();
}

__device__ __forceinline__ float2 pdj_bf2_to_f2(uint32_t u) {
  // bf16 -> fp32 is a 16-bit shift.
  return make_float2(__uint_as_float(u << 16), __uint_as_float(u & 0xffff0000u));
}
__device__ __forceinline__ uint32_t pdj_f2_to_bf2(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

#define PDJ_CUDA_CHECK(expr)                                                       \
  do {                                                                             \
    cudaError_t _e = (expr);                                                       \
    if (_e != cudaSuccess) {                                                       \
      pdj_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
      return (int)_e ? (int)_e : -1;                                               \
    }                                                                              \
  } while (0)

// Host: last error text (thread-local), see pdj_last_error().
void pdj_set_error(const char* fmt, ...);
// Host: encodes a 2-D bf16 TMA tensor map (CUtensorMap, 128 bytes at `map_out`) for a row-major
// [rows, cols] tensor with a row pitch of ld_elems; box = box_rows x box_cols.  Defined in tma_util.cu.
int pdj_tma_map_2d_bf16(void* map_out, const void* ptr, int rows, int cols, int ld_elems, int box_cols, int box_rows,
                        int swizzle128);
