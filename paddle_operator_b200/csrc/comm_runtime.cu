// Symmetric-memory runtime: per-rank arenas that every rank of a job maps, and
// (when the fabric supports it) one NVLS multicast mapping over all of them.
//
// Reference parity: the reference publishes peers as "<ip>:2379" strings in a
// ConfigMap (controllers/paddlejob_helper.go:215-279) and leaves transport to
// Paddle/NCCL.  Here the rendezvous object additionally carries one POSIX file
// descriptor per rank (cuMemExportToShareableHandle), sent over a unix socket by
// the Python side; this file only does the CUDA VMM part.
//
// Driver entry points are resolved through cudaGetDriverEntryPoint so the
// library has no link-time dependency on libcuda (it must load on CPU-only
// build hosts).
#include <cuda.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include "common.cuh"
#include "../../include/pdjob_b200.h"

static thread_local char g_err[512] = "";
void pdj_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
extern "C" const char* pdj_last_error() { return g_err; }

namespace {

struct Drv {
  bool ok = false;
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*MemRelease)(CUmemGenericAllocationHandle);
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*MemAddressFree)(CUdeviceptr, size_t);
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*MemUnmap)(CUdeviceptr, size_t);
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
  CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long);
  CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
  CUresult (*GetErrorString)(CUresult, const char**);
};
Drv g_drv;

template <typename F>
bool load_sym(const char* name, F* out) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaError_t e = cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q);
  if (e != cudaSuccess || p == nullptr || q != cudaDriverEntryPointSuccess) {
    pdj_set_error("driver entry point %s unavailable (%s)", name, cudaGetErrorString(e));
    return false;
  }
  *out = reinterpret_cast<F>(p);
  return true;
}

bool drv_init() {
  if (g_drv.ok) return true;
  Drv d;
  bool ok = load_sym("cuMemCreate", &d.MemCreate) && load_sym("cuMemRelease", &d.MemRelease) &&
            load_sym("cuMemAddressReserve", &d.MemAddressReserve) &&
            load_sym("cuMemAddressFree", &d.MemAddressFree) && load_sym("cuMemMap", &d.MemMap) &&
            load_sym("cuMemUnmap", &d.MemUnmap) && load_sym("cuMemSetAccess", &d.MemSetAccess) &&
            load_sym("cuMemGetAllocationGranularity", &d.MemGetAllocationGranularity) &&
            load_sym("cuMemExportToShareableHandle", &d.MemExportToShareableHandle) &&
            load_sym("cuMemImportFromShareableHandle", &d.MemImportFromShareableHandle) &&
            load_sym("cuMulticastCreate", &d.MulticastCreate) &&
            load_sym("cuMulticastAddDevice", &d.MulticastAddDevice) &&
            load_sym("cuMulticastBindMem", &d.MulticastBindMem) &&
            load_sym("cuMulticastGetGranularity", &d.MulticastGetGranularity) &&
            load_sym("cuDeviceGetAttribute", &d.DeviceGetAttribute) &&
            load_sym("cuGetErrorString", &d.GetErrorString);
  if (!ok) return false;
  d.ok = true;
  g_drv = d;
  return true;
}

#define DRV_CHECK(call)                                                              \
  do {                                                                               \
    CUresult _r = (call);                                                            \
    if (_r != CUDA_SUCCESS) {                                                        \
      const char* _s = nullptr;                                                      \
      g_drv.GetErrorString(_r, &_s);                                                 \
      pdj_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, _s ? _s : "?");     \
      return (int)_r;                                                                \
    }                                                                                \
  } while (0)

CUmemAllocationProp make_prop(int device) {
  CUmemAllocationProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}

int map_rw(CUdeviceptr va, size_t bytes, int device) {
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  DRV_CHECK(g_drv.MemSetAccess(va, bytes, &acc, 1));
  return 0;
}

}  // namespace

struct PdjArena {
  int device;         // device that maps it (local device)
  size_t bytes;
  CUmemGenericAllocationHandle handle;
  CUdeviceptr va;
  bool imported;
};

struct PdjMulticast {
  size_t bytes;
  CUmemGenericAllocationHandle handle;
  CUdeviceptr va;
  int device;
};

extern "C" {

int pdj_multicast_supported(int device) {
  if (cudaSetDevice(device) != cudaSuccess) return 0;
  cudaFree(0);
  if (!drv_init()) return 0;
  int v = 0;
  if (g_drv.DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, device) != CUDA_SUCCESS) return 0;
  return v;
}

// Rounds `bytes` up so that it is legal both as a cuMemCreate size and (when
// nranks > 1 and multicast exists) as a multicast object size.
int pdj_arena_round_size(int device, size_t bytes, int nranks, size_t* out) {
  PDJ_CUDA_CHECK(cudaSetDevice(device));
  PDJ_CUDA_CHECK(cudaFree(0));
  if (!drv_init()) return -1;
  CUmemAllocationProp prop = make_prop(device);
  size_t gran = 0;
  DRV_CHECK(g_drv.MemGetAllocationGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
  if (nranks > 1 && pdj_multicast_supported(device)) {
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof(mp));
    mp.numDevices = nranks;
    mp.size = bytes;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t mg = 0;
    if (g_drv.MulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg > gran)
      gran = mg;
  }
  *out = (bytes + gran - 1) / gran * gran;
  return 0;
}

int pdj_arena_create(int device, size_t bytes, PdjArena** out) {
  PDJ_CUDA_CHECK(cudaSetDevice(device));
  PDJ_CUDA_CHECK(cudaFree(0));
  if (!drv_init()) return -1;
  CUmemAllocationProp prop = make_prop(device);
  PdjArena* a = new PdjArena();
  a->device = device;
  a->bytes = bytes;
  a->imported = false;
  DRV_CHECK(g_drv.MemCreate(&a->handle, bytes, &prop, 0));
  DRV_CHECK(g_drv.MemAddressReserve(&a->va, bytes, 0, 0, 0));
  DRV_CHECK(g_drv.MemMap(a->va, bytes, 0, a->handle, 0));
  if (int r = map_rw(a->va, bytes, device)) return r;
  PDJ_CUDA_CHECK(cudaMemset(reinterpret_cast<void*>(a->va), 0, bytes));
  PDJ_CUDA_CHECK(cudaDeviceSynchronize());
  *out = a;
  return 0;
}

int pdj_arena_export_fd(PdjArena* a, int* fd) {
  if (!drv_init()) return -1;
  int f = -1;
  DRV_CHECK(g_drv.MemExportToShareableHandle(&f, a->handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *fd = f;
  return 0;
}

// Maps a peer's arena (received as a file descriptor) for access from `device`.
int pdj_arena_import_fd(int device, int fd, size_t bytes, PdjArena** out) {
  PDJ_CUDA_CHECK(cudaSetDevice(device));
  if (!drv_init()) return -1;
  PdjArena* a = new PdjArena();
  a->device = device;
  a->bytes = bytes;
  a->imported = true;
  DRV_CHECK(g_drv.MemImportFromShareableHandle(&a->handle, (void*)(uintptr_t)fd,
                                               CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  DRV_CHECK(g_drv.MemAddressReserve(&a->va, bytes, 0, 0, 0));
  DRV_CHECK(g_drv.MemMap(a->va, bytes, 0, a->handle, 0));
  if (int r = map_rw(a->va, bytes, device)) return r;
  *out = a;
  return 0;
}

void* pdj_arena_ptr(PdjArena* a) { return reinterpret_cast<void*>(a->va); }
size_t pdj_arena_bytes(PdjArena* a) { return a->bytes; }

int pdj_arena_destroy(PdjArena* a) {
  if (!a) return 0;
  if (!drv_init()) return -1;
  cudaSetDevice(a->device);
  cudaDeviceSynchronize();
  g_drv.MemUnmap(a->va, a->bytes);
  g_drv.MemAddressFree(a->va, a->bytes);
  g_drv.MemRelease(a->handle);
  delete a;
  return 0;
}

// --- NVLS multicast --------------------------------------------------------
// Protocol (driven from Python): rank 0 creates the object and exports an fd;
// every rank imports it, adds its device, barriers, binds its own arena at
// offset 0, and maps the multicast object.

int pdj_multicast_create(int device, int nranks, size_t bytes, PdjMulticast** out, int* fd) {
  PDJ_CUDA_CHECK(cudaSetDevice(device));
  if (!drv_init()) return -1;
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  mp.numDevices = nranks;
  mp.size = bytes;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  PdjMulticast* m = new PdjMulticast();
  m->bytes = bytes;
  m->va = 0;
  m->device = device;
  DRV_CHECK(g_drv.MulticastCreate(&m->handle, &mp));
  int f = -1;
  DRV_CHECK(g_drv.MemExportToShareableHandle(&f, m->handle, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0));
  *fd = f;
  *out = m;
  return 0;
}

int pdj_multicast_import(int device, int fd, size_t bytes, PdjMulticast** out) {
  PDJ_CUDA_CHECK(cudaSetDevice(device));
  if (!drv_init()) return -1;
  PdjMulticast* m = new PdjMulticast();
  m->bytes = bytes;
  m->va = 0;
  m->device = device;
  DRV_CHECK(g_drv.MemImportFromShareableHandle(&m->handle, (void*)(uintptr_t)fd,
                                               CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR));
  *out = m;
  return 0;
}

int pdj_multicast_add_device(PdjMulticast* m) {
  if (!drv_init()) return -1;
  DRV_CHECK(g_drv.MulticastAddDevice(m->handle, m->device));
  return 0;
}

// Call only after every rank has added its device.
int pdj_multicast_bind(PdjMulticast* m, PdjArena* local, void** mc_ptr) {
  if (!drv_init()) return -1;
  DRV_CHECK(g_drv.MulticastBindMem(m->handle, 0, local->handle, 0, m->bytes, 0));
  DRV_CHECK(g_drv.MemAddressReserve(&m->va, m->bytes, 0, 0, 0));
  DRV_CHECK(g_drv.MemMap(m->va, m->bytes, 0, m->handle, 0));
  if (int r = map_rw(m->va, m->bytes, m->device)) return r;
  *mc_ptr = reinterpret_cast<void*>(m->va);
  return 0;
}

int pdj_multicast_destroy(PdjMulticast* m) {
  if (!m) return 0;
  if (!drv_init()) return -1;
  if (m->va) {
    g_drv.MemUnmap(m->va, m->bytes);
    g_drv.MemAddressFree(m->va, m->bytes);
  }
  g_drv.MemRelease(m->handle);
  delete m;
  return 0;
}

int pdj_device_count() {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
  return n;
}

int pdj_can_access_peer(int dev, int peer) {
  int v = 0;
  if (cudaDeviceCanAccessPeer(&v, dev, peer) != cudaSuccess) return 0;
  return v;
}

}  // extern "C"
