// vmm.cu — peer-mapped shard memory through the CUDA virtual-memory-management API.
//
// Every rank allocates its shard as one physical allocation exportable as a POSIX file
// descriptor, and maps its own and every peer's allocation side by side into ONE reserved
// virtual range:  [ shard 0 | shard 1 | ... | shard W-1 ],  stride = shard bytes.
// Kernels then address any shard with ordinary pointers (NVLink 5 / NVSwitch carries the
// loads, stores, `red` and `atom` operations), and the whole distributed table is a single 2-D
// tensor that one TMA tensor map can describe.
//
// This replaces the reference's transport layer (ref: core/transfer/transfer.h:55-268,
// Route.h:20-112: one ZeroMQ PUSH socket per peer over TCP); file descriptors are exchanged
// by the Python layer over Unix sockets (SCM_RIGHTS), the analogue of the master's route
// broadcast (system/master/init.h:74-99).
//
// The driver API is resolved at run time with cudaGetDriverEntryPoint so the library has no
// link-time dependency on libcuda (it must load on the CPU-only build box too).
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <unistd.h>
#include <cstring>
#include "../../include/ssb200.h"

namespace {

template <typename Fn>
int resolve(const char* name, Fn* out) {
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaError_t e = cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q);
  if (e != cudaSuccess || p == nullptr) return e != cudaSuccess ? (int)e : (int)cudaErrorSymbolNotFound;
  *out = reinterpret_cast<Fn>(p);
  return 0;
}

struct Api {
  CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*MemRelease)(CUmemGenericAllocationHandle);
  CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
  CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*MemAddressFree)(CUdeviceptr, size_t);
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*MemUnmap)(CUdeviceptr, size_t);
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  bool ok = false;
};

Api g_api;

int load_api() {
  if (g_api.ok) return 0;
  cudaFree(0);  // make sure the primary context exists
  int e;
  if ((e = resolve("cuMemGetAllocationGranularity", &g_api.MemGetAllocationGranularity))) return e;
  if ((e = resolve("cuMemCreate", &g_api.MemCreate))) return e;
  if ((e = resolve("cuMemRelease", &g_api.MemRelease))) return e;
  if ((e = resolve("cuMemExportToShareableHandle", &g_api.MemExportToShareableHandle))) return e;
  if ((e = resolve("cuMemImportFromShareableHandle", &g_api.MemImportFromShareableHandle))) return e;
  if ((e = resolve("cuMemAddressReserve", &g_api.MemAddressReserve))) return e;
  if ((e = resolve("cuMemAddressFree", &g_api.MemAddressFree))) return e;
  if ((e = resolve("cuMemMap", &g_api.MemMap))) return e;
  if ((e = resolve("cuMemUnmap", &g_api.MemUnmap))) return e;
  if ((e = resolve("cuMemSetAccess", &g_api.MemSetAccess))) return e;
  g_api.ok = true;
  return 0;
}

CUmemAllocationProp make_prop(int device) {
  CUmemAllocationProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = device;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return prop;
}

// CUresult values are returned offset so Python can tell them from cudaError_t
inline int cu(CUresult r) { return r == CUDA_SUCCESS ? 0 : 100000 + (int)r; }

}  // namespace

extern "C" {

int ssb_vmm_granularity(int device, uint64_t* out) {
  if (int e = load_api()) return e;
  CUmemAllocationProp prop = make_prop(device);
  size_t g = 0;
  CUresult r = g_api.MemGetAllocationGranularity(&g, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED);
  *out = g;
  return cu(r);
}

// Allocate `bytes` (multiple of the granularity) of device memory; returns the generic handle
// and an exported POSIX fd (caller closes it after the peers imported it).
int ssb_vmm_create(int device, uint64_t bytes, uint64_t* handle, int* fd) {
  if (int e = load_api()) return e;
  CUmemAllocationProp prop = make_prop(device);
  CUmemGenericAllocationHandle h;
  CUresult r = g_api.MemCreate(&h, bytes, &prop, 0);
  if (r != CUDA_SUCCESS) return cu(r);
  int f = -1;
  r = g_api.MemExportToShareableHandle(&f, h, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) { g_api.MemRelease(h); return cu(r); }
  *handle = (uint64_t)h;
  *fd = f;
  return 0;
}

int ssb_vmm_import(int fd, uint64_t* handle) {
  if (int e = load_api()) return e;
  CUmemGenericAllocationHandle h;
  CUresult r = g_api.MemImportFromShareableHandle(&h, (void*)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  if (r != CUDA_SUCCESS) return cu(r);
  *handle = (uint64_t)h;
  return 0;
}

int ssb_vmm_reserve(uint64_t bytes, uint64_t align, uint64_t* va) {
  if (int e = load_api()) return e;
  CUdeviceptr p = 0;
  CUresult r = g_api.MemAddressReserve(&p, bytes, align, 0, 0);
  *va = (uint64_t)p;
  return cu(r);
}

// Map `handle` at `va` and give `device` read/write access to it.
int ssb_vmm_map(uint64_t va, uint64_t bytes, uint64_t handle, int device) {
  if (int e = load_api()) return e;
  CUresult r = g_api.MemMap((CUdeviceptr)va, bytes, 0, (CUmemGenericAllocationHandle)handle, 0);
  if (r != CUDA_SUCCESS) return cu(r);
  CUmemAccessDesc acc;
  memset(&acc, 0, sizeof(acc));
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = device;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = g_api.MemSetAccess((CUdeviceptr)va, bytes, &acc, 1);
  return cu(r);
}

int ssb_vmm_unmap(uint64_t va, uint64_t bytes) {
  if (int e = load_api()) return e;
  return cu(g_api.MemUnmap((CUdeviceptr)va, bytes));
}

int ssb_vmm_release(uint64_t handle) {
  if (int e = load_api()) return e;
  return cu(g_api.MemRelease((CUmemGenericAllocationHandle)handle));
}

int ssb_vmm_free_va(uint64_t va, uint64_t bytes) {
  if (int e = load_api()) return e;
  return cu(g_api.MemAddressFree((CUdeviceptr)va, bytes));
}

int ssb_memset_async(void* p, int value, uint64_t bytes, void* stream) {
  return (int)cudaMemsetAsync(p, value, bytes, (cudaStream_t)stream);
}

}  // extern "C"
