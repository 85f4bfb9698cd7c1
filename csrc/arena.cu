// Symmetric arena on the CUDA driver API: one physical allocation per rank, peer-mapped into every
// process of the node through POSIX file-descriptor handles, plus an NVLS multicast object bound to all
// of them (cuMulticastCreate / AddDevice / BindMem).  This is the memory the collectives of
// csrc/allreduce.cu run on; it replaces what the reference gets from ProcessGroupNCCL's registered
// buffers behind /root/reference/ddp_gpus.py:16,32 (NCCL sets up the same mappings internally).
//
// The file descriptors travel between processes out of band (the Python side passes them over
// AF_UNIX sockets, ops/arena.py); every function here is local to the calling process.  Multicast is
// optional: when the device or the driver refuses it the arena still works for one-/two-shot kernels.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <unistd.h>
#include "../include/b200dt.h"

namespace b200 {

#define B200_DRV_FN(name, ver_name)                                                             \
  static decltype(&name) p_##name() {                                                            \
    static decltype(&name) fn = nullptr;                                                         \
    if (fn == nullptr) {                                                                         \
      void* sym = nullptr;                                                                       \
      cudaDriverEntryPointQueryResult q;                                                         \
      if (cudaGetDriverEntryPoint(ver_name, &sym, cudaEnableDefault, &q) == cudaSuccess &&       \
          q == cudaDriverEntryPointSuccess)                                                      \
        fn = reinterpret_cast<decltype(&name)>(sym);                                             \
    }                                                                                            \
    return fn;                                                                                   \
  }

B200_DRV_FN(cuMemCreate, "cuMemCreate")
B200_DRV_FN(cuMemRelease, "cuMemRelease")
B200_DRV_FN(cuMemGetAllocationGranularity, "cuMemGetAllocationGranularity")
B200_DRV_FN(cuMemExportToShareableHandle, "cuMemExportToShareableHandle")
B200_DRV_FN(cuMemImportFromShareableHandle, "cuMemImportFromShareableHandle")
B200_DRV_FN(cuMemAddressReserve, "cuMemAddressReserve")
B200_DRV_FN(cuMemAddressFree, "cuMemAddressFree")
B200_DRV_FN(cuMemMap, "cuMemMap")
B200_DRV_FN(cuMemUnmap, "cuMemUnmap")
B200_DRV_FN(cuMemSetAccess, "cuMemSetAccess")
B200_DRV_FN(cuMulticastCreate, "cuMulticastCreate")
B200_DRV_FN(cuMulticastAddDevice, "cuMulticastAddDevice")
B200_DRV_FN(cuMulticastBindMem, "cuMulticastBindMem")
B200_DRV_FN(cuMulticastUnbind, "cuMulticastUnbind")
B200_DRV_FN(cuMulticastGetGranularity, "cuMulticastGetGranularity")
B200_DRV_FN(cuDeviceGet, "cuDeviceGet")
B200_DRV_FN(cuDeviceGetAttribute, "cuDeviceGetAttribute")

struct Arena {
  int dev = 0;
  int rank = 0, world = 1;
  size_t size = 0;              // mapped bytes (multiple of the granularity)
  CUmemGenericAllocationHandle local = 0;
  CUmemGenericAllocationHandle peer_h[B200DT_MAX_WORLD] = {0};
  CUdeviceptr va[B200DT_MAX_WORLD] = {0};
  bool mapped[B200DT_MAX_WORLD] = {false};
  CUmemGenericAllocationHandle mc = 0;
  CUdeviceptr mc_va = 0;
  bool mc_bound = false, mc_mapped = false;
  int last_error = 0;           // CUresult of the last failing driver call
};

static int fail(Arena* a, CUresult r) {
  if (a != nullptr) a->last_error = (int)r;
  return -7;
}

static int map_handle(Arena* a, CUmemGenericAllocationHandle h, CUdeviceptr* va) {
  CUresult r = p_cuMemAddressReserve()(va, a->size, 0, 0, 0);
  if (r != CUDA_SUCCESS) return fail(a, r);
  r = p_cuMemMap()(*va, a->size, 0, h, 0);
  if (r != CUDA_SUCCESS) { p_cuMemAddressFree()(*va, a->size); *va = 0; return fail(a, r); }
  CUmemAccessDesc acc{};
  acc.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  acc.location.id = a->dev;
  acc.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  r = p_cuMemSetAccess()(*va, a->size, &acc, 1);
  if (r != CUDA_SUCCESS) {
    p_cuMemUnmap()(*va, a->size); p_cuMemAddressFree()(*va, a->size); *va = 0;
    return fail(a, r);
  }
  return 0;
}

}  // namespace b200

using namespace b200;

// 1 if device `dev` can take part in an NVLS multicast object, 0 if not, < 0 on error.
extern "C" int b200dt_multicast_supported(int dev) {
  if (!p_cuDeviceGet() || !p_cuDeviceGetAttribute() || !p_cuMulticastCreate()) return 0;
  cudaFree(0);                                          // make sure the runtime initialised the driver
  CUdevice d;
  if (p_cuDeviceGet()(&d, dev) != CUDA_SUCCESS) return -7;
  int v = 0;
  if (p_cuDeviceGetAttribute()(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, d) != CUDA_SUCCESS) return -7;
  return v;
}

// Allocate this rank's share: `nbytes` rounded up to the allocation / multicast granularity, exportable
// as a POSIX fd, mapped read-write on `dev`.  *out receives the arena, *size_out the rounded size.
extern "C" int b200dt_arena_create(int dev, int rank, int world, long long nbytes, void** out,
                                   long long* size_out) {
  if (world < 1 || world > B200DT_MAX_WORLD || rank < 0 || rank >= world || nbytes <= 0) return -3;
  if (!p_cuMemCreate() || !p_cuMemMap() || !p_cuMemAddressReserve() || !p_cuMemSetAccess()) return -1;
  cudaError_t ce = cudaSetDevice(dev);
  if (ce != cudaSuccess) return (int)ce;
  cudaFree(0);
  Arena* a = new Arena();
  a->dev = dev; a->rank = rank; a->world = world;
  CUmemAllocationProp prop{};
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = dev;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  size_t gran = 0;
  CUresult r = p_cuMemGetAllocationGranularity()(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED);
  if (r != CUDA_SUCCESS) { int e = fail(a, r); delete a; return e; }
  if (p_cuMulticastGetGranularity() && b200dt_multicast_supported(dev) == 1) {
    CUmulticastObjectProp mp{};
    mp.numDevices = world > 1 ? world : 2;
    mp.size = (size_t)nbytes;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t mg = 0;
    if (p_cuMulticastGetGranularity()(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg > gran)
      gran = mg;
  }
  a->size = ((size_t)nbytes + gran - 1) / gran * gran;
  r = p_cuMemCreate()(&a->local, a->size, &prop, 0);
  if (r != CUDA_SUCCESS) { int e = fail(a, r); delete a; return e; }
  int e = map_handle(a, a->local, &a->va[rank]);
  if (e) { p_cuMemRelease()(a->local); delete a; return e; }
  a->mapped[rank] = true;
  cudaMemset(reinterpret_cast<void*>(a->va[rank]), 0, a->size);
  cudaDeviceSynchronize();
  *out = a;
  if (size_out) *size_out = (long long)a->size;
  return 0;
}

// POSIX fd for this rank's allocation (the caller sends it to the peers and closes it afterwards).
extern "C" int b200dt_arena_export_fd(void* arena, int* fd) {
  Arena* a = reinterpret_cast<Arena*>(arena);
  int f = -1;
  CUresult r = p_cuMemExportToShareableHandle()(&f, a->local, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) return fail(a, r);
  *fd = f;
  return 0;
}

// Map peer `peer`'s allocation from an fd that was received into THIS process.
extern "C" int b200dt_arena_map_peer(void* arena, int peer, int fd) {
  Arena* a = reinterpret_cast<Arena*>(arena);
  if (peer < 0 || peer >= a->world || peer == a->rank || a->mapped[peer]) return -3;
  cudaSetDevice(a->dev);
  CUresult r = p_cuMemImportFromShareableHandle()(&a->peer_h[peer], (void*)(uintptr_t)fd,
                                                   CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  if (r != CUDA_SUCCESS) return fail(a, r);
  int e = map_handle(a, a->peer_h[peer], &a->va[peer]);
  if (e) { p_cuMemRelease()(a->peer_h[peer]); a->peer_h[peer] = 0; return e; }
  a->mapped[peer] = true;
  return 0;
}

// Rank 0: create the multicast object for `world` devices and export it as an fd.
extern "C" int b200dt_arena_mc_create(void* arena, int* fd) {
  Arena* a = reinterpret_cast<Arena*>(arena);
  if (!p_cuMulticastCreate()) return -1;
  cudaSetDevice(a->dev);
  CUmulticastObjectProp mp{};
  mp.numDevices = a->world;
  mp.size = a->size;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  CUresult r = p_cuMulticastCreate()(&a->mc, &mp);
  if (r != CUDA_SUCCESS) { a->mc = 0; return fail(a, r); }
  int f = -1;
  r = p_cuMemExportToShareableHandle()(&f, a->mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
  if (r != CUDA_SUCCESS) { p_cuMemRelease()(a->mc); a->mc = 0; return fail(a, r); }
  *fd = f;
  return 0;
}

// Other ranks: adopt the multicast object from a received fd.
extern "C" int b200dt_arena_mc_import(void* arena, int fd) {
  Arena* a = reinterpret_cast<Arena*>(arena);
  cudaSetDevice(a->dev);
  CUresult r = p_cuMemImportFromShareableHandle()(&a->mc, (void*)(uintptr_t)fd,
                                                   CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  if (r != CUDA_SUCCESS) { a->mc = 0; return fail(a, r); }
  return 0;
}

// Every rank: add its device to the object.  ALL ranks must have returned from this call (barrier)
// before any of them calls b200dt_arena_mc_bind.
extern "C" int b200dt_arena_mc_add_device(void* arena) {
  Arena* a = reinterpret_cast<Arena*>(arena);
  if (a->mc == 0) return -3;
  cudaSetDevice(a->dev);
  CUdevice d;
  CUresult r = p_cuDeviceGet()(&d, a->dev);
  if (r != CUDA_SUCCESS) return fail(a, r);
  r = p_cuMulticastAddDevice()(a->mc, d);
  if (r != CUDA_SUCCESS) return fail(a, r);
  return 0;
}

// Every rank: bind its physical allocation at offset 0 of the object and map the multicast address.
extern "C" int b200dt_arena_mc_bind(void* arena) {
  Arena* a = reinterpret_cast<Arena*>(arena);
  if (a->mc == 0) return -3;
  cudaSetDevice(a->dev);
  CUresult r = p_cuMulticastBindMem()(a->mc, 0, a->local, 0, a->size, 0);
  if (r != CUDA_SUCCESS) return fail(a, r);
  a->mc_bound = true;
  int e = map_handle(a, a->mc, &a->mc_va);
  if (e) return e;
  a->mc_mapped = true;
  return 0;
}

extern "C" unsigned long long b200dt_arena_ptr(void* arena, int peer) {
  Arena* a = reinterpret_cast<Arena*>(arena);
  return (peer >= 0 && peer < a->world && a->mapped[peer]) ? (unsigned long long)a->va[peer] : 0ull;
}
extern "C" unsigned long long b200dt_arena_mc_ptr(void* arena) {
  Arena* a = reinterpret_cast<Arena*>(arena);
  return a->mc_mapped ? (unsigned long long)a->mc_va : 0ull;
}
extern "C" int b200dt_arena_last_error(void* arena) { return reinterpret_cast<Arena*>(arena)->last_error; }

extern "C" int b200dt_arena_destroy(void* arena) {
  Arena* a = reinterpret_cast<Arena*>(arena);
  if (a == nullptr) return 0;
  cudaSetDevice(a->dev);
  cudaDeviceSynchronize();
  if (a->mc_mapped) { p_cuMemUnmap()(a->mc_va, a->size); p_cuMemAddressFree()(a->mc_va, a->size); }
  if (a->mc_bound) {
    CUdevice d;
    if (p_cuDeviceGet()(&d, a->dev) == CUDA_SUCCESS) p_cuMulticastUnbind()(a->mc, d, 0, a->size);
  }
  if (a->mc) p_cuMemRelease()(a->mc);
  for (int r = 0; r < a->world; ++r) {
    if (!a->mapped[r]) continue;
    p_cuMemUnmap()(a->va[r], a->size);
    p_cuMemAddressFree()(a->va[r], a->size);
    if (r != a->rank && a->peer_h[r]) p_cuMemRelease()(a->peer_h[r]);
  }
  if (a->local) p_cuMemRelease()(a->local);
  delete a;
  return 0;
}

// Status word the collectives report time-outs through: pinned host memory mapped into the device,
// so the host can read it at any time without touching a stream.
extern "C" int b200dt_status_alloc(int** host_ptr, unsigned long long* dev_ptr) {
  int* h = nullptr;
  cudaError_t e = cudaHostAlloc(reinterpret_cast<void**>(&h), 64, cudaHostAllocMapped | cudaHostAllocPortable);
  if (e != cudaSuccess) return (int)e;
  memset(h, 0, 64);
  void* d = nullptr;
  e = cudaHostGetDevicePointer(&d, h, 0);
  if (e != cudaSuccess) { cudaFreeHost(h); return (int)e; }
  *host_ptr = h;
  *dev_ptr = (unsigned long long)d;
  return 0;
}
extern "C" int b200dt_status_free(int* host_ptr) { return (int)cudaFreeHost(host_ptr); }
