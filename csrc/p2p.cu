// Single-process peer-memory kernels behind DataParallel: every pointer is a plain device pointer
// of ONE process with peer access enabled, so a kernel running on GPU a can load from / store to
// GPU b over NVLink directly.  Ordering between devices is by CUDA events on the host side.
// Replaces torch._C._broadcast_coalesced, _scatter, _gather and torch.cuda.nccl.reduce, which
// nn.DataParallel calls on every iteration (/root/reference/01.data_parallel.ipynb:275,482,486).
#include "common.cuh"
#include "ptx.cuh"
#include "../include/b200dt.h"

namespace b200 {

struct BcastArgs {
  void* dst[8];
  int ndst;
};

// One read of the source, up to eight peer stores per 16-byte vector (store-based push:
// NVLink writes are posted, so the root never stalls on a peer round trip).
__global__ void __launch_bounds__(512)
p2p_bcast_kernel(const uint4* __restrict__ src, BcastArgs a, long long nvec, long long nbytes) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nvec; i += stride) {
    const uint4 v = ld_nc_v4(src + i);
#pragma unroll
    for (int d = 0; d < 8; ++d)
      if (d < a.ndst) st_v4(reinterpret_cast<uint4*>(a.dst[d]) + i, v);
  }
  // byte tail
  if (blockIdx.x == 0) {
    const unsigned char* s = reinterpret_cast<const unsigned char*>(src);
    for (long long b = nvec * 16 + threadIdx.x; b < nbytes; b += blockDim.x)
      for (int d = 0; d < a.ndst; ++d) reinterpret_cast<unsigned char*>(a.dst[d])[b] = s[b];
  }
}

struct ReduceArgs {
  const float* src[8];
  int nsrc;
};

// Root pulls: dst = (accumulate ? dst : 0) + scale * sum_k src_k, fp32, 16-byte vectors.
__global__ void __launch_bounds__(512)
p2p_reduce_kernel(float* __restrict__ dst, ReduceArgs a, long long n, float scale, int accumulate,
                  int vec) {
  const long long nvec = vec ? n / 4 : 0;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nvec; i += stride) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int k = 0; k < 8; ++k)
      if (k < a.nsrc) {
        const uint4 u = ld_nc_v4(reinterpret_cast<const uint4*>(a.src[k]) + i);
        acc.x += __uint_as_float(u.x); acc.y += __uint_as_float(u.y);
        acc.z += __uint_as_float(u.z); acc.w += __uint_as_float(u.w);
      }
    acc.x *= scale; acc.y *= scale; acc.z *= scale; acc.w *= scale;
    float4* d = reinterpret_cast<float4*>(dst) + i;
    if (accumulate) {
      const float4 o = *d;
      acc.x += o.x; acc.y += o.y; acc.z += o.z; acc.w += o.w;
    }
    *d = acc;
  }
  for (long long j = nvec * 4 + blockIdx.x * (long long)blockDim.x + threadIdx.x; j < n;
       j += stride) {
      float s = 0.f;
      for (int k = 0; k < a.nsrc; ++k) s += a.src[k][j];
      dst[j] = (accumulate ? dst[j] : 0.f) + s * scale;
    }
}

// Gather one replica's output rows into the root tensor and (optionally) fold sum(src) into a
// root-side partial: the output gather fused with the loss reduction.
__global__ void __launch_bounds__(512)
p2p_gather_kernel(const unsigned char* __restrict__ src, unsigned char* __restrict__ dst,
                  long long nbytes, float* partial, int is_f32) {
  __shared__ float red[32];
  const bool vec = (((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0);
  const long long nvec = vec ? nbytes / 16 : 0;
  const long long stride = (long long)gridDim.x * blockDim.x;
  float acc = 0.f;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < nvec; i += stride) {
    const uint4 v = ld_nc_v4(reinterpret_cast<const uint4*>(src) + i);
    st_v4(reinterpret_cast<uint4*>(dst) + i, v);
    if (partial != nullptr) {
      if (is_f32) {
        acc += __uint_as_float(v.x) + __uint_as_float(v.y) + __uint_as_float(v.z) +
               __uint_as_float(v.w);
      } else {
        float f[8];
        unpack8(v, f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc += f[j];
      }
    }
  }
  const int esz = is_f32 ? 4 : 2;
  for (long long b = nvec * 16 + (blockIdx.x * (long long)blockDim.x + threadIdx.x) * esz;
       b < nbytes; b += stride * esz) {
    for (int k = 0; k < esz; ++k) dst[b + k] = src[b + k];
    if (partial != nullptr)
      acc += is_f32 ? *reinterpret_cast<const float*>(src + b)
                    : __bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(src + b));
  }
  if (partial != nullptr) {
    acc = block_sum(acc, red);
    if (threadIdx.x == 0) atomicAdd_system(partial, acc);
  }
}

// Up to eight independent (src, dst, bytes) segments copied by ONE launch: the scatter of DataParallel
// (root slices -> one buffer per peer, stores over NVLink), its gather (peer buffers -> slices of the root
// tensor, loads over NVLink) and the slice hand-out of the two-level broadcast are all this kernel.
struct SegArgs {
  const unsigned char* src[8];
  unsigned char* dst[8];
  long long begin[9];       // prefix sums of the segment sizes in 16-byte vectors (vec) or bytes (!vec)
  int n;
};

__global__ void __launch_bounds__(512)
p2p_multi_copy_kernel(SegArgs a, int vec) {
  const long long total = a.begin[a.n];
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += stride) {
    int s = 0;
#pragma unroll
    for (int k = 1; k < 8; ++k)
      if (k < a.n && i >= a.begin[k]) s = k;
    const long long off = i - a.begin[s];
    if (vec) st_v4(reinterpret_cast<uint4*>(a.dst[s]) + off, ld_nc_v4(reinterpret_cast<const uint4*>(a.src[s]) + off));
    else a.dst[s][off] = a.src[s][off];
  }
}

static int blocks_for(long long nvec, int cap) {
  long long b = (nvec + 511) / 512;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

}  // namespace b200

using namespace b200;

extern "C" int b200dt_enable_peer_access(int dev, int peer) {
  int prev = 0;
  cudaGetDevice(&prev);
  cudaError_t e = cudaSetDevice(dev);
  if (e != cudaSuccess) return (int)e;
  int can = 0;
  cudaDeviceCanAccessPeer(&can, dev, peer);
  if (!can) { cudaSetDevice(prev); return -6; }
  e = cudaDeviceEnablePeerAccess(peer, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); e = cudaSuccess; }
  cudaSetDevice(prev);
  return (int)e;
}

extern "C" int b200dt_p2p_broadcast(const void* src, const uint64_t* dsts_host, int ndst,
                                    long long nbytes, void* stream) {
  if (ndst <= 0 || nbytes <= 0) return 0;
  if (ndst > 8) return -3;
  BcastArgs a{};
  a.ndst = ndst;
  bool vec = (reinterpret_cast<uintptr_t>(src) & 15) == 0;
  for (int d = 0; d < ndst; ++d) {
    a.dst[d] = reinterpret_cast<void*>(dsts_host[d]);
    vec = vec && ((dsts_host[d] & 15) == 0);
  }
  const long long nvec = vec ? nbytes / 16 : 0;
  p2p_bcast_kernel<<<blocks_for(nvec, 64), 512, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const uint4*>(src), a, nvec, nbytes);
  return (int)cudaGetLastError();
}

extern "C" int b200dt_p2p_reduce_f32(float* dst, const uint64_t* srcs_host, int nsrc, long long n,
                                     float scale, int accumulate, void* stream) {
  if (nsrc <= 0 || n <= 0) return 0;
  if (nsrc > 8) return -3;
  ReduceArgs a{};
  a.nsrc = nsrc;
  bool vec = (reinterpret_cast<uintptr_t>(dst) & 15) == 0;
  for (int k = 0; k < nsrc; ++k) {
    a.src[k] = reinterpret_cast<const float*>(srcs_host[k]);
    vec = vec && ((srcs_host[k] & 15) == 0);
  }
  p2p_reduce_kernel<<<blocks_for(vec ? n / 4 : n, 64), 512, 0, (cudaStream_t)stream>>>(
      dst, a, n, scale, accumulate, vec ? 1 : 0);
  return (int)cudaGetLastError();
}

extern "C" int b200dt_p2p_gather_rows(const void* src, void* dst_peer, long long nbytes,
                                      float* partial_peer, int is_f32, void* stream) {
  if (nbytes <= 0) return 0;
  p2p_gather_kernel<<<blocks_for(nbytes / 16, 32), 512, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const unsigned char*>(src), reinterpret_cast<unsigned char*>(dst_peer), nbytes,
      partial_peer, is_f32);
  return (int)cudaGetLastError();
}

extern "C" int b200dt_p2p_multi_copy(const uint64_t* srcs_host, const uint64_t* dsts_host,
                                     const long long* nbytes_host, int n, void* stream) {
  if (n <= 0) return 0;
  if (n > 8) return -3;
  SegArgs a{};
  bool vec = true;
  for (int k = 0; k < n; ++k) {
    if (nbytes_host[k] < 0) return -3;
    vec = vec && ((srcs_host[k] & 15) == 0) && ((dsts_host[k] & 15) == 0) && (nbytes_host[k] % 16 == 0);
  }
  a.n = n;
  a.begin[0] = 0;
  for (int k = 0; k < n; ++k) {
    a.src[k] = reinterpret_cast<const unsigned char*>(srcs_host[k]);
    a.dst[k] = reinterpret_cast<unsigned char*>(dsts_host[k]);
    a.begin[k + 1] = a.begin[k] + (vec ? nbytes_host[k] / 16 : nbytes_host[k]);
  }
  if (a.begin[n] == 0) return 0;
  p2p_multi_copy_kernel<<<blocks_for(a.begin[n], 64), 512, 0, (cudaStream_t)stream>>>(a, vec ? 1 : 0);
  return (int)cudaGetLastError();
}

extern "C" int b200dt_device_sm_count(void) {
  int dev = 0, n = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return -1;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return -1;
  return n;
}

// Plain cudaMemcpyAsync host->device for the small pointer / chunk tables.  Called through the C
// ABI so that torch's pinned-memory allocator never records an event for the source buffer: an
// event recorded while a CUDA graph is being captured cannot be queried later and would poison
// that allocator.
extern "C" int b200dt_memcpy_h2d_async(void* dst, const void* src_pinned, long long nbytes, void* stream) {
  if (nbytes <= 0) return 0;
  return (int)cudaMemcpyAsync(dst, src_pinned, (size_t)nbytes, cudaMemcpyHostToDevice,
                              (cudaStream_t)stream);
}

extern "C" const char* b200dt_error_string(int code) {
  return cudaGetErrorString(static_cast<cudaError_t>(code));
}

extern "C" const char* b200dt_version(void) { return "b200dt 0.1 (sm_100a)"; }
