// Small device helpers shared by every kernel file.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

namespace b200 {

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float2 unpack_bf16x2(uint32_t u) {
  __nv_bfloat162 v = *reinterpret_cast<__nv_bfloat162*>(&u);
  return __bfloat1622float2(v);
}
__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  float2 a = unpack_bf16x2(u.x), b = unpack_bf16x2(u.y), c = unpack_bf16x2(u.z),
         d = unpack_bf16x2(u.w);
  f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint4 u;
  u.x = pack_bf16x2(f[0], f[1]);
  u.y = pack_bf16x2(f[2], f[3]);
  u.z = pack_bf16x2(f[4], f[5]);
  u.w = pack_bf16x2(f[6], f[7]);
  return u;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Block-wide sum; `red` must hold >= 32 floats. Result valid in every thread.
__device__ __forceinline__ float block_sum(float v, float* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  float r = (threadIdx.x < nw) ? red[threadIdx.x] : 0.f;
  if (warp == 0) {
    r = warp_sum(r);
    if (lane == 0) red[0] = r;
  }
  __syncthreads();
  return red[0];
}

// Last-CTA fold of the partial-sum copies behind a BatchNorm `sums` buffer: copy 0 = sum over copies
// 0..nrep-1 (ascending: a fixed order) of n floats each (n % 4 == 0, 16-byte aligned copies, stride n).
// 16-byte L2 loads, eight in flight per thread: the fold used to be a serial chain of scalar loads on one
// CTA and cost 20-40 us per launch on the 1024/2048-channel layers.
__device__ __forceinline__ void fold_sum_copies(float* sums, int n, int nrep, int tid, int nthreads) {
  if (n & 3) {                                       // odd channel counts (never a conv): scalar loads
    for (int i = tid; i < n; i += nthreads) {
      float t = 0.f;
      for (int r = 0; r < nrep; ++r) t += __ldcg(sums + (size_t)r * n + i);
      sums[i] = t;
    }
    return;
  }
  for (int i = tid * 4; i < n; i += nthreads * 4) {
    float4 t = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int r0 = 0; r0 < nrep; r0 += 8) {
      float4 v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u)
        v[u] = (r0 + u < nrep) ? __ldcg(reinterpret_cast<const float4*>(sums + (size_t)(r0 + u) * n + i))
                               : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int u = 0; u < 8; ++u) { t.x += v[u].x; t.y += v[u].y; t.z += v[u].z; t.w += v[u].w; }
    }
    *reinterpret_cast<float4*>(sums + i) = t;
  }
}

inline int ceil_div(long long a, long long b) { return static_cast<int>((a + b - 1) / b); }

// Library-wide switches (csrc/config.cu).  deterministic(): every floating-point sum that is normally
// combined with atomics (BatchNorm statistics, split-K weight gradients, bias gradients) is combined in
// a fixed order instead, so the same launch on the same inputs gives the same bits.
int deterministic();
// Copies of the partial-sum array behind a BatchNorm `sums` buffer: B200DT_BN_REPLICAS normally, one
// per possible CTA slot (B200DT_BN_REPLICAS_DET) in deterministic mode.
int bn_replicas();
// copies a C-channel reduction spreads its atomics over (<= bn_replicas()); the ticket word follows them
int bn_replicas_used(int C);
// the caller hands in zeroed `sums` buffers (b200dt_set_sums_prezeroed): the entry points skip their memset
int sums_prezeroed();

// kernels launched through launch_pdl may begin (prologue only: they call griddep_wait() before touching memory)
// while the previous kernel of the stream drains.  B200DT_PDL is a mask over kernel classes (kPdlGemm,
// kPdlBn); 0 turns the attribute off everywhere.
int pdl_enabled();
constexpr int kPdlGemm = 1, kPdlBn = 2;

// Launch with programmatic stream serialization: the kernel's CTAs may become resident while the tail of the
// previous kernel in the stream is still running, and sit in griddepcontrol.wait until that kernel has completed.
// Hides launch latency and the prologue (barrier init, TMEM allocation, descriptor prefetch) of the ~340 short
// kernels of a ResNet-50 step; captured into CUDA graphs as programmatic edges.  The kernel MUST execute
// griddep_wait() in every thread before its first global-memory access.
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(int cls, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                              cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = (pdl_enabled() & cls) ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

}  // namespace b200
