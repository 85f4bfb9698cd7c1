// Loss kernels: each call produces the scalar loss AND d(loss)/d(input) in one pass.
// Replaces F.cross_entropy with probability targets (/root/reference/ddp_gpus.py:37), nn.MSELoss
// (/root/reference/03.model_parallel.ipynb:533,976) and output.sum() (01.data_parallel.ipynb:485).
//
// The scalar is reduced in two levels without a second launch: every CTA writes its partial to
// the scratch area, the last CTA to finish (ticket counter in the scratch area) adds the partials
// in a fixed order, so the result is deterministic.  The counter is left at zero for the next call.
#include "common.cuh"
#include "ptx.cuh"
#include "../include/b200dt.h"

namespace b200 {

constexpr int kLossMaxBlocks = 1024;
struct LossScratch {
  float partial[kLossMaxBlocks];
  unsigned int ticket;
  unsigned int pad[3];
};

// Called by all threads of the CTA with the CTA's partial in `v` (valid in thread 0).
__device__ __forceinline__ void finish_scalar(float v, LossScratch* sc, float* out, float scale,
                                              float* red) {
  __shared__ bool is_last;
  if (threadIdx.x == 0) {
    sc->partial[blockIdx.x] = v;
    __threadfence();
    const unsigned int t = atomicAdd(&sc->ticket, 1u);
    is_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (is_last) {
    __threadfence();
    float s = 0.f;
    for (int i = threadIdx.x; i < (int)gridDim.x; i += blockDim.x)
      s += *reinterpret_cast<volatile float*>(&sc->partial[i]);
    s = block_sum(s, red);
    if (threadIdx.x == 0) {
      *out = s * scale;
      sc->ticket = 0;
    }
  }
}

template <typename T> __device__ __forceinline__ float ldf(const T* p, long long i);
template <> __device__ __forceinline__ float ldf<float>(const float* p, long long i) { return p[i]; }
template <> __device__ __forceinline__ float ldf<__nv_bfloat16>(const __nv_bfloat16* p, long long i) {
  return __bfloat162float(p[i]);
}
template <typename T> __device__ __forceinline__ void stf(T* p, long long i, float v);
template <> __device__ __forceinline__ void stf<float>(float* p, long long i, float v) { p[i] = v; }
template <> __device__ __forceinline__ void stf<__nv_bfloat16>(__nv_bfloat16* p, long long i, float v) {
  p[i] = __float2bfloat16(v);
}

template <typename T>
__global__ void __launch_bounds__(256)
mse_kernel(const T* __restrict__ pred, const T* __restrict__ target, T* __restrict__ grad,
           float* loss, long long n, float gscale, LossScratch* sc) {
  __shared__ float red[32];
  float acc = 0.f;
  const float gs = 2.f * gscale / (float)n;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    const float d = ldf(pred, i) - ldf(target, i);
    acc += d * d;
    if (grad != nullptr) stf(grad, i, d * gs);
  }
  acc = block_sum(acc, red);
  finish_scalar(acc, sc, loss, 1.f / (float)n, red);
}

template <typename T>
__global__ void __launch_bounds__(256)
sum_kernel(const T* __restrict__ x, T* __restrict__ grad, float* loss, long long n, float gscale,
           LossScratch* sc) {
  __shared__ float red[32];
  float acc = 0.f;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    acc += ldf(x, i);
    if (grad != nullptr) stf(grad, i, gscale);
  }
  acc = block_sum(acc, red);
  finish_scalar(acc, sc, loss, 1.f, red);
}

// One warp per row. kIndex: labels are class indices, else probability rows.
template <typename T, bool kIndex>
__global__ void __launch_bounds__(256)
ce_kernel(const T* __restrict__ logits, const float* __restrict__ probs,
          const long long* __restrict__ labels, T* __restrict__ grad, float* loss, int B, int C,
          float gscale, LossScratch* sc) {
  __shared__ float red[32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
  float acc = 0.f;
  const float gs = gscale / (float)B;
  for (int row = blockIdx.x * wpb + warp; row < B; row += gridDim.x * wpb) {
    const T* x = logits + (long long)row * C;
    float m = -INFINITY;
    for (int c = lane; c < C; c += 32) m = fmaxf(m, ldf(x, c));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    float se = 0.f, tsum = 0.f, tx = 0.f;
    const long long lab = kIndex ? labels[row] : 0;
    for (int c = lane; c < C; c += 32) {
      const float v = ldf(x, c);
      se += expf(v - m);
      if (kIndex) {
        if (c == lab) { tsum = 1.f; tx = v; }
      } else {
        const float t = probs[(long long)row * C + c];
        tsum += t;
        tx += t * v;
      }
    }
    se = warp_sum(se);
    tsum = warp_sum(tsum);
    tx = warp_sum(tx);
    const float lse = m + logf(se);
    // loss_row = sum_c t_c * (lse - x_c), accumulated term by term: with a single class the term
    // is t * 0, so the reference's degenerate case (ddp_gpus.py:37) is exactly zero.
    const float inv = 1.f / se;
    float row_loss = 0.f;
    for (int c = lane; c < C; c += 32) {
      const float v = ldf(x, c);
      float t;
      if (kIndex) t = (c == lab) ? 1.f : 0.f;
      else t = probs[(long long)row * C + c];
      row_loss += t * (lse - v);
      if (grad != nullptr) {
        const float sm = expf(v - m) * inv;
        stf(grad + (long long)row * C, c, (sm * tsum - t) * gs);
      }
    }
    acc += row_loss;
  }
  acc = block_sum(acc, red);
  finish_scalar(acc, sc, loss, 1.f / (float)B, red);
}

static int loss_grid(long long work_items) {
  long long b = (work_items + 255) / 256;
  if (b > 148 * 4) b = 148 * 4;
  if (b > kLossMaxBlocks) b = kLossMaxBlocks;
  if (b < 1) b = 1;
  return (int)b;
}

}  // namespace b200

using namespace b200;

extern "C" long long b200dt_loss_scratch_bytes(void) { return (long long)sizeof(LossScratch); }

extern "C" int b200dt_mse_loss(const void* pred, const void* target, void* grad, float* loss,
                               long long n, float grad_scale, int is_bf16, void* scratch,
                               void* stream) {
  if (n <= 0 || scratch == nullptr) return -3;
  const int g = loss_grid(n);
  cudaStream_t st = (cudaStream_t)stream;
  if (is_bf16)
    mse_kernel<__nv_bfloat16><<<g, 256, 0, st>>>((const __nv_bfloat16*)pred,
                                                 (const __nv_bfloat16*)target, (__nv_bfloat16*)grad,
                                                 loss, n, grad_scale, (LossScratch*)scratch);
  else
    mse_kernel<float><<<g, 256, 0, st>>>((const float*)pred, (const float*)target, (float*)grad,
                                         loss, n, grad_scale, (LossScratch*)scratch);
  return (int)cudaGetLastError();
}

extern "C" int b200dt_sum_loss(const void* x, void* grad, float* loss, long long n,
                               float grad_scale, int is_bf16, void* scratch, void* stream) {
  if (n <= 0 || scratch == nullptr) return -3;
  const int g = loss_grid(n);
  cudaStream_t st = (cudaStream_t)stream;
  if (is_bf16)
    sum_kernel<__nv_bfloat16><<<g, 256, 0, st>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)grad,
                                                 loss, n, grad_scale, (LossScratch*)scratch);
  else
    sum_kernel<float><<<g, 256, 0, st>>>((const float*)x, (float*)grad, loss, n, grad_scale,
                                         (LossScratch*)scratch);
  return (int)cudaGetLastError();
}

extern "C" int b200dt_ce_prob_loss(const void* logits, const float* probs, void* grad, float* loss,
                                   int B, int C, float grad_scale, int is_bf16, void* scratch,
                                   void* stream) {
  if (B <= 0 || C <= 0 || scratch == nullptr) return -3;
  const int g = loss_grid((long long)B * 32);
  cudaStream_t st = (cudaStream_t)stream;
  if (is_bf16)
    ce_kernel<__nv_bfloat16, false><<<g, 256, 0, st>>>((const __nv_bfloat16*)logits, probs, nullptr,
                                                       (__nv_bfloat16*)grad, loss, B, C, grad_scale,
                                                       (LossScratch*)scratch);
  else
    ce_kernel<float, false><<<g, 256, 0, st>>>((const float*)logits, probs, nullptr, (float*)grad,
                                               loss, B, C, grad_scale, (LossScratch*)scratch);
  return (int)cudaGetLastError();
}

extern "C" int b200dt_ce_index_loss(const void* logits, const long long* labels, void* grad,
                                    float* loss, int B, int C, float grad_scale, int is_bf16,
                                    void* scratch, void* stream) {
  if (B <= 0 || C <= 0 || scratch == nullptr) return -3;
  const int g = loss_grid((long long)B * 32);
  cudaStream_t st = (cudaStream_t)stream;
  if (is_bf16)
    ce_kernel<__nv_bfloat16, true><<<g, 256, 0, st>>>((const __nv_bfloat16*)logits, nullptr, labels,
                                                      (__nv_bfloat16*)grad, loss, B, C, grad_scale,
                                                      (LossScratch*)scratch);
  else
    ce_kernel<float, true><<<g, 256, 0, st>>>((const float*)logits, nullptr, labels, (float*)grad,
                                              loss, B, C, grad_scale, (LossScratch*)scratch);
  return (int)cudaGetLastError();
}

// ------------------------------------------------------------------------------------------
// MSE loss for callers that must not allocate.
//
// Some users of this library run where cudaMalloc is off limits: a training loop whose memory
// was carved up once at start-up and handed out by its own arena, a plugin living inside
// another runtime that owns the device heap, or code running while allocation is forbidden
// (for example inside a region that has been frozen for replay).  b200dt_mse_loss() needs a
// scratch buffer from its caller; the variant below keeps that scratch inside the library image
// instead, in a fixed array in device global memory.  Every GPU that loads the library pays for
// the array (B200DT_MSE_STATIC_SCRATCH_BYTES, about 1 KB) until the process exits, which is why it
// is kept small and why the entry point only accepts up to B200DT_MSE_STATIC_MAX_ELEMS elements:
// one partial-sum slot per 4096-element tile, 256 slots.
namespace b200 {

constexpr int kStaticTile = 4096;                 // elements reduced by one CTA
constexpr int kStaticSlots = 256;                 // partial sums held in the static scratch
struct StaticLossScratch {
  float partial[kStaticSlots];
  unsigned int ticket;
  unsigned int pad[3];
};
static_assert(sizeof(StaticLossScratch) == B200DT_MSE_STATIC_SCRATCH_BYTES, "header and kernel disagree");
static_assert((long long)kStaticTile * kStaticSlots == B200DT_MSE_STATIC_MAX_ELEMS, "header and kernel disagree");

// The scratch itself.  In host code this name is only a handle for the runtime's symbol table,
// not an address the GPU understands; the address is looked up per device further down.
static __device__ __align__(256) unsigned char g_mse_static_scratch[sizeof(StaticLossScratch)];

template <typename T>
__global__ void __launch_bounds__(256)
mse_static_kernel(const T* __restrict__ pred, const T* __restrict__ target, T* __restrict__ grad,
                  float* loss, long long n, float gscale, StaticLossScratch* sc) {
  __shared__ float red[32];
  __shared__ bool is_last;
  float acc = 0.f;
  const float gs = 2.f * gscale / (float)n;
  const long long begin = (long long)blockIdx.x * kStaticTile;
  const long long end = begin + kStaticTile < n ? begin + kStaticTile : n;
  for (long long i = begin + threadIdx.x; i < end; i += blockDim.x) {
    const float d = ldf(pred, i) - ldf(target, i);
    acc += d * d;
    if (grad != nullptr) stf(grad, i, d * gs);
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) {
    sc->partial[blockIdx.x] = acc;
    __threadfence();
    is_last = (atomicAdd(&sc->ticket, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (is_last) {
    __threadfence();
    float s = 0.f;
    for (int i = threadIdx.x; i < (int)gridDim.x; i += blockDim.x)
      s += *reinterpret_cast<volatile float*>(&sc->partial[i]);
    s = block_sum(s, red);
    if (threadIdx.x == 0) *loss = s / (float)n;
  }
}

constexpr int kMaxDevices = 64;
static void* g_mse_static_ptr[kMaxDevices] = {nullptr};

// Device address of the scratch on the CURRENT device, resolved once per device.
static int mse_static_scratch_for_current_device(StaticLossScratch** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= kMaxDevices) return -8;
  void* p = __atomic_load_n(&g_mse_static_ptr[dev], __ATOMIC_ACQUIRE);
  if (p == nullptr) {
    e = cudaGetSymbolAddress(&p, g_mse_static_scratch);
    if (e != cudaSuccess) return (int)e;
    if (p == nullptr) return -8;
    // two threads racing here resolve the same address; either store is fine
    __atomic_store_n(&g_mse_static_ptr[dev], p, __ATOMIC_RELEASE);
  }
  *out = reinterpret_cast<StaticLossScratch*>(p);
  return 0;
}

}  // namespace b200

extern "C" int b200dt_mse_loss_static(const void* pred, const void* target, void* grad, float* loss,
                                      long long n, float grad_scale, int is_bf16, void* stream) {
  if (n <= 0 || pred == nullptr || target == nullptr || loss == nullptr) return -3;
  if (n > B200DT_MSE_STATIC_MAX_ELEMS) return -9;      // more tiles than partial-sum slots
  StaticLossScratch* sc = nullptr;
  int rc = mse_static_scratch_for_current_device(&sc);
  if (rc != 0) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  // The array still holds whatever the previous call left; only the ticket has to start at zero
  // (every partial slot that is read is written first by this launch).
  cudaError_t e = cudaMemsetAsync(&sc->ticket, 0, sizeof(unsigned int), st);
  if (e != cudaSuccess) return (int)e;
  const int grid = (int)((n + kStaticTile - 1) / kStaticTile);
  if (is_bf16)
    mse_static_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(
        (const __nv_bfloat16*)pred, (const __nv_bfloat16*)target, (__nv_bfloat16*)grad, loss, n,
        grad_scale, sc);
  else
    mse_static_kernel<float><<<grid, 256, 0, st>>>((const float*)pred, (const float*)target,
                                                   (float*)grad, loss, n, grad_scale, sc);
  return (int)cudaGetLastError();
}
