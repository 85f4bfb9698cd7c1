// fp32 SIMT Linear for the tutorial's degenerate layers: Linear(20,1) (ddp_gpus.py:81), Linear(32,2)
// (01.data_parallel.ipynb:171), Linear(10000,10) and Linear(10,5) (03.model_parallel.ipynb:444-446).
// Their feature counts are not multiples of 8 (no TMA row pitch) and a step produces 20..200 outputs:
// there is no tensor-core work in them, only launch latency, so each pass is ONE small kernel with one
// thread per output element (forward: per K-slice, combined with atomics) -- this replaces the cuBLAS
// GEMV / GEMM calls nn.Linear makes for them.
//
// Every kernel body is an `elem` function of a flat index that also compiles for the host; the
// `*_host` entry points run exactly that code in a serial loop, so the index arithmetic is unit-tested
// on machines without a GPU (tests/test_plumbing_cpu.py) against torch.
#include <cuda_runtime.h>
#include <stdint.h>
#include "../include/b200dt.h"

namespace b200 {
int deterministic();        // csrc/config.cu

#define B200_HD __host__ __device__ __forceinline__

B200_HD void atomic_add_f32(float* p, float v) {
#ifdef __CUDA_ARCH__
  atomicAdd(p, v);
#else
  *p += v;             // host emulation is serial
#endif
}

// Forward. idx enumerates (b, n, c): slice c of `slices` owns k = c, c + slices, c + 2*slices, ...
// (consecutive threads read consecutive k).  y must be zero on entry; slice 0 adds the bias.
B200_HD void skinny_fwd_elem(const float* x, const float* w, const float* bias, float* y, int B, int K,
                             int N, int slices, long long idx) {
  const int c = (int)(idx % slices);
  const long long bn = idx / slices;
  const int n = (int)(bn % N);
  const long long b = bn / N;
  if (b >= B) return;
  const float* xr = x + b * K;
  const float* wr = w + (long long)n * K;
  float acc = 0.f;
  for (int k = c; k < K; k += slices) acc += xr[k] * wr[k];
  if (c == 0 && bias != nullptr) acc += bias[n];
  atomic_add_f32(y + b * N + n, acc);
}

// Data gradient: dx[b,k] = sum_n dy[b,n] * w[n,k]; one thread per (b, k).
B200_HD void skinny_dgrad_elem(const float* dy, const float* w, float* dx, int B, int K, int N,
                               long long idx) {
  const int k = (int)(idx % K);
  const long long b = idx / K;
  if (b >= B) return;
  float acc = 0.f;
  for (int n = 0; n < N; ++n) acc += dy[b * N + n] * w[(long long)n * K + k];
  dx[b * K + k] = acc;
}

// Weight gradient: dw[n,k] = sum_b dy[b,n] * x[b,k]; one thread per (n, k, c) with the batch
// interleaved over `slices`; dw must be zero on entry.
B200_HD void skinny_wgrad_elem(const float* dy, const float* x, float* dw, int B, int K, int N,
                               int slices, long long idx) {
  const int c = (int)(idx % slices);
  const long long nk = idx / slices;
  const int k = (int)(nk % K);
  const long long n = nk / K;
  if (n >= N) return;
  float acc = 0.f;
  for (int b = c; b < B; b += slices) acc += dy[(long long)b * N + n] * x[(long long)b * K + k];
  atomic_add_f32(dw + n * K + k, acc);
}

// Bias gradient: db[n] = sum_b dy[b,n]; one thread per (n, c); db must be zero on entry.
B200_HD void skinny_bgrad_elem(const float* dy, float* db, int B, int N, int slices, long long idx) {
  const int c = (int)(idx % slices);
  const long long n = idx / slices;
  if (n >= N) return;
  float acc = 0.f;
  for (int b = c; b < B; b += slices) acc += dy[(long long)b * N + n];
  atomic_add_f32(db + n, acc);
}

__global__ void skinny_fwd_kernel(const float* x, const float* w, const float* bias, float* y, int B,
                                  int K, int N, int slices, long long total) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x)
    skinny_fwd_elem(x, w, bias, y, B, K, N, slices, i);
}
__global__ void skinny_dgrad_kernel(const float* dy, const float* w, float* dx, int B, int K, int N,
                                    long long total) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x)
    skinny_dgrad_elem(dy, w, dx, B, K, N, i);
}
__global__ void skinny_wgrad_kernel(const float* dy, const float* x, float* dw, int B, int K, int N,
                                    int slices, long long total) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x)
    skinny_wgrad_elem(dy, x, dw, B, K, N, slices, i);
}
__global__ void skinny_bgrad_kernel(const float* dy, float* db, int B, int N, int slices,
                                    long long total) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x)
    skinny_bgrad_elem(dy, db, B, N, slices, i);
}

// Number of reduction slices: enough threads to fill the machine for long reductions, one for short.
static int pick_slices(long long outputs, int reduce_len) {
  if (deterministic()) return 1;     // one thread per output: no atomics to order
  int s = 1;
  while (s < 256 && (long long)s * outputs < 148LL * 256 && s * 8 < reduce_len) s *= 2;
  return s;
}
static int grid_for_total(long long total) {
  long long b = (total + 255) / 256;
  if (b > 148 * 8) b = 148 * 8;
  if (b < 1) b = 1;
  return (int)b;
}

}  // namespace b200

using namespace b200;

static bool bad_dims(int B, int K, int N) { return B <= 0 || K <= 0 || N <= 0; }

extern "C" int b200dt_skinny_linear_fwd(const float* x, const float* w, const float* bias, float* y,
                                        int B, int K, int N, void* stream) {
  if (bad_dims(B, K, N)) return B == 0 ? 0 : -3;
  cudaStream_t st = (cudaStream_t)stream;
  cudaError_t e = cudaMemsetAsync(y, 0, sizeof(float) * (size_t)B * N, st);
  if (e != cudaSuccess) return (int)e;
  const int slices = pick_slices((long long)B * N, K);
  const long long total = (long long)B * N * slices;
  skinny_fwd_kernel<<<grid_for_total(total), 256, 0, st>>>(x, w, bias, y, B, K, N, slices, total);
  return (int)cudaGetLastError();
}

extern "C" int b200dt_skinny_linear_bwd(const float* dy, const float* x, const float* w, float* dx,
                                        float* dw, float* db, int B, int K, int N, void* stream) {
  if (bad_dims(B, K, N)) return B == 0 ? 0 : -3;
  cudaStream_t st = (cudaStream_t)stream;
  if (dx != nullptr) {
    const long long total = (long long)B * K;
    skinny_dgrad_kernel<<<grid_for_total(total), 256, 0, st>>>(dy, w, dx, B, K, N, total);
  }
  if (dw != nullptr) {
    cudaError_t e = cudaMemsetAsync(dw, 0, sizeof(float) * (size_t)N * K, st);
    if (e != cudaSuccess) return (int)e;
    const int slices = pick_slices((long long)N * K, B);
    const long long total = (long long)N * K * slices;
    skinny_wgrad_kernel<<<grid_for_total(total), 256, 0, st>>>(dy, x, dw, B, K, N, slices, total);
  }
  if (db != nullptr) {
    cudaError_t e = cudaMemsetAsync(db, 0, sizeof(float) * (size_t)N, st);
    if (e != cudaSuccess) return (int)e;
    const int slices = pick_slices(N, B);
    const long long total = (long long)N * slices;
    skinny_bgrad_kernel<<<grid_for_total(total), 256, 0, st>>>(dy, db, B, N, slices, total);
  }
  return (int)cudaGetLastError();
}

// ---- host emulation of the very same element functions (no GPU needed; see the header comment)
extern "C" int b200dt_skinny_linear_fwd_host(const float* x, const float* w, const float* bias,
                                             float* y, int B, int K, int N) {
  if (bad_dims(B, K, N)) return B == 0 ? 0 : -3;
  for (long long i = 0; i < (long long)B * N; ++i) y[i] = 0.f;
  const int slices = pick_slices((long long)B * N, K);
  const long long total = (long long)B * N * slices;
  for (long long i = 0; i < total; ++i) skinny_fwd_elem(x, w, bias, y, B, K, N, slices, i);
  return 0;
}

extern "C" int b200dt_skinny_linear_bwd_host(const float* dy, const float* x, const float* w,
                                             float* dx, float* dw, float* db, int B, int K, int N) {
  if (bad_dims(B, K, N)) return B == 0 ? 0 : -3;
  if (dx != nullptr)
    for (long long i = 0; i < (long long)B * K; ++i) skinny_dgrad_elem(dy, w, dx, B, K, N, i);
  if (dw != nullptr) {
    for (long long i = 0; i < (long long)N * K; ++i) dw[i] = 0.f;
    const int slices = pick_slices((long long)N * K, B);
    for (long long i = 0; i < (long long)N * K * slices; ++i)
      skinny_wgrad_elem(dy, x, dw, B, K, N, slices, i);
  }
  if (db != nullptr) {
    for (int i = 0; i < N; ++i) db[i] = 0.f;
    const int slices = pick_slices(N, B);
    for (long long i = 0; i < (long long)N * slices; ++i) skinny_bgrad_elem(dy, db, B, N, slices, i);
  }
  return 0;
}
