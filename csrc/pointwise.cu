// Fused pointwise backward of `y = relu(x W^T + b)`: g = dy * (y > 0) and db = column sums of g in
// ONE pass (torch spends four kernels on it: compare, multiply, fp32 cast, sum).  Replaces the ATen
// pointwise/reduction kernels behind nn.Linear + nn.ReLU backward of the reference's MLP-style models
// (/root/reference/03.model_parallel.ipynb:440-450, 01.data_parallel.ipynb:171).
#include "common.cuh"
#include "ptx.cuh"
#include "../include/b200dt.h"

namespace b200 {

constexpr int kPwThreads = 256;
constexpr int kPwUnroll = 4;

// Thread (tx, ty) owns the 8 channels of 16-byte column `blockIdx.x * cols + tx` and walks the rows
// ty, ty + lanes, ... of its row slab; partial column sums are combined through shared memory and
// one atomicAdd per (column, CTA) lands in db.
__global__ void __launch_bounds__(kPwThreads)
relu_bwd_bias_kernel(const uint4* __restrict__ dy, const uint4* __restrict__ y,
                     uint4* __restrict__ g_out, float* __restrict__ db, long long rows, int C,
                     int cols, int lanes) {
  __shared__ float red[kPwThreads][9];
  const int cv = C / 8;
  const int tx = threadIdx.x % cols, ty = threadIdx.x / cols;
  const int col = blockIdx.x * cols + tx;
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  if (col < cv && ty < lanes) {
    const long long stride = (long long)gridDim.y * lanes;
    for (long long r = (long long)blockIdx.y * lanes + ty; r < rows; r += kPwUnroll * stride) {
      uint4 vg[kPwUnroll], vy[kPwUnroll];
#pragma unroll
      for (int u = 0; u < kPwUnroll; ++u) {
        const long long rr = r + u * stride;
        if (rr < rows) {
          vg[u] = ld_nc_v4(dy + rr * cv + col);
          if (y != nullptr) vy[u] = ld_nc_v4(y + rr * cv + col);
        }
      }
#pragma unroll
      for (int u = 0; u < kPwUnroll; ++u) {
        const long long rr = r + u * stride;
        if (rr >= rows) break;
        float g[8];
        unpack8(vg[u], g);
        if (y != nullptr) {
          float yv[8];
          unpack8(vy[u], yv);
#pragma unroll
          for (int j = 0; j < 8; ++j) g[j] = yv[j] > 0.f ? g[j] : 0.f;
          g_out[rr * cv + col] = pack8(g);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += g[j];
      }
    }
  }
  if (db == nullptr) return;
#pragma unroll
  for (int j = 0; j < 8; ++j) red[threadIdx.x][j] = acc[j];
  __syncthreads();
  if (ty == 0 && col < cv) {
    for (int l = 1; l < lanes; ++l)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += red[l * cols + tx][j];
#pragma unroll
    for (int j = 0; j < 8; ++j) atomicAdd(db + col * 8 + j, acc[j]);
  }
}

}  // namespace b200

using namespace b200;

extern "C" int b200dt_relu_bwd_bias(const void* dy, const void* y, void* g, float* db, long long rows,
                                    int C, void* stream) {
  if (rows <= 0 || C <= 0) return 0;
  if (C % 8) return -3;
  if ((y == nullptr) != (g == nullptr)) return -3;
  if (y == nullptr && db == nullptr) return 0;
  cudaStream_t st = (cudaStream_t)stream;
  if (db != nullptr) {
    cudaError_t e = cudaMemsetAsync(db, 0, sizeof(float) * (size_t)C, st);
    if (e != cudaSuccess) return (int)e;
  }
  const int cv = C / 8;
  const int cols = cv < 32 ? cv : 32;
  const int lanes = kPwThreads / cols;
  const int gx = (cv + cols - 1) / cols;
  long long gy = (148LL * 8 + gx - 1) / gx;
  const long long max_gy = (rows + (long long)lanes * kPwUnroll - 1) / ((long long)lanes * kPwUnroll);
  if (gy > max_gy) gy = max_gy;
  if (gy < 1 || (db != nullptr && deterministic())) gy = 1;   // det: one CTA per column group, one add per column
  relu_bwd_bias_kernel<<<dim3(gx, (unsigned)gy), kPwThreads, 0, st>>>(
      reinterpret_cast<const uint4*>(dy), reinterpret_cast<const uint4*>(y),
      reinterpret_cast<uint4*>(g), db, rows, C, cols, lanes);
  return (int)cudaGetLastError();
}
