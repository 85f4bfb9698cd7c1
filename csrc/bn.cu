// BatchNorm (training) for NHWC bf16 activations, fused with ReLU and the residual add.
// Replaces the cuDNN BatchNorm / ATen relu / add chain of torchvision's Bottleneck that the
// reference runs in /root/reference/03.model_parallel.ipynb:807-833 (tv:models/resnet.py:108-163).
//
// Every kernel streams a [rows, C] matrix with 16-byte (8-channel) accesses.  A thread owns one
// 8-channel group for its whole life, so per-channel constants live in registers; column sums are
// kept in registers across a strip of rows, folded through shared memory and finished with one
// fp32 atomic per channel per CTA.
#include "common.cuh"
#include "ptx.cuh"
#include "../include/b200dt.h"

namespace b200 {

constexpr int kBnThreads = 256;
constexpr int kUnroll = 4;
constexpr int kUnrollBwd = 4;
// The partial-sum array behind `sums` has b200::bn_replicas_used(C) copies: up to 16 (spreads the atomics
// of ~900 CTAs; fewer for wide layers), one per row-slab CTA in deterministic mode (plain stores, folded in
// slab order), then one ticket word.

struct Slab {
  int cols_per_block;   // vector (8-channel) columns handled by one CTA
  int rows_per_iter;    // rows covered by one CTA iteration
  dim3 grid;
};

static Slab make_slab(long long rows, int C, int target_blocks) {
  Slab s;
  const int cv = C / 8;
  s.cols_per_block = cv < 32 ? cv : 32;
  // cols_per_block must divide 256 for the thread mapping below: C/8 in {8,16,32,...} does.
  while (kBnThreads % s.cols_per_block) --s.cols_per_block;
  s.rows_per_iter = kBnThreads / s.cols_per_block;
  const int gx = (cv + s.cols_per_block - 1) / s.cols_per_block;
  long long gy = target_blocks / gx;
  const long long max_gy = (rows + s.rows_per_iter - 1) / s.rows_per_iter;
  if (gy > max_gy) gy = max_gy;
  if (deterministic() && gy > bn_replicas()) gy = bn_replicas();     // one partial-sum copy per row slab
  if (gy < 1) gy = 1;
  s.grid = dim3(gx, (unsigned)gy, 1);
  return s;
}

// Fold the per-thread accumulators acc[NACC][8] of a CTA and add the CTA totals to
// out[a * C + channel].  Lanes of a warp that own the same channel group are combined with
// shuffles, the eight warps through shared memory, and the final sums / atomics are spread over
// all threads (no serial tail on a few threads).
template <int NACC>
__device__ __forceinline__ void fold_and_add(float (&acc)[NACC][8], int tx, int cols, int C,
                                             float* smem, float* __restrict__ out, int reps, int det) {
  constexpr int NV = NACC * 8;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nwarps = kBnThreads / 32;
  for (int off = cols; off < 32; off <<= 1)
#pragma unroll
    for (int a = 0; a < NACC; ++a)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[a][j] += __shfl_xor_sync(0xffffffffu, acc[a][j], off);
  // with cols == 32 every warp is one row-lane; otherwise lanes < cols hold the warp partials
  if (lane < cols) {
#pragma unroll
    for (int a = 0; a < NACC; ++a)
#pragma unroll
      for (int j = 0; j < 8; ++j) smem[(warp * NV + a * 8 + j) * cols + tx] = acc[a][j];
  }
  __syncthreads();
  const int cv = C / 8;
  for (int t = threadIdx.x; t < NV * cols; t += kBnThreads) {
    const int v = t / cols, x = t - v * cols;
    const int col = blockIdx.x * cols + x;
    if (col >= cv) continue;
    float s = 0.f;
    for (int w = 0; w < nwarps; ++w) s += smem[(w * NV + v) * cols + x];
    if (det) out[(size_t)blockIdx.y * 2 * C + (v >> 3) * C + col * 8 + (v & 7)] = s;   // slab-owned copy
    else atomicAdd(out + (size_t)(blockIdx.y % reps) * 2 * C + (v >> 3) * C + col * 8 + (v & 7), s);
  }
  // The last CTA to finish folds the replicas into copy 0, so consumers read one plain array.
  // `reps` = copies in use (bn_replicas_used(C)); the ticket word sits right behind them.
  if (!det && reps == 1) return;                  // a single copy IS the result
  __shared__ bool is_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int* ticket = reinterpret_cast<unsigned int*>(out + (size_t)reps * 2 * C);
    is_last = (atomicAdd(ticket, 1u) == gridDim.x * gridDim.y - 1);
  }
  __syncthreads();
  if (is_last) {
    __threadfence();
    fold_sum_copies(out, 2 * C, det ? (int)gridDim.y : reps, threadIdx.x, kBnThreads);
  }
}

__global__ void __launch_bounds__(kBnThreads)
bn_stats_kernel(const uint4* __restrict__ x, float* __restrict__ sums, long long rows, int C,
                int cols, int rowlanes, int reps, int det) {
  extern __shared__ float smem[];
  griddep_launch_dependents();
  griddep_wait();
  const int cv = C / 8;
  const int tx = threadIdx.x % cols, ty = threadIdx.x / cols;
  const int col = blockIdx.x * cols + tx;
  float acc[2][8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[0][j] = acc[1][j] = 0.f;
  if (col < cv) {
    const long long stride = (long long)gridDim.y * rowlanes;
    long long r = (long long)blockIdx.y * rowlanes + ty;
    for (; r + (kUnroll - 1) * stride < rows; r += kUnroll * stride) {
      uint4 v[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) v[u] = ld_nc_v4(x + (r + u * stride) * cv + col);
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        float f[8];
        unpack8(v[u], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) { acc[0][j] += f[j]; acc[1][j] += f[j] * f[j]; }
      }
    }
    for (; r < rows; r += stride) {
      float f[8];
      unpack8(ld_nc_v4(x + r * cv + col), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) { acc[0][j] += f[j]; acc[1][j] += f[j] * f[j]; }
    }
  }
  fold_and_add<2>(acc, tx, cols, C, smem, sums, reps, det);
}

template <bool kTrain>
__global__ void __launch_bounds__(kBnThreads)
bn_apply_kernel(const uint4* __restrict__ x, const uint4* __restrict__ res, uint4* __restrict__ y,
                const float* __restrict__ sums, const float* __restrict__ gamma,
                const float* __restrict__ beta, float* running_mean, float* running_var,
                float* save_mean, float* save_invstd, const float* __restrict__ in_mean,
                const float* __restrict__ in_var, long long rows, int C, float eps, float momentum,
                int relu, int cols, int rowlanes, long long* batches_tracked,
                unsigned char* __restrict__ relu_mask) {
  griddep_launch_dependents();
  griddep_wait();
  const int cv = C / 8;
  const int tx = threadIdx.x % cols, ty = threadIdx.x / cols;
  const int col = blockIdx.x * cols + tx;
  if (kTrain && batches_tracked != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0)
    *batches_tracked += 1;   // nn.BatchNorm's num_batches_tracked, without a launch of its own
  if (col >= cv) return;
  float scale[8], shift[8];
  const float inv_rows = 1.f / (float)rows;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = col * 8 + j;
    float mean, var;
    if (kTrain) {
      mean = sums[c] * inv_rows;
      var = fmaxf(sums[C + c] * inv_rows - mean * mean, 0.f);
    } else {
      mean = in_mean[c];
      var = in_var[c];
    }
    const float invstd = rsqrtf(var + eps);
    scale[j] = gamma[c] * invstd;
    shift[j] = beta[c] - mean * scale[j];
    if (kTrain && blockIdx.y == 0 && ty == 0) {
      save_mean[c] = mean;
      save_invstd[c] = invstd;
      if (running_mean != nullptr) {
        const float unbiased = rows > 1 ? var * ((float)rows / (float)(rows - 1)) : var;
        running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
        running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
      }
    }
  }
  const long long stride = (long long)gridDim.y * rowlanes;
  long long r = (long long)blockIdx.y * rowlanes + ty;
  for (; r < rows; r += kUnroll * stride) {
    uint4 vx[kUnroll], vr[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const long long rr = r + u * stride;
      if (rr < rows) {
        vx[u] = ld_nc_v4(x + rr * cv + col);
        if (res != nullptr) vr[u] = ld_nc_v4(res + rr * cv + col);
      }
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const long long rr = r + u * stride;
      if (rr >= rows) break;
      float f[8];
      unpack8(vx[u], f);
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] = f[j] * scale[j] + shift[j];
      if (res != nullptr) {
        float g[8];
        unpack8(vr[u], g);
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] += g[j];
      }
      if (relu) {
        if (relu_mask != nullptr) {
          // one bit per element: backward masks the gradient from this byte instead of re-reading y
          unsigned m = 0;
#pragma unroll
          for (int j = 0; j < 8; ++j) m |= (f[j] > 0.f ? 1u : 0u) << j;
          relu_mask[rr * cv + col] = (unsigned char)m;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = fmaxf(f[j], 0.f);
      }
      y[rr * cv + col] = pack8(f);
    }
  }
}

// kMask: 0 = no ReLU, 1 = mask from the saved output y, 2 = mask recomputed from x,
// 3 = mask from the forward's bit mask (y points at one byte per 8 channels)
template <int kMask>
__global__ void __launch_bounds__(kBnThreads)
bn_bwd_reduce_kernel(const uint4* __restrict__ dy, const uint4* __restrict__ y,
                     const uint4* __restrict__ x, const float* __restrict__ save_mean,
                     const float* __restrict__ save_invstd, const float* __restrict__ gamma,
                     const float* __restrict__ beta, float* __restrict__ sums,
                     uint4* __restrict__ dres, long long rows, int C, int relu, int cols,
                     int rowlanes, int reps, int det) {
  extern __shared__ float smem[];
  griddep_launch_dependents();
  griddep_wait();
  const int cv = C / 8;
  const int tx = threadIdx.x % cols, ty = threadIdx.x / cols;
  const int col = blockIdx.x * cols + tx;
  float acc[2][8];
  float mean[8], invstd[8], fscale[8], fshift[8];
  // y == nullptr with relu: the ReLU mask is recomputed from x with the forward's arithmetic
  // (y = x * scale + shift), which saves reading y (only valid without a residual input)
  constexpr bool remask = (kMask == 2);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    acc[0][j] = acc[1][j] = 0.f;
    const int c = (col < cv ? col : 0) * 8 + j;
    mean[j] = save_mean[c];
    invstd[j] = save_invstd[c];
    fscale[j] = remask ? gamma[c] * invstd[j] : 0.f;
    fshift[j] = remask ? beta[c] - mean[j] * fscale[j] : 0.f;
  }
  if (col < cv) {
    const long long stride = (long long)gridDim.y * rowlanes;
    for (long long r = (long long)blockIdx.y * rowlanes + ty; r < rows; r += kUnrollBwd * stride) {
      uint4 vg[kUnrollBwd], vx[kUnrollBwd], vy[kUnrollBwd];
#pragma unroll
      for (int u = 0; u < kUnrollBwd; ++u) {
        const long long rr = r + u * stride;
        if (rr < rows) {
          vg[u] = ld_nc_v4(dy + rr * cv + col);
          vx[u] = ld_nc_v4(x + rr * cv + col);
          if (kMask == 1) vy[u] = ld_nc_v4(y + rr * cv + col);
          if (kMask == 3) vy[u].x = reinterpret_cast<const unsigned char*>(y)[rr * cv + col];
        }
      }
#pragma unroll
      for (int u = 0; u < kUnrollBwd; ++u) {
        const long long rr = r + u * stride;
        if (rr >= rows) break;
        float g[8], xv[8];
        unpack8(vg[u], g);
        unpack8(vx[u], xv);
        if (remask) {
#pragma unroll
          for (int j = 0; j < 8; ++j) g[j] = (xv[j] * fscale[j] + fshift[j]) > 0.f ? g[j] : 0.f;
        } else if (kMask == 1) {
          float yv[8];
          unpack8(vy[u], yv);
#pragma unroll
          for (int j = 0; j < 8; ++j) g[j] = yv[j] > 0.f ? g[j] : 0.f;
        } else if (kMask == 3) {
#pragma unroll
          for (int j = 0; j < 8; ++j) g[j] = ((vy[u].x >> j) & 1u) ? g[j] : 0.f;
        }
        if (dres != nullptr) dres[rr * cv + col] = pack8(g);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          acc[0][j] += g[j];
          acc[1][j] += g[j] * (xv[j] - mean[j]) * invstd[j];
        }
      }
    }
  }
  fold_and_add<2>(acc, tx, cols, C, smem, sums, reps, det);
}

template <int kMask>
__global__ void __launch_bounds__(kBnThreads)
bn_bwd_apply_kernel(const uint4* __restrict__ dy, const uint4* __restrict__ y,
                    const uint4* __restrict__ x, const float* __restrict__ save_mean,
                    const float* __restrict__ save_invstd, const float* __restrict__ gamma,
                    const float* __restrict__ beta, const float* __restrict__ sums,
                    uint4* __restrict__ dx, float* dgamma, float* dbeta, long long rows, int C,
                    int relu, int cols, int rowlanes) {
  griddep_launch_dependents();
  griddep_wait();
  const int cv = C / 8;
  const int tx = threadIdx.x % cols, ty = threadIdx.x / cols;
  const int col = blockIdx.x * cols + tx;
  if (col >= cv) return;
  float mean[8], invstd[8], scale[8], mg[8], mgx[8], fshift[8];
  const float inv_rows = 1.f / (float)rows;
  constexpr bool remask = (kMask == 2);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = col * 8 + j;
    mean[j] = save_mean[c];
    invstd[j] = save_invstd[c];
    scale[j] = gamma[c] * invstd[j];
    fshift[j] = remask ? beta[c] - mean[j] * scale[j] : 0.f;
    const float sg = sums[c], sgx = sums[C + c];
    mg[j] = sg * inv_rows;
    mgx[j] = sgx * inv_rows;
    if (blockIdx.y == 0 && ty == 0) {
      if (dbeta != nullptr) dbeta[c] = sg;
      if (dgamma != nullptr) dgamma[c] = sgx;
    }
  }
  const long long stride = (long long)gridDim.y * rowlanes;
  for (long long r = (long long)blockIdx.y * rowlanes + ty; r < rows; r += kUnrollBwd * stride) {
    uint4 vg[kUnrollBwd], vx[kUnrollBwd], vy[kUnrollBwd];
#pragma unroll
    for (int u = 0; u < kUnrollBwd; ++u) {
      const long long rr = r + u * stride;
      if (rr < rows) {
        vg[u] = ld_nc_v4(dy + rr * cv + col);
        vx[u] = ld_nc_v4(x + rr * cv + col);
        if (kMask == 1) vy[u] = ld_nc_v4(y + rr * cv + col);
      }
    }
#pragma unroll
    for (int u = 0; u < kUnrollBwd; ++u) {
      const long long rr = r + u * stride;
      if (rr >= rows) break;
      float g[8], xv[8];
      unpack8(vg[u], g);
      unpack8(vx[u], xv);
      if (remask) {
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] = (xv[j] * scale[j] + fshift[j]) > 0.f ? g[j] : 0.f;
      } else if (kMask == 1) {
        float yv[8];
        unpack8(vy[u], yv);
#pragma unroll
        for (int j = 0; j < 8; ++j) g[j] = yv[j] > 0.f ? g[j] : 0.f;
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float xhat = (xv[j] - mean[j]) * invstd[j];
        g[j] = scale[j] * (g[j] - mg[j] - xhat * mgx[j]);
      }
      dx[rr * cv + col] = pack8(g);
    }
  }
}

static int target_blocks() { return 148 * 6; }

// Zero what the reduction adds into: the copies in use and the ticket word behind them -- unless the caller
// has promised zeroed buffers (b200dt_set_sums_prezeroed; the Python layer carves them from zero-filled chunks,
// which removes ~160 memset nodes from a ResNet-50 step).
static cudaError_t clear_sums(float* sums, int C, int used, cudaStream_t st) {
  if (sums_prezeroed()) return cudaSuccess;
  return cudaMemsetAsync(sums, 0, sizeof(float) * ((size_t)2 * C * used + 1), st);
}

}  // namespace b200

using namespace b200;

extern "C" int b200dt_bn_stats(const void* x, float* sums, long long rows, int C, void* stream) {
  if (C % 8 || rows <= 0) return -3;
  cudaStream_t st = (cudaStream_t)stream;
  const int used = bn_replicas_used(C);
  cudaError_t e = clear_sums(sums, C, used, st);
  if (e != cudaSuccess) return (int)e;
  Slab s = make_slab(rows, C, target_blocks());
  const size_t smem = sizeof(float) * kBnThreads * 16;
  e = launch_pdl(kPdlBn, bn_stats_kernel, s.grid, dim3(kBnThreads), smem, st, reinterpret_cast<const uint4*>(x), sums,
                 rows, C, s.cols_per_block, s.rows_per_iter, used, deterministic());
  return e != cudaSuccess ? (int)e : (int)cudaGetLastError();
}

extern "C" int b200dt_bn_apply(const void* x, const void* residual, void* y, const float* sums,
                               const float* gamma, const float* beta, float* running_mean,
                               float* running_var, float* save_mean, float* save_invstd,
                               long long rows, int C, float eps, float momentum, int relu,
                               long long* batches_tracked, unsigned char* relu_mask, void* stream) {
  if (C % 8 || rows <= 0) return -3;
  Slab s = make_slab(rows, C, target_blocks());
  cudaError_t e = launch_pdl(
      kPdlBn, bn_apply_kernel<true>, s.grid, dim3(kBnThreads), 0, (cudaStream_t)stream,
      reinterpret_cast<const uint4*>(x), reinterpret_cast<const uint4*>(residual),
      reinterpret_cast<uint4*>(y), sums, gamma, beta, running_mean, running_var, save_mean,
      save_invstd, nullptr, nullptr, rows, C, eps, momentum, relu, s.cols_per_block,
      s.rows_per_iter, batches_tracked, relu ? relu_mask : nullptr);
  return e != cudaSuccess ? (int)e : (int)cudaGetLastError();
}

extern "C" int b200dt_bn_apply_eval(const void* x, const void* residual, void* y,
                                    const float* gamma, const float* beta, const float* mean,
                                    const float* var, long long rows, int C, float eps, int relu,
                                    void* stream) {
  if (C % 8 || rows <= 0) return -3;
  Slab s = make_slab(rows, C, target_blocks());
  cudaError_t e = launch_pdl(
      kPdlBn, bn_apply_kernel<false>, s.grid, dim3(kBnThreads), 0, (cudaStream_t)stream,
      reinterpret_cast<const uint4*>(x), reinterpret_cast<const uint4*>(residual),
      reinterpret_cast<uint4*>(y), nullptr, gamma, beta, nullptr, nullptr, nullptr, nullptr, mean,
      var, rows, C, eps, 0.f, relu, s.cols_per_block, s.rows_per_iter, nullptr, nullptr);
  return e != cudaSuccess ? (int)e : (int)cudaGetLastError();
}

extern "C" int b200dt_bn_bwd_reduce(const void* dy, const void* y, const void* x,
                                    const float* save_mean, const float* save_invstd,
                                    const float* gamma, const float* beta, float* sums, void* dres,
                                    long long rows, int C, int relu, void* stream) {
  if (C % 8 || rows <= 0) return -3;
  if (relu && y == nullptr && (gamma == nullptr || beta == nullptr)) return -3;
  if (relu == 2 && (y == nullptr || dres == nullptr)) return -3;   // bit-mask mode materialises the masked gradient
  cudaStream_t st = (cudaStream_t)stream;
  const int used = bn_replicas_used(C);
  cudaError_t e = clear_sums(sums, C, used, st);
  if (e != cudaSuccess) return (int)e;
  Slab s = make_slab(rows, C, target_blocks());
  const size_t smem = sizeof(float) * kBnThreads * 16;
  const int mask = !relu ? 0 : (relu == 2 ? 3 : (y != nullptr ? 1 : 2));
#define B200_LAUNCH_REDUCE(M)                                                                      \
  e = launch_pdl(kPdlBn, bn_bwd_reduce_kernel<M>, s.grid, dim3(kBnThreads), smem, st,                      \
      reinterpret_cast<const uint4*>(dy), reinterpret_cast<const uint4*>(y),                       \
      reinterpret_cast<const uint4*>(x), save_mean, save_invstd, gamma, beta, sums,                \
      reinterpret_cast<uint4*>(dres), rows, C, relu, s.cols_per_block, s.rows_per_iter, used,      \
      deterministic())
  if (mask == 0) B200_LAUNCH_REDUCE(0); else if (mask == 1) B200_LAUNCH_REDUCE(1);
  else if (mask == 2) B200_LAUNCH_REDUCE(2); else B200_LAUNCH_REDUCE(3);
#undef B200_LAUNCH_REDUCE
  return e != cudaSuccess ? (int)e : (int)cudaGetLastError();
}

extern "C" int b200dt_bn_bwd_apply(const void* dy, const void* y, const void* x,
                                   const float* save_mean, const float* save_invstd,
                                   const float* gamma, const float* beta, const float* sums,
                                   void* dx, float* dgamma, float* dbeta, long long rows, int C,
                                   int relu, void* stream) {
  if (C % 8 || rows <= 0) return -3;
  if (relu && y == nullptr && beta == nullptr) return -3;
  Slab s = make_slab(rows, C, target_blocks());
  const int mask = !relu ? 0 : (y != nullptr ? 1 : 2);
#define B200_LAUNCH_APPLY(M)                                                                       \
  e = launch_pdl(kPdlBn, bn_bwd_apply_kernel<M>, s.grid, dim3(kBnThreads), 0, (cudaStream_t)stream,        \
      reinterpret_cast<const uint4*>(dy), reinterpret_cast<const uint4*>(y),                       \
      reinterpret_cast<const uint4*>(x), save_mean, save_invstd, gamma, beta, sums,                \
      reinterpret_cast<uint4*>(dx), dgamma, dbeta, rows, C, relu, s.cols_per_block, s.rows_per_iter)
  cudaError_t e = cudaSuccess;
  if (mask == 0) B200_LAUNCH_APPLY(0); else if (mask == 1) B200_LAUNCH_APPLY(1); else B200_LAUNCH_APPLY(2);
#undef B200_LAUNCH_APPLY
  return e != cudaSuccess ? (int)e : (int)cudaGetLastError();
}
