// NHWC bf16 pooling: MaxPool 3x3 stride 2 pad 1 (with a 1-byte arg-max so the backward needs no
// atomics) and global average pooling.  Replaces nn.MaxPool2d / nn.AdaptiveAvgPool2d of the
// ResNet-50 in /root/reference/03.model_parallel.ipynb:815,824.
#include "common.cuh"
#include "ptx.cuh"
#include "../include/b200dt.h"

namespace b200 {

__global__ void __launch_bounds__(256)
maxpool_fwd_kernel(const uint4* __restrict__ x, uint4* __restrict__ y, uint2* __restrict__ argmax,
                   int NB, int H, int W, int C) {
  const int Ho = H / 2, Wo = W / 2, cv = C / 8;
  const long long total = (long long)NB * Ho * Wo * cv;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % cv);
    long long pix = i / cv;
    const int wo = (int)(pix % Wo); pix /= Wo;
    const int ho = (int)(pix % Ho);
    const int n = (int)(pix / Ho);
    float best[8];
    uint32_t arg[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) { best[j] = -INFINITY; arg[j] = 0; }
    // all nine loads are issued before the first compare: out-of-range taps read a clamped (valid) address
    // and are discarded afterwards, so no load sits behind a branch
    uint4 v[9];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int h = min(max(ho * 2 - 1 + r, 0), H - 1);
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const int w = min(max(wo * 2 - 1 + s, 0), W - 1);
        v[r * 3 + s] = ld_nc_v4(x + (((long long)n * H + h) * W + w) * cv + c);
      }
    }
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int h = ho * 2 - 1 + r;
#pragma unroll
      for (int s = 0; s < 3; ++s) {
        const int w = wo * 2 - 1 + s;
        const bool inside = h >= 0 && h < H && w >= 0 && w < W;
        float f[8];
        unpack8(v[r * 3 + s], f);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (inside && f[j] > best[j]) { best[j] = f[j]; arg[j] = r * 3 + s; }
      }
    }
    y[i] = pack8(best);
    uint2 a;
    a.x = arg[0] | (arg[1] << 8) | (arg[2] << 16) | (arg[3] << 24);
    a.y = arg[4] | (arg[5] << 8) | (arg[6] << 16) | (arg[7] << 24);
    argmax[i] = a;
  }
}

// One thread per 2x2 input patch x 8 channels.  The patch with origin (2k, 2l) is covered by the
// windows (k, l), (k, l+1), (k+1, l), (k+1, l+1) only, so four (arg-max, dy) reads produce four
// outputs (the per-pixel formulation needs nine); no atomics: every input pixel has one writer.
__global__ void __launch_bounds__(256)
maxpool_bwd_kernel(const uint4* __restrict__ dy, const uint2* __restrict__ argmax,
                   uint4* __restrict__ dx, int NB, int H, int W, int C) {
  const int Ho = H / 2, Wo = W / 2, cv = C / 8;
  const long long total = (long long)NB * Ho * Wo * cv;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % cv);
    long long pix = i / cv;
    const int l = (int)(pix % Wo); pix /= Wo;
    const int k = (int)(pix % Ho);
    const int n = (int)(pix / Ho);
    float g[4][8];
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
      for (int j = 0; j < 8; ++j) g[q][j] = 0.f;
    // window (k + a, l + b) reaches patch pixel q = 2*py + px with arg-max code kCode[a][b][q]
    // (code = r*3 + s of the window element; 255 = the window does not cover that pixel)
    // the four (arg-max, dy) pairs are loaded up front from clamped (always valid) addresses; windows that do
    // not exist contribute nothing
    uint2 amv[4];
    uint4 dyv[4];
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        const long long o = (((long long)n * Ho + min(k + a, Ho - 1)) * Wo + min(l + b, Wo - 1)) * cv + c;
        amv[a * 2 + b] = argmax[o];
        dyv[a * 2 + b] = ld_nc_v4(dy + o);
      }
#pragma unroll
    for (int a = 0; a < 2; ++a) {
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        const bool exists = (k + a < Ho) && (l + b < Wo);
        const uint2 am = amv[a * 2 + b];
        float d[8];
        unpack8(dyv[a * 2 + b], d);
#pragma unroll
        for (int py = 0; py < 2; ++py) {
          const int r = 2 * (0 - a) + py + 1;          // row of the pixel inside window k + a
          if (r < 0 || r > 2 || (a == 1 && py == 0)) continue;
#pragma unroll
          for (int px = 0; px < 2; ++px) {
            const int sidx = 2 * (0 - b) + px + 1;
            if (sidx < 0 || sidx > 2 || (b == 1 && px == 0)) continue;
            const uint32_t code = r * 3 + sidx;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const uint32_t aj = ((j < 4 ? am.x : am.y) >> ((j & 3) * 8)) & 0xff;
              if (exists && aj == code) g[py * 2 + px][j] += d[j];
            }
          }
        }
      }
    }
#pragma unroll
    for (int py = 0; py < 2; ++py)
#pragma unroll
      for (int px = 0; px < 2; ++px)
        dx[(((long long)n * H + 2 * k + py) * W + 2 * l + px) * cv + c] = pack8(g[py * 2 + px]);
  }
}

// y[n, c] = mean over HW of x[n, hw, c]. One CTA per (n, 32-vector slab of channels).
__global__ void __launch_bounds__(256)
avgpool_fwd_kernel(const uint4* __restrict__ x, uint4* __restrict__ y, int HW, int C) {
  __shared__ float red[8][32][9];
  const int cv = C / 8;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int col = blockIdx.x * 32 + tx;
  const int n = blockIdx.y;
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  if (col < cv)
    for (int p = ty; p < HW; p += 8) {
      float f[8];
      unpack8(ld_nc_v4(x + ((long long)n * HW + p) * cv + col), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += f[j];
    }
#pragma unroll
  for (int j = 0; j < 8; ++j) red[ty][tx][j] = acc[j];
  __syncthreads();
  if (ty == 0 && col < cv) {
    for (int r = 1; r < 8; ++r)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += red[r][tx][j];
    const float inv = 1.f / (float)HW;
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] *= inv;
    y[(long long)n * cv + col] = pack8(acc);
  }
}

__global__ void __launch_bounds__(256)
avgpool_bwd_kernel(const uint4* __restrict__ dy, uint4* __restrict__ dx, int NB, int HW, int C) {
  const int cv = C / 8;
  const long long total = (long long)NB * HW * cv;
  const float inv = 1.f / (float)HW;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % cv);
    const int n = (int)(i / ((long long)HW * cv));
    float f[8];
    unpack8(dy[(long long)n * cv + c], f);
#pragma unroll
    for (int j = 0; j < 8; ++j) f[j] *= inv;
    dx[i] = pack8(f);
  }
}

static int grid_for(long long total) {
  long long b = (total + 255) / 256;
  if (b > 148 * 16) b = 148 * 16;
  if (b < 1) b = 1;
  return (int)b;
}

}  // namespace b200

using namespace b200;

extern "C" int b200dt_maxpool3x3s2_fwd(const void* x, void* y, uint8_t* argmax, int NB, int H, int W,
                                       int C, void* stream) {
  if ((C % 8) || ((H | W) & 1)) return -3;
  const long long total = (long long)NB * (H / 2) * (W / 2) * (C / 8);
  maxpool_fwd_kernel<<<grid_for(total), 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const uint4*>(x), reinterpret_cast<uint4*>(y),
      reinterpret_cast<uint2*>(argmax), NB, H, W, C);
  return (int)cudaGetLastError();
}

extern "C" int b200dt_maxpool3x3s2_bwd(const void* dy, const uint8_t* argmax, void* dx, int NB,
                                       int H, int W, int C, void* stream) {
  if ((C % 8) || ((H | W) & 1)) return -3;
  const long long total = (long long)NB * (H / 2) * (W / 2) * (C / 8);   // one thread per 2x2 patch
  maxpool_bwd_kernel<<<grid_for(total), 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const uint4*>(dy), reinterpret_cast<const uint2*>(argmax),
      reinterpret_cast<uint4*>(dx), NB, H, W, C);
  return (int)cudaGetLastError();
}

extern "C" int b200dt_avgpool_fwd(const void* x, void* y, int NB, int HW, int C, void* stream) {
  if (C % 8) return -3;
  dim3 grid((C / 8 + 31) / 32, NB);
  avgpool_fwd_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const uint4*>(x),
                                                             reinterpret_cast<uint4*>(y), HW, C);
  return (int)cudaGetLastError();
}

extern "C" int b200dt_avgpool_bwd(const void* dy, void* dx, int NB, int HW, int C, void* stream) {
  if (C % 8) return -3;
  const long long total = (long long)NB * HW * (C / 8);
  avgpool_bwd_kernel<<<grid_for(total), 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const uint4*>(dy), reinterpret_cast<uint4*>(dx), NB, HW, C);
  return (int)cudaGetLastError();
}
