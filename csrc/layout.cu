// Layout / dtype helpers: fp32 -> bf16 shadow casts, 2-D transposes (K-major weight copies for
// dgrad), NCHW fp32 -> NHWC bf16, and the 7x7 stride-2 stem gather that turns the 3-channel input
// into a K-major matrix for the tcgen05 GEMM (replaces the gather inside cuDNN's stem conv,
// /root/reference/03.model_parallel.ipynb:812).
#include "common.cuh"
#include "ptx.cuh"
#include "../include/b200dt.h"

namespace b200 {

__global__ void cast_f32_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst,
                                     long long n) {
  const long long n8 = n / 8;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const bool aligned = ((reinterpret_cast<uintptr_t>(src) & 15) == 0) &&
                       ((reinterpret_cast<uintptr_t>(dst) & 15) == 0);
  long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (aligned) {
    for (; i < n8; i += stride) {
      const float4 a = reinterpret_cast<const float4*>(src)[2 * i];
      const float4 b = reinterpret_cast<const float4*>(src)[2 * i + 1];
      uint4 u;
      u.x = pack_bf16x2(a.x, a.y); u.y = pack_bf16x2(a.z, a.w);
      u.z = pack_bf16x2(b.x, b.y); u.w = pack_bf16x2(b.z, b.w);
      reinterpret_cast<uint4*>(dst)[i] = u;
    }
    for (long long j = n8 * 8 + blockIdx.x * (long long)blockDim.x + threadIdx.x; j < n; j += stride)
      dst[j] = __float2bfloat16(src[j]);
  } else {
    for (; i < n; i += stride) dst[i] = __float2bfloat16(src[i]);
  }
}

// 32x32 tile transpose through padded shared memory.
__global__ void transpose_bf16_kernel(const __nv_bfloat16* __restrict__ src,
                                      __nv_bfloat16* __restrict__ dst, int rows, int cols) {
  __shared__ __nv_bfloat16 tile[32][33];
  const int bx = blockIdx.x * 32, by = blockIdx.y * 32;
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    const int r = by + j, c = bx + threadIdx.x;
    if (r < rows && c < cols) tile[j][threadIdx.x] = src[(long long)r * cols + c];
  }
  __syncthreads();
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    const int c = bx + j, r = by + threadIdx.x;
    if (r < rows && c < cols) dst[(long long)c * rows + r] = tile[threadIdx.x][j];
  }
}

// One CTA per (n, h): reads the C rows of W floats coalesced, writes W*C bf16 contiguous.
__global__ void nchw_to_nhwc_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ y,
                                    int C, int H, int W) {
  extern __shared__ float s[];   // [C][W+1]
  const int n = blockIdx.y, h = blockIdx.x;
  for (int i = threadIdx.x; i < C * W; i += blockDim.x) {
    const int c = i / W, w = i % W;
    s[c * (W + 1) + w] = x[(((long long)n * C + c) * H + h) * W + w];
  }
  __syncthreads();
  __nv_bfloat16* out = y + ((long long)n * H + h) * W * C;
  for (int i = threadIdx.x; i < C * W; i += blockDim.x) {
    const int w = i / C, c = i % C;
    out[i] = __float2bfloat16(s[c * (W + 1) + w]);
  }
}

// Stem gather. Output row = (n, ho, wo); column k = (r*7 + s)*3 + c for k < 147, zero up to ldk.
// One CTA handles kStemRows consecutive output rows of one image: the 2*kStemRows + 5 input rows
// they share are staged ONCE in shared memory as bf16, channel-interleaved ([row][w + 3][c], zero
// padded), so that the 21 values of one filter row are contiguous; then thread t owns one 16-byte
// piece (v8 = t % (ldk/8)) of the K-major output rows, keeps its eight source offsets in registers
// and walks over the output pixels: eight 2-byte shared loads, one coalesced 16-byte store.
constexpr int kStemRows = 4;

__global__ void __launch_bounds__(256)
stem_im2col_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ cols, int H, int W,
                   int ldk) {
  extern __shared__ __align__(16) unsigned short sb[];   // [2*kStemRows + 5][(W + 6) * 3] bf16 bits
  const int Ho = H / 2, Wo = W / 2;
  const int n = blockIdx.y, ho0 = blockIdx.x * kStemRows;
  const int PW3 = (W + 6) * 3;
  const int nrows = 2 * kStemRows + 5;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // zero the left / right padding columns (3 pixels x 3 channels on both sides of every row)
  for (int i = threadIdx.x; i < nrows * 18; i += blockDim.x) {
    const int r = i / 18, e = i % 18;
    sb[r * PW3 + (e < 9 ? e : PW3 - 18 + e)] = 0;
  }
  // (row, channel) pairs are spread over the warps; a warp reads its W floats coalesced
  for (int rc = warp; rc < nrows * 3; rc += 8) {
    const int r = rc / 3, c = rc - r * 3;
    const int h = ho0 * 2 - 3 + r;
    unsigned short* dst = sb + r * PW3 + 9 + c;
    if (h >= 0 && h < H) {
      const float* src = x + (((long long)n * 3 + c) * H + h) * W;
      for (int w = lane; w < W; w += 32)
        dst[w * 3] = __bfloat16_as_ushort(__float2bfloat16(__ldg(src + w)));
    } else {
      for (int w = lane; w < W; w += 32) dst[w * 3] = 0;
    }
  }
  __syncthreads();
  const int kv = ldk / 8;                       // 16-byte vectors per output row
  const int per_pass = blockDim.x / kv;         // output pixels per pass over the CTA
  const int v8 = threadIdx.x % kv, slot = threadIdx.x / kv;
  if (slot >= per_pass) return;
  int off[8];                                   // k -> offset inside the staged rows, -1 = padding
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int k = v8 * 8 + j;
    off[j] = k < 147 ? (k / 21) * PW3 + (k % 21) : -1;
  }
  const int total = kStemRows * Wo;
  for (int q = slot; q < total; q += per_pass) {
    const int rr = q / Wo, wo = q - rr * Wo;
    const int ho = ho0 + rr;
    if (ho >= Ho) break;
    const unsigned short* base = sb + (2 * rr) * PW3 + wo * 6;
    uint32_t u[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint32_t lo = off[2 * j] >= 0 ? base[off[2 * j]] : 0u;
      const uint32_t hi = off[2 * j + 1] >= 0 ? base[off[2 * j + 1]] : 0u;
      u[j] = lo | (hi << 16);
    }
    *reinterpret_cast<uint4*>(cols + (((long long)n * Ho + ho) * Wo + wo) * ldk + v8 * 8) =
        make_uint4(u[0], u[1], u[2], u[3]);
  }
}

}  // namespace b200

using namespace b200;

extern "C" int b200dt_cast_f32_to_bf16(const float* src, void* dst, long long n, void* stream) {
  if (n <= 0) return 0;
  long long blocks = (n / 8 + 255) / 256 + 1;
  if (blocks > 148 * 16) blocks = 148 * 16;
  cast_f32_bf16_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(
      src, reinterpret_cast<__nv_bfloat16*>(dst), n);
  return (int)cudaGetLastError();
}

extern "C" int b200dt_transpose_bf16(const void* src, void* dst, int rows, int cols, void* stream) {
  if (rows <= 0 || cols <= 0) return 0;
  dim3 grid((cols + 31) / 32, (rows + 31) / 32);
  transpose_bf16_kernel<<<grid, dim3(32, 8), 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const __nv_bfloat16*>(src), reinterpret_cast<__nv_bfloat16*>(dst), rows, cols);
  return (int)cudaGetLastError();
}

extern "C" int b200dt_nchw_f32_to_nhwc_bf16(const float* x, void* y, int NB, int C, int H, int W,
                                            void* stream) {
  const size_t smem = sizeof(float) * C * (W + 1);
  if (smem > 48 * 1024) return -3;
  nchw_to_nhwc_kernel<<<dim3(H, NB), 256, smem, (cudaStream_t)stream>>>(
      x, reinterpret_cast<__nv_bfloat16*>(y), C, H, W);
  return (int)cudaGetLastError();
}

extern "C" int b200dt_stem_im2col(const float* x, void* cols, int NB, int H, int W, int ldk,
                                  void* stream) {
  if ((H | W) & 1 || ldk < 152 || (ldk % 8)) return -3;
  const size_t smem = sizeof(unsigned short) * (2 * kStemRows + 5) * 3 * (W + 6);
  if (smem > 48 * 1024 || ldk / 8 > 256) return -3;
  stem_im2col_kernel<<<dim3((H / 2 + kStemRows - 1) / kStemRows, NB), 256, smem, (cudaStream_t)stream>>>(
      x, reinterpret_cast<__nv_bfloat16*>(cols), H, W, ldk);
  return (int)cudaGetLastError();
}
