// ResNet stem (7x7, stride 2, pad 3, Cin = 3 -> Cout = 64) as an IMPLICIT GEMM on tcgen05: the patch
// matrix is built tile by tile in shared memory, straight in the 128-byte-swizzled operand layout the tensor
// core reads, and never exists in global memory (the explicit im2col of csrc/layout.cu materialised
// NB*112*112 x 160 bf16 = 1 GB at batch 256 and the GEMM and the weight gradient each read it back).
// Replaces the cuDNN stem convolution the reference reaches through torchvision's ResNet
// (/root/reference/03.model_parallel.ipynb:812, tv:models/resnet.py conv1).
//
// Forward : y[n,oh,ow,:] (NHWC bf16) = sum_k patch[(n,oh,ow), k] * w[:, k],  k = (ky*7 + kx)*3 + c  (147 -> 160)
//           one CTA = tiles of 16 x 8 output pixels; A = patch tile (K-major, 128 rows), B = filter (K-major).
// Wgrad   : dw[co, k] = sum_pixels dy[pixel, co] * patch[pixel, k]; the same patch tile is the MN-major
//           operand (rows = reduction index), D = [k (2 x 128 rows), co (64 columns)] in TMEM, one partial per
//           CTA written to a scratch slab and folded in CTA order by a second kernel (deterministic).
// Two CTAs per SM: while one builds its next patch tile the other runs its MMAs / epilogue.
#include "common.cuh"
#include "ptx.cuh"
#include "../include/b200dt.h"

namespace b200 {

constexpr int kStemThreads = 256;
constexpr int kStemK = 147;            // 7 * 7 * 3
constexpr int kStemChunks = 19;        // 16-byte (8-element) pieces of a patch row that hold data (19 * 8 = 152)
constexpr int kStemTW = 16;            // output pixels per tile row

// Staged input: bf16, channel-interleaved [row][(w + 3) * 3 + c], zero padded, so that the 21 values of one
// filter row of one output pixel are contiguous: patch element k = ky * 21 + j lives at
// stage[(2*py + ky) * pitch + 2*px*3 + j].
template <int kRowsOut>
struct StemGeom {
  static constexpr int kInRows = 2 * kRowsOut + 5;
  static constexpr int kInCols = 2 * kStemTW + 5;          // 37
  static constexpr int kPitch = kInCols * 3 + 1;           // 112 bf16 per staged row
  static constexpr int kStageBytes = ((kInRows * kPitch * 2 + 127) / 128) * 128;
  static constexpr int kPixels = kRowsOut * kStemTW;
};

// Input window of one tile, software-pipelined: every thread owns a fixed set of window elements (their
// (row, col, channel) decomposition never changes, so no division per tile); `fetch` issues the global loads of
// the NEXT tile into registers while the current tile is being built / multiplied, `store` writes them (as
// bf16) into the staging buffer at the top of the next iteration.
template <int kRowsOut>
struct StemWindow {
  using G = StemGeom<kRowsOut>;
  static constexpr int kElems = G::kInRows * 3 * G::kInCols;
  static constexpr int kPer = (kElems + kStemThreads - 1) / kStemThreads;
  int rc[kPer];        // row << 8 | col   (-1: this thread has no such element)
  int goff[kPer];      // (c * H + row) * W + col, relative to the window origin
  int soff[kPer];      // staging offset row * pitch + col * 3 + c
  float val[kPer];

  __device__ __forceinline__ void init(int H, int W) {
#pragma unroll
    for (int j = 0; j < kPer; ++j) {
      const int i = threadIdx.x + j * kStemThreads;
      if (i < kElems) {
        const int col = i % G::kInCols, q = i / G::kInCols;
        const int c = q % 3, r = q / 3;
        rc[j] = (r << 8) | col;
        goff[j] = (c * H + r) * W + col;
        soff[j] = r * G::kPitch + col * 3 + c;
      } else {
        rc[j] = -1; goff[j] = 0; soff[j] = 0;
      }
    }
  }
  __device__ __forceinline__ void fetch(const float* __restrict__ x, int n, int oh0, int ow0, int H, int W) {
    const int ih0 = 2 * oh0 - 3, iw0 = 2 * ow0 - 3;
    const float* base = x + ((long long)n * 3 * H + ih0) * W + iw0;
#pragma unroll
    for (int j = 0; j < kPer; ++j) {
      const int ih = ih0 + (rc[j] >> 8), iw = iw0 + (rc[j] & 255);
      const bool ok = rc[j] >= 0 && ih >= 0 && ih < H && iw >= 0 && iw < W;
      val[j] = ok ? __ldg(base + goff[j]) : 0.f;
    }
  }
  __device__ __forceinline__ void store(unsigned short* stage) const {
#pragma unroll
    for (int j = 0; j < kPer; ++j)
      if (rc[j] >= 0) stage[soff[j]] = __bfloat16_as_ushort(__float2bfloat16(val[j]));
  }
};

// Build the patch tile [pixels][160] from the staged window into 64-element-wide, 128-byte-swizzled blocks:
// block kb (k = 64*kb ..) starts at tile + kb * block_bytes, pixel row p at p * 128, 16-byte chunk (k%64)/8
// XOR (p & 7).  Thread t owns chunk t % 19 (its eight source offsets never change) and walks over the pixels.
template <int kRowsOut>
__device__ __forceinline__ void stem_build_patch(const unsigned short* stage, uint8_t* tile, int block_bytes) {
  using G = StemGeom<kRowsOut>;
  constexpr int kLanes = kStemThreads / kStemChunks;       // 13 pixel lanes
  const int v8 = threadIdx.x % kStemChunks;
  const int lane = threadIdx.x / kStemChunks;
  if (lane >= kLanes) return;
  int off[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const int k = v8 * 8 + e;
    off[e] = k < kStemK ? (k / 21) * G::kPitch + (k % 21) : -1;
  }
  const int kb = v8 >> 3, ch = v8 & 7;
  uint8_t* blk = tile + kb * block_bytes;
  for (int p = lane; p < G::kPixels; p += kLanes) {
    const int py = p / kStemTW, px = p % kStemTW;
    const unsigned short* src = stage + (2 * py) * G::kPitch + (2 * px) * 3;
    uint32_t w[4];
#pragma unroll
    for (int e = 0; e < 8; e += 2) {
      const uint32_t lo = off[e] >= 0 ? src[off[e]] : 0u;
      const uint32_t hi = off[e + 1] >= 0 ? src[off[e + 1]] : 0u;
      w[e >> 1] = lo | (hi << 16);
    }
    *reinterpret_cast<uint4*>(blk + p * 128 + ((ch ^ (p & 7)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
  }
}

__device__ __forceinline__ void zero_smem(uint8_t* p, int bytes) {
  for (int i = threadIdx.x * 16; i < bytes; i += kStemThreads * 16)
    *reinterpret_cast<uint4*>(p + i) = make_uint4(0, 0, 0, 0);
}

struct StemBars {
  uint64_t mma_done;
  uint32_t tmem_base;
  uint32_t pad;
};

// ------------------------------------------------------------------------------------------ forward
constexpr int kFwdRows = 8;                                   // 16 x 8 = 128 output pixels per tile
constexpr int kFwdABytes = 3 * 16384, kFwdBBytes = 3 * 8192, kFwdOutBytes = 4 * 4096;
constexpr int kFwdStage = StemGeom<kFwdRows>::kStageBytes;
constexpr int kFwdSmem = kFwdABytes + kFwdBBytes + kFwdOutBytes + kFwdStage + 4 * 2 * 64 * 4 + 64 + 1024;

__global__ void __launch_bounds__(kStemThreads, 2)
stem_fprop_kernel(const float* __restrict__ x, const float* __restrict__ w, __nv_bfloat16* __restrict__ y,
                  float* bn_sums, int bn_reps, int det, int NB, int H, int W) {
  extern __shared__ uint8_t smem_raw[];
  // 1024-byte alignment for the swizzle atoms, computed as an OFFSET so the compiler keeps the shared address space
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sA = smem;
  uint8_t* sB = sA + kFwdABytes;
  uint8_t* sOut = sB + kFwdBBytes;
  unsigned short* stage = reinterpret_cast<unsigned short*>(sOut + kFwdOutBytes);
  float* cs = reinterpret_cast<float*>(sOut + kFwdOutBytes + kFwdStage);      // [4 warps][2][64]
  StemBars* bars = reinterpret_cast<StemBars*>(cs + 4 * 2 * 64);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int Ho = H / 2, Wo = W / 2;
  const int tiles_x = Wo / kStemTW, tiles_y = Ho / kFwdRows;
  const int ntiles = NB * tiles_x * tiles_y;

  if (threadIdx.x == 0) {
    mbar_init(&bars->mma_done, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<64>(&bars->tmem_base);
  zero_smem(sA, kFwdABytes + kFwdBBytes);
  __syncthreads();
  // filter -> bf16 K-major blocks [64 rows][64 k], row n at n * 128, chunk (k % 64) / 8 XOR (n & 7)
  for (int i = threadIdx.x; i < 64 * kStemChunks; i += kStemThreads) {
    const int n = i / kStemChunks, v8 = i % kStemChunks;
    uint32_t wv[4];
#pragma unroll
    for (int e = 0; e < 8; e += 2) {
      const int k = v8 * 8 + e;
      const float a = k < kStemK ? __ldg(w + n * kStemK + k) : 0.f;
      const float b = k + 1 < kStemK ? __ldg(w + n * kStemK + k + 1) : 0.f;
      wv[e >> 1] = pack_bf16x2(a, b);
    }
    *reinterpret_cast<uint4*>(sB + (v8 >> 3) * 8192 + n * 128 + (((v8 & 7) ^ (n & 7)) << 4)) =
        make_uint4(wv[0], wv[1], wv[2], wv[3]);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;
  const uint32_t idesc = umma_idesc_bf16(128, 64, false, false);

  float2 sm2[4], sq2[4];                       // running column partials of this lane (stats)
#pragma unroll
  for (int k = 0; k < 4; ++k) { sm2[k] = make_float2(0.f, 0.f); sq2[k] = make_float2(0.f, 0.f); }
  uint32_t phase = 0;
  StemWindow<kFwdRows> win;
  win.init(H, W);
  if (blockIdx.x < ntiles) {
    const int t = blockIdx.x;
    win.fetch(x, t / (tiles_x * tiles_y), ((t / tiles_x) % tiles_y) * kFwdRows, (t % tiles_x) * kStemTW, H, W);
  }
  for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
    const int tx = t % tiles_x, ty = (t / tiles_x) % tiles_y, n = t / (tiles_x * tiles_y);
    const int oh0 = ty * kFwdRows, ow0 = tx * kStemTW;
    win.store(stage);
    tc_fence_before();
    __syncthreads();              // window staged; last tile's TMEM reads and MMA operand reads are over
    {
      const int tn = t + gridDim.x;          // next tile's loads fly while this one is built and multiplied
      if (tn < ntiles)
        win.fetch(x, tn / (tiles_x * tiles_y), ((tn / tiles_x) % tiles_y) * kFwdRows, (tn % tiles_x) * kStemTW, H, W);
    }
    stem_build_patch<kFwdRows>(stage, sA, 16384);
    fence_proxy_async();          // generic-proxy writes of the operand tile -> visible to the tensor core
    __syncthreads();
    if (threadIdx.x == 0) {
      tc_fence_after();
      const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB);
#pragma unroll
      for (int kb = 0; kb < 3; ++kb) {
        const int ks_n = kb < 2 ? 4 : 2;                          // K = 160 = 64 + 64 + 32
        for (int ks = 0; ks < ks_n; ++ks)
          umma_f16(tmem_base, umma_desc_sw128(a0 + kb * 16384 + ks * 32, 16, 1024),
                   umma_desc_sw128(b0 + kb * 8192 + ks * 32, 16, 1024), idesc, (kb | ks) ? 1u : 0u);
      }
      umma_commit(&bars->mma_done);
    }
    mbar_wait(&bars->mma_done, phase);
    phase ^= 1;
    tc_fence_after();
    if (warp < 4) {
      uint32_t v0[32], v1[32];
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(warp * 32) << 16);
      tmem_ld_32x32(taddr, v0);
      tmem_ld_32x32(taddr + 32, v1);
      tmem_ld_wait();
      uint8_t* buf = sOut + warp * 4096;
      uint8_t* my_row = buf + lane * 128;
      const int sw = lane & 7;
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const uint32_t* v = half ? v1 : v0;
#pragma unroll
        for (int j = 0; j < 32; j += 8) {
          float f[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(v[j + e]);
          *reinterpret_cast<uint4*>(my_row + (((half * 4 + (j >> 3)) ^ sw) << 4)) = pack8(f);
        }
      }
      __syncwarp();
      const int c8 = lane & 7, rg = lane >> 3;
      if (bn_sums != nullptr) {
        const uint8_t* rowp = buf + rg * (8 * 128);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const uint4 u = *reinterpret_cast<const uint4*>(rowp + i * 128 + ((c8 ^ i) << 4));
          const uint32_t wv[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const float2 f = make_float2(__uint_as_float(wv[k] << 16), __uint_as_float(wv[k] & 0xffff0000u));
            sm2[k] = __fadd2_rn(sm2[k], f);
            sq2[k] = __ffma2_rn(f, f, sq2[k]);
          }
        }
      }
      // rows 32*warp .. +31 of the tile = output rows 2*warp, 2*warp + 1 (16 pixels x 128 B contiguous each)
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int row = i * 4 + rg;
        const uint4 u = *reinterpret_cast<const uint4*>(buf + row * 128 + ((c8 ^ (row & 7)) << 4));
        const int py = 2 * warp + (row >> 4), px = row & 15;
        __nv_bfloat16* dst = y + ((((long long)n * Ho + oh0 + py) * Wo + ow0 + px) * 64 + c8 * 8);
        *reinterpret_cast<uint4*>(dst) = u;
      }
      __syncwarp();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<64>(tmem_base);
  }
  if (bn_sums == nullptr) return;
  // per-CTA column statistics: row groups by shuffle, the four epilogue warps through shared memory (fixed order)
  if (warp < 4) {
#pragma unroll
    for (int off = 8; off <= 16; off <<= 1)
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        sm2[k].x += __shfl_xor_sync(0xffffffffu, sm2[k].x, off);
        sm2[k].y += __shfl_xor_sync(0xffffffffu, sm2[k].y, off);
        sq2[k].x += __shfl_xor_sync(0xffffffffu, sq2[k].x, off);
        sq2[k].y += __shfl_xor_sync(0xffffffffu, sq2[k].y, off);
      }
    if (lane < 8) {
      float* d = cs + warp * 128 + lane * 8;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        d[2 * k] = sm2[k].x; d[2 * k + 1] = sm2[k].y;
        d[64 + 2 * k] = sq2[k].x; d[64 + 2 * k + 1] = sq2[k].y;
      }
    }
  }
  __syncthreads();
  if (threadIdx.x < 128) {
    const int j = threadIdx.x;                 // 0..63 sums, 64..127 sums of squares
    const float v = ((cs[j] + cs[128 + j]) + cs[256 + j]) + cs[384 + j];
    float* dst = bn_sums + (size_t)(det ? blockIdx.x : blockIdx.x % bn_reps) * 128 + j;
    if (det) *dst = v; else atomicAdd(dst, v);
  }
  __shared__ bool is_last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned int* ticket = reinterpret_cast<unsigned int*>(bn_sums + (size_t)bn_reps * 128);
    is_last = atomicAdd(ticket, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (is_last) {
    __threadfence();
    fold_sum_copies(bn_sums, 128, det ? min(bn_reps, (int)gridDim.x) : bn_reps, threadIdx.x, blockDim.x);
  }
}

// ------------------------------------------------------------------------------------------ weight gradient
// One k-block = 64 output pixels (16 x 4).  A = patch block, MN-major: rows = pixels (reduction index), four
// 64-wide chunks of patch elements (the fourth is all zero: M is padded to 2 x 128); B = dy block, MN-major:
// rows = pixels, 64 output channels.  D0 = patch elements 0..127, D1 = 128..255 (only 128..146 are real).
constexpr int kWgRows = 4;
constexpr int kWgABytes = 4 * 8192, kWgBBytes = 8192;
constexpr int kWgStage = StemGeom<kWgRows>::kStageBytes;
constexpr int kWgSmem = kWgABytes + kWgBBytes + kWgStage + 64 + 1024;

__global__ void __launch_bounds__(kStemThreads, 2)
stem_wgrad_kernel(const float* __restrict__ x, const __nv_bfloat16* __restrict__ dy, float* __restrict__ partial,
                  int NB, int H, int W) {
  extern __shared__ uint8_t smem_raw[];
  // 1024-byte alignment for the swizzle atoms, computed as an OFFSET so the compiler keeps the shared address space
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* sA = smem;
  uint8_t* sB = sA + kWgABytes;
  unsigned short* stage = reinterpret_cast<unsigned short*>(sB + kWgBBytes);
  StemBars* bars = reinterpret_cast<StemBars*>(sB + kWgBBytes + kWgStage);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int Ho = H / 2, Wo = W / 2;
  const int tiles_x = Wo / kStemTW, tiles_y = Ho / kWgRows;
  const int nblocks = NB * tiles_x * tiles_y;

  if (threadIdx.x == 0) {
    mbar_init(&bars->mma_done, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<128>(&bars->tmem_base);
  zero_smem(sA, kWgABytes);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;
  const uint32_t idesc = umma_idesc_bf16(128, 64, true, true);       // both operands MN-major
  uint32_t phase = 0;
  bool first = true;
  StemWindow<kWgRows> win;
  win.init(H, W);
  uint4 dyv[2];                   // this thread's two 16-byte pieces of the next dy block
  auto fetch_dy = [&](int tt) {
    const int n = tt / (tiles_x * tiles_y), oh0 = ((tt / tiles_x) % tiles_y) * kWgRows, ow0 = (tt % tiles_x) * kStemTW;
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int i = threadIdx.x + j * kStemThreads;
      const int p = i >> 3, c = i & 7;
      dyv[j] = ld_nc_v4(dy + ((((long long)n * Ho + oh0 + p / kStemTW) * Wo + ow0 + p % kStemTW) * 64 + c * 8));
    }
  };
  if (blockIdx.x < nblocks) {
    const int t = blockIdx.x;
    win.fetch(x, t / (tiles_x * tiles_y), ((t / tiles_x) % tiles_y) * kWgRows, (t % tiles_x) * kStemTW, H, W);
    fetch_dy(t);
  }
  for (int t = blockIdx.x; t < nblocks; t += gridDim.x) {
    win.store(stage);
    if (!first) {                 // the previous block's MMAs must have consumed sA / sB
      mbar_wait(&bars->mma_done, phase);
      phase ^= 1;
    }
    __syncthreads();
    // dy block: pixel p = (py, px) -> row p (128 B = 64 channels), chunk c XOR (p & 7)
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int i = threadIdx.x + j * kStemThreads;
      const int p = i >> 3, c = i & 7;
      *reinterpret_cast<uint4*>(sB + p * 128 + ((c ^ (p & 7)) << 4)) = dyv[j];
    }
    {
      const int tn = t + gridDim.x;          // next block's loads fly while this one is built and multiplied
      if (tn < nblocks) {
        win.fetch(x, tn / (tiles_x * tiles_y), ((tn / tiles_x) % tiles_y) * kWgRows, (tn % tiles_x) * kStemTW, H, W);
        fetch_dy(tn);
      }
    }
    stem_build_patch<kWgRows>(stage, sA, 8192);
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x == 0) {
      tc_fence_after();
      const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB);
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {                                 // 16 pixels (k-rows) per MMA step
        const uint64_t db = umma_desc_sw128(b0 + ks * 2048, 8192, 1024);
        umma_f16(tmem_base, umma_desc_sw128(a0 + ks * 2048, 8192, 1024), db, idesc, (!first || ks) ? 1u : 0u);
        umma_f16(tmem_base + 64, umma_desc_sw128(a0 + 16384 + ks * 2048, 8192, 1024), db, idesc,
                 (!first || ks) ? 1u : 0u);
      }
      umma_commit(&bars->mma_done);
    }
    first = false;
  }
  if (!first) mbar_wait(&bars->mma_done, phase);
  tc_fence_after();
  // partial[cta][k][co]: lane = patch element, 64 output channels per lane (D1 only below k = 147)
  float* mine = partial + (size_t)blockIdx.x * (kStemK * 64);
  if (warp < 4) {
    for (int half = 0; half < 2; ++half) {
      const int k = half * 128 + warp * 32 + lane;
      if (half == 1 && warp * 32 >= kStemK - 128) continue;            // warp-uniform: nothing real in these lanes
      uint32_t v0[32], v1[32];
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(warp * 32) << 16) + half * 64;
      tmem_ld_32x32(taddr, v0);
      tmem_ld_32x32(taddr + 32, v1);
      tmem_ld_wait();
      if (k < kStemK) {
        float4* d = reinterpret_cast<float4*>(mine + (size_t)k * 64);
#pragma unroll
        for (int j = 0; j < 32; j += 4) {
          d[j >> 2] = make_float4(first ? 0.f : __uint_as_float(v0[j]), first ? 0.f : __uint_as_float(v0[j + 1]),
                                  first ? 0.f : __uint_as_float(v0[j + 2]), first ? 0.f : __uint_as_float(v0[j + 3]));
          d[8 + (j >> 2)] = make_float4(first ? 0.f : __uint_as_float(v1[j]), first ? 0.f : __uint_as_float(v1[j + 1]),
                                        first ? 0.f : __uint_as_float(v1[j + 2]), first ? 0.f : __uint_as_float(v1[j + 3]));
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<128>(tmem_base);
  }
}

// dw[co][k] = sum over CTAs of partial[cta][k][co], in a fixed order: one block per k, thread (slice, co) adds
// every fourth partial (coalesced 256-byte rows, eight loads in flight), the four slices are added in slice order.
__global__ void __launch_bounds__(256)
stem_wgrad_fold_kernel(const float* __restrict__ partial, float* __restrict__ dw, int nparts) {
  __shared__ float red[4][64];
  const int k = blockIdx.x, co = threadIdx.x & 63, slice = threadIdx.x >> 6;
  const float* src = partial + (size_t)k * 64 + co;
  constexpr size_t kPart = (size_t)kStemK * 64;
  float acc = 0.f;
  int c = slice;
  for (; c + 28 < nparts; c += 32) {
    float v[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) v[u] = src[(size_t)(c + 4 * u) * kPart];
#pragma unroll
    for (int u = 0; u < 8; ++u) acc += v[u];
  }
  for (; c < nparts; c += 4) acc += src[(size_t)c * kPart];
  red[slice][co] = acc;
  __syncthreads();
  if (slice == 0) dw[co * kStemK + k] = ((red[0][co] + red[1][co]) + red[2][co]) + red[3][co];
}

}  // namespace b200

using namespace b200;

static int stem_shape_ok(int NB, int H, int W) {
  return NB > 0 && H > 0 && W > 0 && (H % 2 == 0) && (W % 2 == 0) && ((W / 2) % kStemTW == 0) &&
         ((H / 2) % kFwdRows == 0);
}
static int stem_grid(int work, int cap_reps) {
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  int g = 2 * sms;
  if (cap_reps > 0 && g > cap_reps) g = cap_reps;
  return g < work ? g : work;
}

// y[NB,H/2,W/2,64] bf16 NHWC = conv7x7 s2 p3 of x[NB,3,H,W] fp32 NCHW with w[64,7,7,3] fp32 (OHWI).
// Returns -3 for shapes the tiling cannot take (caller falls back to im2col + GEMM).
extern "C" int b200dt_stem_conv_fprop(const float* x, const float* w, void* y, int NB, int H, int W,
                                      float* bn_sums, void* stream) {
  if (!stem_shape_ok(NB, H, W)) return -3;
  cudaStream_t st = (cudaStream_t)stream;
  static bool configured[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !configured[dev]) {
    cudaError_t e = cudaFuncSetAttribute(stem_fprop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kFwdSmem);
    if (e != cudaSuccess) return (int)e;
    e = cudaFuncSetAttribute(stem_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kWgSmem);
    if (e != cudaSuccess) return (int)e;
    configured[dev] = true;
  }
  const int det = deterministic(), reps = bn_replicas_used(64);
  if (bn_sums != nullptr) {
    cudaError_t e = sums_prezeroed() ? cudaSuccess
                                     : cudaMemsetAsync(bn_sums, 0, sizeof(float) * ((size_t)2 * 64 * reps + 1), st);
    if (e != cudaSuccess) return (int)e;
  }
  const int ntiles = NB * ((W / 2) / kStemTW) * ((H / 2) / kFwdRows);
  const int grid = stem_grid(ntiles, (bn_sums != nullptr && det) ? reps : 0);
  stem_fprop_kernel<<<grid, kStemThreads, kFwdSmem, st>>>(x, w, reinterpret_cast<__nv_bfloat16*>(y), bn_sums,
                                                         reps, det, NB, H, W);
  return (int)cudaGetLastError();
}

// Number of floats of scratch b200dt_stem_conv_wgrad needs (one 147 x 64 partial per CTA).
extern "C" long long b200dt_stem_wgrad_scratch_floats(void) {
  return (long long)stem_grid(1 << 30, 0) * kStemK * 64;
}

// dw[64,7,7,3] fp32 = sum over pixels of dy[NB,H/2,W/2,64] (bf16) x patches of x; deterministic.
extern "C" int b200dt_stem_conv_wgrad(const float* x, const void* dy, float* dw, float* scratch, int NB,
                                      int H, int W, void* stream) {
  if (!stem_shape_ok(NB, H, W) || scratch == nullptr) return -3;
  cudaStream_t st = (cudaStream_t)stream;
  static bool configured[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !configured[dev]) {
    cudaError_t e = cudaFuncSetAttribute(stem_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kWgSmem);
    if (e != cudaSuccess) return (int)e;
    configured[dev] = true;
  }
  const int nblocks = NB * ((W / 2) / kStemTW) * ((H / 2) / kWgRows);
  const int grid = stem_grid(nblocks, 0);
  stem_wgrad_kernel<<<grid, kStemThreads, kWgSmem, st>>>(x, reinterpret_cast<const __nv_bfloat16*>(dy), scratch,
                                                        NB, H, W);
  stem_wgrad_fold_kernel<<<kStemK, 256, 0, st>>>(scratch, dw, grid);
  return (int)cudaGetLastError();
}
