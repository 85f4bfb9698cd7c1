// Tiled-pixel implicit GEMM on tcgen05 / TMEM, fed by TMA.  sm_100a only.
//
// One persistent, warp-specialised kernel serves every GEMM-shaped op of the library:
//   * Linear forward / dgrad           (A = rows x K, K-major;  B = weights, K-major)
//   * Conv2d forward / dgrad, NHWC     (A = pixel boxes fetched by 4-D TMA with the filter-tap
//                                        offset folded into the box origin; padding = TMA
//                                        out-of-bounds zero fill; stride 2 = phase views)
//   * Linear / Conv2d weight gradient  (both operands MN-major, reduction over pixels,
//                                        split-K across CTAs with fp32 red.add)
//
// Roles per CTA (192 threads): warp 0 = TMA producer, warp 1 = tcgen05.mma issuer (+TMEM
// allocation), warps 2..5 = epilogue (tcgen05.ld -> registers -> global).  Two TMEM accumulator
// stages let the epilogue of tile i overlap the main loop of tile i+1.
//
// Replaces what the reference reaches through cuBLAS / cuDNN: nn.Linear
// (/root/reference/ddp_gpus.py:81, 01.data_parallel.ipynb:171, 03.model_parallel.ipynb:444-446)
// and the torchvision ResNet-50 convolutions (03.model_parallel.ipynb:807-833).
#include <stdlib.h>
#include "ptx.cuh"
#include "common.cuh"
#include "../include/b200dt.h"

namespace b200 {

constexpr int kBM = 128;          // UMMA M (cta_group::1)
constexpr int kBK = 64;           // K elements per stage (one 128-byte swizzle row)
constexpr int kThreads = 320;          // 2 control warps + 8 epilogue warps
constexpr int kMaxTaps = 9;
constexpr int kMaxMaps = 4;
constexpr int kColAccBytes = 4 * 2 * 256 * 4;  // fused BN statistics: {sum, sumsq}[256] of the current tile, one copy
                                               // per TMEM lane quarter (plain stores, no shared atomics); only
                                               // reserved when the launch asks for statistics
constexpr int kStageOutBytes = 128 * 128;   // 8 epilogue warps x (32 rows x 128 B); x2 = budget kept for staging

struct Tap {
  int8_t map, dw, dh, pad;
};

struct GemmParams {
  int kind;                 // 0 = FPROP-like (A pixels x K), 1 = WGRAD-like (reduce over pixels)
  int BN;                   // UMMA N: 64, 128 or 256 (multiple of 64)
  int stages;
  int tiles_m, tiles_n, splits;
  int kb_total, kb_per_split;
  // pixel tiling of the A operand (FPROP) / of the reduction (WGRAD)
  int bw, bh, bn;           // box extents; bw*bh*bn == 128 (FPROP) or 64 (WGRAD)
  int tiles_w, tiles_h;     // boxes per row / per column of one image
  int W, H, NB;             // valid extents for masking
  int cin_blocks;           // FPROP: Cin / 64 ; WGRAD: 64-wide N chunks per tap
  int ntaps;
  Tap taps[kMaxTaps];
  // epilogue
  void* C;
  long long sN, sH, sW;     // FPROP: element strides of an output pixel; WGRAD: sW = ldc
  int n_valid;              // valid output columns
  int m_valid;              // WGRAD: valid output rows (Cout)
  int out_fp32, relu, atomic;
  int qbox[3];              // per-warp store sub-box (w, h, n): a quarter (32 rows) of the tile
  int qoff[4][3];           // origin of quarter q inside the tile box (w, h, n)
  float* bn_sums;           // optional: fused BatchNorm statistics of the bf16 output (see bn.cu layout)
  int bn_C;                 // channels of that BN (== n_valid)
  int b_mn;                 // kind 0 only: B operand is MN-major (dgrad reads the filter un-transposed)
  int tma_store;            // epilogue stages tiles in smem and stores them with TMA (mapC)
  int colacc_bytes;         // kColAccBytes when bn_sums != nullptr, else 0
  // Tile index as mixed-radix digits (no integer division per tile in any warp role: a tile of a short-K
  // convolution lasts ~1.5 us and the single producer thread used to spend a third of that dividing).
  //   kind 0: (n_blk | tw | th | tn) with radices (tiles_n, tiles_w, tiles_h, -);
  //   kind 1: (n_blk | m_blk | split | 0) with radices (tiles_n, tiles_m, -, -).
  // step[] are the digits of gridDim.x, i.e. what one iteration of the persistent loop adds.
  int rad[3];
  int step[4];
  // Halo mode (3x3, stride 1, Cin == Cout == 64): ONE TMA box of (bw+2) x (bh+2) pixels is staged per tile and the
  // nine taps are nine shifted operand views of it (descriptor start = any 128-byte row, 8-row groups (bw+2)*128
  // bytes apart: the swizzle is a function of the absolute shared address, bench/probes/umma_shift_probe.cu);
  // the 9 x 64 x 64 filter stays resident in shared memory for the whole kernel.
  int halo;
  const float* bias;
  // deterministic mode (b200::deterministic()): sums are combined in a fixed order
  int det;                  // 1: no floating-point atomics anywhere in this launch
  int bn_reps;              // deterministic mode: CTA slots behind bn_sums
  int bn_used;              // copies of {sum, sumsq}[bn_C] the atomics are spread over (ticket word follows them;
                            // 1 outside deterministic mode: no ticket, no fold)
  long long split_stride;   // WGRAD, det: split s writes its partial tile to C + s * split_stride
};

struct TmaMaps {
  CUtensorMap a[kMaxMaps];
  CUtensorMap b[kMaxMaps];
  CUtensorMap c;
};

struct alignas(8) Barriers {
  uint64_t full[8];
  uint64_t empty[8];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint64_t wfull;
  uint32_t tmem_base;
  uint32_t pad;
};

struct TileCoord { int d0, d1, d2, d3; };
__device__ __forceinline__ void tile_init(const GemmParams& p, int t, TileCoord& c) {
  c.d0 = t % p.rad[0]; t /= p.rad[0];
  c.d1 = t % p.rad[1]; t /= p.rad[1];
  c.d2 = t % p.rad[2];
  c.d3 = t / p.rad[2];
}
__device__ __forceinline__ void tile_next(const GemmParams& p, TileCoord& c) {
  c.d0 += p.step[0];
  int k = c.d0 >= p.rad[0] ? 1 : 0;
  c.d0 -= k ? p.rad[0] : 0;
  c.d1 += p.step[1] + k;
  k = c.d1 >= p.rad[1] ? 1 : 0;
  c.d1 -= k ? p.rad[1] : 0;
  c.d2 += p.step[2] + k;
  k = c.d2 >= p.rad[2] ? 1 : 0;
  c.d2 -= k ? p.rad[2] : 0;
  c.d3 += p.step[3] + k;
}

__device__ __forceinline__ void store_chunk(const GemmParams& p, const float (&f)[32],
                                            long long row_off, int col0, bool row_ok) {
  if (!row_ok) return;
  const int nrem = p.n_valid - col0;
  if (nrem <= 0) return;
  if (p.out_fp32) {
    float* dst = reinterpret_cast<float*>(p.C) + row_off + col0;
    if (p.atomic) {
      if (nrem >= 32 && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
#pragma unroll
        for (int j = 0; j < 32; j += 4)
          asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + j), "f"(f[j]),
                       "f"(f[j + 1]), "f"(f[j + 2]), "f"(f[j + 3])
                       : "memory");
      } else {
        for (int j = 0; j < 32 && j < nrem; ++j) atomicAdd(dst + j, f[j]);
      }
    } else if (nrem >= 32 && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
#pragma unroll
      for (int j = 0; j < 32; j += 4)
        *reinterpret_cast<float4*>(dst + j) = make_float4(f[j], f[j + 1], f[j + 2], f[j + 3]);
    } else {
      for (int j = 0; j < 32 && j < nrem; ++j) dst[j] = f[j];
    }
  } else {
    __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(p.C) + row_off + col0;
    if (nrem >= 32 && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
#pragma unroll
      for (int j = 0; j < 32; j += 8) {
        uint4 u;
        u.x = pack_bf16x2(f[j], f[j + 1]);
        u.y = pack_bf16x2(f[j + 2], f[j + 3]);
        u.z = pack_bf16x2(f[j + 4], f[j + 5]);
        u.w = pack_bf16x2(f[j + 6], f[j + 7]);
        *reinterpret_cast<uint4*>(dst + j) = u;
      }
    } else {
      for (int j = 0; j < 32 && j < nrem; ++j) dst[j] = __float2bfloat16(f[j]);
    }
  }
}

__global__ void __launch_bounds__(kThreads, 1)
tc_gemm_kernel(const __grid_constant__ TmaMaps maps, const __grid_constant__ GemmParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // 1024-byte alignment is required by the 128-byte swizzle atoms; the kernel has no static
  // shared memory, so the dynamic window starts at the (aligned) base of the CTA's shared space.
  uint8_t* smem = smem_raw;
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  const int a_stage_bytes = kBM * kBK * 2;            // 16 KB
  const int b_stage_bytes = p.BN * kBK * 2;           // 8..32 KB
  constexpr int kHaloW = 9 * 8192;                     // resident filter: 9 taps x (64 x 64) bf16
  constexpr int kHaloStage = 23 * 1024;                // (8+2) x (16+2) pixels x 128 B = 23040, rounded to 1 KB
  const int stage_bytes = p.halo ? kHaloStage : a_stage_bytes + b_stage_bytes;
  if (p.halo) smem += kHaloW;                          // stages follow the resident filter
  uint8_t* stage_out = smem + p.stages * stage_bytes;          // 2 x 16 KB epilogue staging
  float* col_acc = reinterpret_cast<float*>(stage_out + 2 * kStageOutBytes);     // [4 quarters][2][256]
  Barriers* bars = reinterpret_cast<Barriers*>(stage_out + 2 * kStageOutBytes + p.colacc_bytes);

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);     // provably warp-uniform
  const int lane = threadIdx.x & 31;
  const int total_tiles = p.tiles_m * p.tiles_n * p.splits;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&maps.a[0]);
    tma_prefetch_desc(&maps.b[0]);
    if (p.tma_store) tma_prefetch_desc(&maps.c);
    for (int i = 0; i < p.stages; ++i) {
      mbar_init(&bars->full[i], 1);
      mbar_init(&bars->empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bars->tmem_full[i], 1);
      mbar_init(&bars->tmem_empty[i], 8);        // one arrive per epilogue warp
    }
    mbar_init(&bars->wfull, 1);
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc<512>(&bars->tmem_base);
  }
  // everything above touched only parameters, shared memory and TMEM: it may overlap the previous kernel's tail
  griddep_launch_dependents();
  griddep_wait();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    // The whole warp runs the loops (every value is warp-uniform) and one elected lane issues each TMA /
    // expect-tx; the loops carry coordinates incrementally (no division per k-block) and use pre-computed
    // 32-bit shared addresses, because the instruction count per TMA issue bounds the rate at which
    // short-K tiles can be fed.
    {
      const uint32_t smem0 = smem_u32(smem);
      const uint32_t full0 = smem_u32(&bars->full[0]);
      int stage = 0;
      uint32_t phase = 0;
      TileCoord tc;
      tile_init(p, blockIdx.x, tc);
      if (p.halo) {
        // resident filter: nine 64 x 64 blocks, loaded once per CTA
        const uint32_t sw0 = smem0 - kHaloW, wb = smem_u32(&bars->wfull);
        mbar_expect_tx_e(wb, kHaloW);
        for (int tap = 0; tap < 9; ++tap) {
          if (!p.b_mn) tma_load_2d_e(sw0 + tap * 8192, &maps.b[0], wb, tap * 64, 0);
          else tma_load_2d_e(sw0 + tap * 8192, &maps.b[0], wb, p.taps[tap].pad * 64, 0);
        }
        const uint32_t tx = (uint32_t)((p.bw + 2) * (p.bh + 2) * 128);
        for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, tile_next(p, tc)) {
          mbar_wait(&bars->empty[stage], phase ^ 1);
          const uint32_t fb = full0 + stage * 8;
          mbar_expect_tx_e(fb, tx);
          // one box with a one-pixel halo all around; outside the image TMA fills zeros (= the padding)
          tma_load_4d_e(smem0 + stage * stage_bytes, &maps.a[1], fb, 0, tc.d1 * p.bw - 1, tc.d2 * p.bh - 1, tc.d3);
          if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
      } else
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, tile_next(p, tc)) {
        const int n_blk = tc.d0;
        const int m_blk = tc.d1;                          // kind 1 only
        const int split = p.kind == 1 ? tc.d2 : 0;
        const int kb0 = split * p.kb_per_split;
        const int kb1 = min(p.kb_total, kb0 + p.kb_per_split);
        if (p.kind == 0) {
          const int tw = tc.d1, th = tc.d2, tn = tc.d3;
          const int w0 = tw * p.bw, h0 = th * p.bh, n0 = tn * p.bn;
          const int nb0 = n_blk * p.BN;
          int kcoord = 0;
#pragma unroll 1
          for (int tap = 0; tap < p.ntaps; ++tap) {
            const Tap tp = p.taps[tap];
            const CUtensorMap* ma = &maps.a[tp.map];
            const int wt = w0 + tp.dw, ht = h0 + tp.dh;
#pragma unroll 1
            for (int cb = 0; cb < p.cin_blocks; ++cb, kcoord += kBK) {
              mbar_wait(&bars->empty[stage], phase ^ 1);
              const uint32_t sa = smem0 + stage * stage_bytes;
              const uint32_t fb = full0 + stage * 8;
              mbar_expect_tx_e(fb, stage_bytes);
              tma_load_4d_e(sa, ma, fb, cb * kBK, wt, ht, n0);
              if (!p.b_mn) {
                tma_load_2d_e(sa + a_stage_bytes, &maps.b[0], fb, kcoord, nb0);
              } else {
                // MN-major B: rows = reduction index (this 64-block of it), columns = output
                // channels of this tile, fetched as 64-wide chunks; tp.pad = filter tap index.
                const int col0 = tp.pad * p.n_valid + nb0;
#pragma unroll 1
                for (int c = 0; c < (p.BN >> 6); ++c)
                  tma_load_2d_e(sa + a_stage_bytes + c * 8192, &maps.b[0], fb, col0 + c * 64, cb * kBK);
              }
              if (++stage == p.stages) { stage = 0; phase ^= 1; }
            }
          }
        } else {
          // N is the stacked (tap, input-channel) axis in 64-wide chunks; a tile covers BN/64
          // consecutive chunks, each with its own tap shift (and phase map for stride 2).
          const int nchunks = p.BN >> 6;
          const int gtotal = p.ntaps * p.cin_blocks;       // cin_blocks = 64-wide chunks per tap
          const CUtensorMap* mb[4];
          int cc0[4], cdw[4], cdh[4];
          int nvalid = 0;
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const int g = n_blk * nchunks + c;
            const bool ok = (c < nchunks) && (g < gtotal);
            const int tap = ok ? g / p.cin_blocks : 0;
            const Tap tp = p.taps[tap];
            mb[c] = &maps.b[tp.map];
            cc0[c] = ok ? (g - tap * p.cin_blocks) * 64 : 0;
            cdw[c] = tp.dw;
            cdh[c] = tp.dh;
            if (ok) nvalid = c + 1;
          }
          const uint32_t tx_bytes = a_stage_bytes + nvalid * 8192;
          const int m0 = m_blk * kBM;
          // pixel box origin of k-block kb0, then advanced incrementally
          int w0 = (kb0 % p.tiles_w) * p.bw;
          int h0 = ((kb0 / p.tiles_w) % p.tiles_h) * p.bh;
          int n0 = (kb0 / (p.tiles_w * p.tiles_h)) * p.bn;
          const int wend = p.tiles_w * p.bw, hend = p.tiles_h * p.bh;
#pragma unroll 1
          for (int kb = kb0; kb < kb1; ++kb) {
            mbar_wait(&bars->empty[stage], phase ^ 1);
            const uint32_t sa = smem0 + stage * stage_bytes;
            const uint32_t sb = sa + a_stage_bytes;
            const uint32_t fb = full0 + stage * 8;
            mbar_expect_tx_e(fb, tx_bytes);
            // A = dY^T : two 64-wide chunks of output channels, 64 pixels deep.
            tma_load_4d_e(sa, &maps.a[0], fb, m0, w0, h0, n0);
            tma_load_4d_e(sa + 8192, &maps.a[0], fb, m0 + 64, w0, h0, n0);
#pragma unroll
            for (int c = 0; c < 4; ++c)
              if (c < nvalid) tma_load_4d_e(sb + c * 8192, mb[c], fb, cc0[c], w0 + cdw[c], h0 + cdh[c], n0);
            w0 += p.bw;
            if (w0 >= wend) { w0 = 0; h0 += p.bh; if (h0 >= hend) { h0 = 0; n0 += p.bn; } }
            if (++stage == p.stages) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer
    // The whole warp runs the loop (every value below is warp-uniform) and ONE elected lane issues each
    // tcgen05.mma / commit: see umma_f16_elect.
    {
      const bool a_mn = p.kind == 1;
      const bool b_mn = p.kind == 1 || p.b_mn;
      const uint32_t idesc = umma_idesc_bf16(kBM, p.BN, a_mn, b_mn);
      const uint32_t lbo_a = a_mn ? 8192u : 16u, lbo_b = b_mn ? 8192u : 16u;
      const uint32_t kstep_a = a_mn ? 2048u : 32u, kstep_b = b_mn ? 2048u : 32u;
      const uint32_t smem0_mma = smem_u32(smem);
      const uint64_t desc_a0 = umma_desc_sw128(0, lbo_a, p.halo ? (uint32_t)(p.bw + 2) * 128u : 1024u);
      const uint64_t desc_b0 = umma_desc_sw128(0, lbo_b, 1024);
      uint32_t halo_aoff[9];                   // (row offset of tap's view inside the staged patch) >> 4
#pragma unroll
      for (int tap = 0; tap < 9; ++tap)
        halo_aoff[tap] = (uint32_t)((p.taps[tap].dh + 1) * (p.bw + 2) + (p.taps[tap].dw + 1)) * 8u;
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      TileCoord tc;
      tile_init(p, blockIdx.x, tc);
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, tile_next(p, tc)) {
        const int split = p.kind == 1 ? tc.d2 : 0;
        const int kb0 = split * p.kb_per_split;
        const int kb1 = min(p.kb_total, kb0 + p.kb_per_split);
        mbar_wait(&bars->tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * p.BN;
        if (p.halo) {
          if (t == (int)blockIdx.x) { mbar_wait(&bars->wfull, 0); tc_fence_after(); }
          mbar_wait(&bars->full[stage], phase);
          tc_fence_after();
          const uint32_t sa = (smem0_mma + stage * stage_bytes) >> 4;
          const uint32_t sw0 = (smem0_mma - kHaloW) >> 4;
#pragma unroll
          for (int tap = 0; tap < 9; ++tap) {
            // the tap's operand = the staged patch seen from pixel (dh + 1, dw + 1)
#pragma unroll
            for (int k = 0; k < kBK / 16; ++k)
              umma_f16_elect(tmem_d, desc_a0 + (sa + halo_aoff[tap] + k * 2), desc_b0 + (sw0 + tap * 512 + k * (kstep_b >> 4)),
                       idesc, (tap > 0 || k > 0) ? 1u : 0u);
          }
          umma_commit_elect(&bars->empty[stage]);
          if (++stage == p.stages) { stage = 0; phase ^= 1; }
        } else
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&bars->full[stage], phase);
          tc_fence_after();
          // descriptor = constant high part + (shared address >> 4): one add per operand per instruction
          const uint32_t sa = (smem0_mma + stage * stage_bytes) >> 4;
          const uint32_t sb = sa + (a_stage_bytes >> 4);
#pragma unroll
          for (int k = 0; k < kBK / 16; ++k)
            umma_f16_elect(tmem_d, desc_a0 + (sa + k * (kstep_a >> 4)), desc_b0 + (sb + k * (kstep_b >> 4)), idesc,
                     (kb > kb0 || k > 0) ? 1u : 0u);
          umma_commit_elect(&bars->empty[stage]);
          if (++stage == p.stages) { stage = 0; phase ^= 1; }
        }
        umma_commit_elect(&bars->tmem_full[acc]);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------ epilogue (warps 2..9)
    // Two warps per TMEM lane quarter; each warp is an independent pipeline over its column
    // slabs: tcgen05.ld -> (bias / ReLU) -> swizzled 4 KB smem tile -> its own TMA store of a
    // 32-row sub-box.  No CTA-wide barrier is needed anywhere in the epilogue.
    const int q = warp & 3;                // TMEM lane quarter this warp may touch
    const int pi = (warp - 2) >> 2;        // which of the two warps of the quarter
    const int row = q * 32 + lane;
    uint8_t* buf = stage_out + (warp - 2) * 4096;
    int acc = 0;
    uint32_t acc_phase = 0;
    // which copy of the statistics this CTA adds into (one modulo per kernel, not per tile)
    const size_t stat_copy_off =
        p.bn_sums != nullptr ? (size_t)(p.det ? blockIdx.x : blockIdx.x % p.bn_used) * 2 * p.bn_C : 0;
    TileCoord tc;
    tile_init(p, blockIdx.x, tc);
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, tile_next(p, tc)) {
      const int n_blk = tc.d0;
      const int m_blk = tc.d1;              // kind 1 only
      const int split = p.kind == 1 ? tc.d2 : 0;
      const int col_base = n_blk * p.BN;   // WGRAD: taps are consecutive, tap * Cin + ct * BN
      int tw = 0, th = 0, tn = 0;
      if (p.kind == 0) { tw = tc.d1; th = tc.d2; tn = tc.d3; }
      mbar_wait(&bars->tmem_full[acc], acc_phase);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * p.BN;
      if (p.tma_store) {
        const int cols_per_slab = p.out_fp32 ? 32 : 64;
        const int nslabs = p.out_fp32 ? (p.BN >> 5) : (p.BN >> 6);        // (no division in the tile loop)
        int c1, c2, c3;                    // origin of this warp's 32-row sub-box in the output map
        if (p.kind == 0) {
          c1 = tw * p.bw + p.qoff[q][0]; c2 = th * p.bh + p.qoff[q][1]; c3 = tn * p.bn + p.qoff[q][2];
        } else {
          c1 = m_blk * kBM + q * 32; c2 = p.split_stride ? split : 0; c3 = 0;
        }
        uint8_t* my_row = buf + lane * 128;
        const int sw = lane & 7;
        // halo tiles may hang over the image edge: those rows are clipped by the TMA store, but they must not
        // enter the fused statistics (their 3x3 windows still touch valid pixels), so they are staged as zeros
        // (halo tiles are always 8 pixels wide: shifts, not divisions, once per tile per thread)
        const bool zero_row = p.halo && (th * p.bh + (row >> 3) >= p.H || tw * 8 + (row & 7) >= p.W);
        const bool plain = (p.bias == nullptr) && !p.relu;       // uniform across the CTA
        for (int sl = pi; sl < nslabs; sl += 2) {
          const int col0 = col_base + sl * cols_per_slab;
          if (col0 >= p.n_valid) break;    // uniform: whole slab is outside the tensor
          const bool full = (col0 + cols_per_slab <= p.n_valid);
          if (p.out_fp32) {
            uint32_t v[32];
            tmem_ld_32x32(taddr + sl * 32, v);
            // the previous TMA store of this warp must have drained the staging tile
            if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            __syncwarp();
            tmem_ld_wait();
            if (plain) {
#pragma unroll
              for (int j = 0; j < 32; j += 4)
                *reinterpret_cast<uint4*>(my_row + (((j >> 2) ^ sw) << 4)) =
                    make_uint4(v[j], v[j + 1], v[j + 2], v[j + 3]);
            } else {
#pragma unroll
              for (int j = 0; j < 32; j += 4) {
                float f[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) f[e] = __uint_as_float(v[j + e]);
                if (p.bias != nullptr) {
                  if (full) {
                    const float4 bv = __ldg(reinterpret_cast<const float4*>(p.bias + col0 + j));
                    f[0] += bv.x; f[1] += bv.y; f[2] += bv.z; f[3] += bv.w;
                  } else {
#pragma unroll
                    for (int e = 0; e < 4; ++e)
                      if (col0 + j + e < p.n_valid) f[e] += __ldg(p.bias + col0 + j + e);
                  }
                }
                if (p.relu) {
#pragma unroll
                  for (int e = 0; e < 4; ++e) f[e] = fmaxf(f[e], 0.f);
                }
                *reinterpret_cast<float4*>(my_row + (((j >> 2) ^ sw) << 4)) =
                    make_float4(f[0], f[1], f[2], f[3]);
              }
            }
          } else {
            uint32_t v0[32], v1[32];
            tmem_ld_32x32(taddr + sl * 64, v0);            // both halves in flight before the wait
            tmem_ld_32x32(taddr + sl * 64 + 32, v1);
            // the previous TMA store of this warp must have drained the staging tile
            if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            __syncwarp();
            tmem_ld_wait();
            if (zero_row) {                    // (halo tiles only, edge rows only: skipped by whole warps otherwise)
#pragma unroll
              for (int j = 0; j < 32; ++j) { v0[j] = 0u; v1[j] = 0u; }
            }
#pragma unroll
            for (int half = 0; half < 2; ++half) {
              const uint32_t* v = half ? v1 : v0;
              const int cb = col0 + half * 32;
#pragma unroll
              for (int j = 0; j < 32; j += 8) {
                float f[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) f[e] = __uint_as_float(v[j + e]);
                if (!plain) {
                  if (p.bias != nullptr) {
                    if (full) {
                      const float4 b0 = __ldg(reinterpret_cast<const float4*>(p.bias + cb + j));
                      const float4 b1 = __ldg(reinterpret_cast<const float4*>(p.bias + cb + j + 4));
                      f[0] += b0.x; f[1] += b0.y; f[2] += b0.z; f[3] += b0.w;
                      f[4] += b1.x; f[5] += b1.y; f[6] += b1.z; f[7] += b1.w;
                    } else {
#pragma unroll
                      for (int e = 0; e < 8; ++e)
                        if (cb + j + e < p.n_valid) f[e] += __ldg(p.bias + cb + j + e);
                    }
                  }
                  if (p.relu) {
#pragma unroll
                    for (int e = 0; e < 8; ++e) f[e] = fmaxf(f[e], 0.f);
                  }
                }
                const int chunk = half * 4 + (j >> 3);
                *reinterpret_cast<uint4*>(my_row + ((chunk ^ sw) << 4)) = pack8(f);
              }
            }
          }
          fence_proxy_async();             // generic-proxy smem writes -> visible to TMA
          __syncwarp();
          if (p.bn_sums != nullptr && !p.out_fp32) {
            // Fused BatchNorm statistics of this warp's 32 rows x 64 columns, straight from the staged (already
            // bf16-rounded) tile; rows outside the tensor hold exact zeros (TMA zero fill), so they do not
            // count.  Lane l reads the 16-byte chunk (l & 7) = 8 columns of rows (l >> 3) * 8 .. + 7 with
            // LDS.128 (conflict-free under the 128-byte swizzle) and accumulates with packed fp32x2 math;
            // the four row groups are then combined with two shuffle rounds.
            const int c8 = lane & 7, rg = lane >> 3;
            float2 sm2[4], sq2[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) { sm2[k] = make_float2(0.f, 0.f); sq2[k] = make_float2(0.f, 0.f); }
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
#pragma unroll
            for (int off = 8; off <= 16; off <<= 1) {
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                sm2[k].x += __shfl_xor_sync(0xffffffffu, sm2[k].x, off);
                sm2[k].y += __shfl_xor_sync(0xffffffffu, sm2[k].y, off);
                sq2[k].x += __shfl_xor_sync(0xffffffffu, sq2[k].x, off);
                sq2[k].y += __shfl_xor_sync(0xffffffffu, sq2[k].y, off);
              }
            }
            if (rg == 0) {
              // every (quarter, column) pair has exactly one writer per tile: plain stores, fixed order later
              float* accb = col_acc + q * 512 + sl * 64 + c8 * 8;
              *reinterpret_cast<float4*>(accb) = make_float4(sm2[0].x, sm2[0].y, sm2[1].x, sm2[1].y);
              *reinterpret_cast<float4*>(accb + 4) = make_float4(sm2[2].x, sm2[2].y, sm2[3].x, sm2[3].y);
              *reinterpret_cast<float4*>(accb + 256) = make_float4(sq2[0].x, sq2[0].y, sq2[1].x, sq2[1].y);
              *reinterpret_cast<float4*>(accb + 260) = make_float4(sq2[2].x, sq2[2].y, sq2[3].x, sq2[3].y);
            }
          }
          if (lane == 0) {
            if (p.atomic)
              asm volatile(
                  "cp.reduce.async.bulk.tensor.4d.global.shared::cta.add.tile.bulk_group"
                  " [%0, {%2, %3, %4, %5}], [%1];" ::"l"(reinterpret_cast<uint64_t>(&maps.c)),
                  "r"(smem_u32(buf)), "r"(col0), "r"(c1), "r"(c2), "r"(c3)
                  : "memory");
            else
              asm volatile(
                  "cp.async.bulk.tensor.4d.global.shared::cta.tile.bulk_group"
                  " [%0, {%2, %3, %4, %5}], [%1];" ::"l"(reinterpret_cast<uint64_t>(&maps.c)),
                  "r"(smem_u32(buf)), "r"(col0), "r"(c1), "r"(c2), "r"(c3)
                  : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
        }
      } else {
        // direct-store fallback (outputs whose pitch is not a multiple of 16 bytes, odd boxes)
        bool row_ok;
        long long row_off;
        if (p.kind == 0) {
          const int dw = row % p.bw;
          const int dh = (row / p.bw) % p.bh;
          const int dn = row / (p.bw * p.bh);
          const int w = tw * p.bw + dw, h = th * p.bh + dh, n = tn * p.bn + dn;
          row_ok = (w < p.W) && (h < p.H) && (n < p.NB);
          row_off = n * p.sN + h * p.sH + w * p.sW;
        } else {
          const int r = m_blk * kBM + row;
          row_ok = r < p.m_valid;
          row_off = static_cast<long long>(r) * p.sW + split * p.split_stride;
        }
        for (int c = pi; c < p.BN / 32; c += 2) {
          uint32_t v[32];
          tmem_ld_32x32(taddr + c * 32, v);
          tmem_ld_wait();
          float f[32];
          const int col0 = col_base + c * 32;
#pragma unroll
          for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
          if (p.bias != nullptr) {
            for (int j = 0; j < 32; ++j)
              if (col0 + j < p.n_valid) f[j] += __ldg(p.bias + col0 + j);
          }
          if (p.relu) {
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.f);
          }
          store_chunk(p, f, row_off, col0, row_ok);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars->tmem_empty[acc]);
      if (p.bn_sums != nullptr) {
        // all eight epilogue warps have added their rows of this tile: flush the tile's column
        // sums to global memory (spread over the replicas) and clear the buffer for tile t + 2
        asm volatile("bar.sync 2, 256;" ::: "memory");
        float* accb = col_acc;
        // det: this CTA owns copy blockIdx.x and column c is always handled by the same thread, so a
        // plain read-modify-write in tile order replaces the atomics
        float* dstb = p.bn_sums + stat_copy_off;
        const int e = (warp - 2) * 32 + lane;          // 0..255: one column per thread
        if (e < p.BN) {
          const int c = col_base + e;
          if (c < p.n_valid) {
            // quarters in a fixed order: the tile's column sums do not depend on warp timing
            const float sv = ((accb[e] + accb[512 + e]) + accb[1024 + e]) + accb[1536 + e];
            const float qv = ((accb[256 + e] + accb[768 + e]) + accb[1280 + e]) + accb[1792 + e];
            if (p.det) { dstb[c] += sv; dstb[p.bn_C + c] += qv; }
            else { atomicAdd(dstb + c, sv); atomicAdd(dstb + p.bn_C + c, qv); }
          }
        }
        asm volatile("bar.sync 2, 256;" ::: "memory");      // all read before the next tile overwrites
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (p.tma_store && lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    if (p.bn_sums != nullptr && (p.det || p.bn_used > 1)) {
      // the last CTA to get here folds the replicas into copy 0 (what bn_apply reads)
      __threadfence();
      asm volatile("bar.sync 2, 256;" ::: "memory");
      uint32_t mine = 0, fold_last = 0;
      if (threadIdx.x == 64) {
        unsigned int* ticket = reinterpret_cast<unsigned int*>(p.bn_sums + (size_t)p.bn_used * 2 * p.bn_C);
        mine = (atomicAdd(ticket, 1u) == gridDim.x - 1) ? 1u : 0u;
      }
      // broadcast the verdict through the barrier's OR-reduction (no static shared memory: the
      // kernel asks for the full 227 KB as dynamic shared memory)
      asm volatile(
          "{\n.reg .pred pin, pout;\nsetp.ne.u32 pin, %1, 0;\n"
          "bar.red.or.pred pout, 2, 256, pin;\nselp.u32 %0, 1, 0, pout;\n}\n"
          : "=r"(fold_last) : "r"(mine) : "memory");
      if (fold_last) {
        __threadfence();
        fold_sum_copies(p.bn_sums, 2 * p.bn_C, p.det ? min(p.bn_reps, (int)gridDim.x) : p.bn_used,
                        (int)threadIdx.x - 64, 256);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

// ------------------------------------------------------------------------------------------
// Host side: tensor-map construction and launch.

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &qres) ==
            cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(sym);
  }
  return fn;
}

// dims: innermost first (C, W, H, N); strides in ELEMENTS for W, H, N.
static int make_map4(CUtensorMap* m, const void* base, const long long dims[4],
                     const long long strides_el[3], const int box[4]) {
  EncodeTiledFn enc = get_encode();
  if (!enc) return -1;
  cuuint64_t gd[4] = {(cuuint64_t)dims[0], (cuuint64_t)dims[1], (cuuint64_t)dims[2],
                      (cuuint64_t)dims[3]};
  cuuint64_t gs[3] = {(cuuint64_t)strides_el[0] * 2, (cuuint64_t)strides_el[1] * 2,
                      (cuuint64_t)strides_el[2] * 2};
  cuuint32_t bx[4] = {(cuuint32_t)box[0], (cuuint32_t)box[1], (cuuint32_t)box[2],
                      (cuuint32_t)box[3]};
  cuuint32_t es[4] = {1, 1, 1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), gd, gs, bx, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : -(int)r - 1000;
}

// Output tensor map (TMA store): bf16 (64-column slabs) or fp32 (32-column slabs).
static int make_map_out(CUtensorMap* m, const void* base, const long long dims[4],
                        const long long strides_el[3], const int box[4], bool f32) {
  EncodeTiledFn enc = get_encode();
  if (!enc) return -1;
  const int esz = f32 ? 4 : 2;
  cuuint64_t gd[4] = {(cuuint64_t)dims[0], (cuuint64_t)dims[1], (cuuint64_t)dims[2],
                      (cuuint64_t)dims[3]};
  cuuint64_t gs[3] = {(cuuint64_t)strides_el[0] * esz, (cuuint64_t)strides_el[1] * esz,
                      (cuuint64_t)strides_el[2] * esz};
  cuuint32_t bx[4] = {(cuuint32_t)(f32 ? 32 : 64), (cuuint32_t)box[1], (cuuint32_t)box[2],
                      (cuuint32_t)box[3]};
  cuuint32_t es[4] = {1, 1, 1, 1};
  CUresult r = enc(m, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4,
                   const_cast<void*>(base), gd, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : -(int)r - 1000;
}
static bool out_map_ok(const void* base, const long long strides_el[3], bool f32) {
  const int esz = f32 ? 4 : 2;
  if (reinterpret_cast<uintptr_t>(base) & 15) return false;
  for (int i = 0; i < 3; ++i)
    if ((strides_el[i] * esz) % 16) return false;
  return true;
}

static int make_map2(CUtensorMap* m, const void* base, long long inner, long long outer,
                     long long ld_el, int box_inner, int box_outer) {
  EncodeTiledFn enc = get_encode();
  if (!enc) return -1;
  cuuint64_t gd[2] = {(cuuint64_t)inner, (cuuint64_t)outer};
  cuuint64_t gs[1] = {(cuuint64_t)ld_el * 2};
  cuuint32_t bx[2] = {(cuuint32_t)box_inner, (cuuint32_t)box_outer};
  cuuint32_t es[2] = {1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gd, gs, bx, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : -(int)r - 1000;
}

static int num_sms() {
  static int cached[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) return 148;
  if (cached[dev] == 0) {
    cudaDeviceGetAttribute(&cached[dev], cudaDevAttrMultiProcessorCount, dev);
    if (cached[dev] <= 0) cached[dev] = 148;
  }
  return cached[dev];
}

static int launch(const CUtensorMap mapsA[kMaxMaps], const CUtensorMap mapsB[kMaxMaps],
                  const CUtensorMap* mapC, GemmParams& p, cudaStream_t stream) {
  const int stage_bytes = p.halo ? 23 * 1024 : kBM * kBK * 2 + p.BN * kBK * 2;
  p.colacc_bytes = p.bn_sums != nullptr ? kColAccBytes : 0;
  const int halo_w = p.halo ? 9 * 8192 : 0;
  const int fixed = 2 * kStageOutBytes + p.colacc_bytes + (int)sizeof(Barriers) + 64 + halo_w;
  int stages = (227 * 1024 - fixed) / stage_bytes;
  if (stages > 8) stages = 8;
  if (stages < 2) return -2;
  p.stages = stages;
  p.tma_store = mapC != nullptr ? 1 : 0;
  p.det = deterministic();
  p.bn_reps = bn_replicas();
  p.bn_used = p.bn_sums != nullptr ? bn_replicas_used(p.bn_C) : 1;

  const int smem = stages * stage_bytes + fixed;
  // function attributes are per device: opt in to the large dynamic smem once on each
  static bool configured[64] = {false};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !configured[dev]) {
    cudaError_t e = cudaFuncSetAttribute(tc_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         227 * 1024);
    if (e != cudaSuccess) return (int)e;
    configured[dev] = true;
  }
  const int total = p.tiles_m * p.tiles_n * p.splits;
  int grid = total < num_sms() ? total : num_sms();
  if (p.det && p.bn_sums != nullptr && grid > p.bn_reps) grid = p.bn_reps;   // one statistics slot per CTA
  if (grid <= 0) return 0;
  p.rad[0] = p.tiles_n;
  p.rad[1] = p.kind == 0 ? p.tiles_w : p.tiles_m;
  p.rad[2] = p.kind == 0 ? p.tiles_h : (1 << 30);
  {
    int g = grid;
    p.step[0] = g % p.rad[0]; g /= p.rad[0];
    p.step[1] = g % p.rad[1]; g /= p.rad[1];
    p.step[2] = g % p.rad[2];
    p.step[3] = g / p.rad[2];
  }
  TmaMaps maps;
  for (int i = 0; i < kMaxMaps; ++i) { maps.a[i] = mapsA[i]; maps.b[i] = mapsB[i]; }
  maps.c = mapC != nullptr ? *mapC : mapsA[0];
  cudaError_t le = launch_pdl(kPdlGemm, tc_gemm_kernel, dim3(grid), dim3(kThreads), (size_t)smem, stream, maps, p);
  if (le != cudaSuccess) return (int)le;
  return (int)cudaGetLastError();
}

// Split the 128-row tile box (bw, bh, bn) into four 32-row sub-boxes, one per TMEM lane quarter
// (rows are ordered w fastest, n slowest, so the split peels the slowest dimensions first).
static bool quarter_box(GemmParams& p) {
  int dims[3] = {p.bw, p.bh, p.bn};          // w, h, n
  int f[3] = {1, 1, 1};
  int need = 4;
  for (int d = 2; d >= 0 && need > 1; --d) {
    int g = 1;
    while (g * 2 <= need && dims[d] % (g * 2) == 0) g *= 2;
    f[d] = g;
    need /= g;
    if (g != dims[d] && need > 1) return false;   // a faster dimension may only be split once
  }                                               // every slower one is fully split
  if (need != 1) return false;
  for (int d = 0; d < 3; ++d) p.qbox[d] = dims[d] / f[d];
  for (int q = 0; q < 4; ++q) {
    const int iw = q % f[0], ih = (q / f[0]) % f[1], in = q / (f[0] * f[1]);
    p.qoff[q][0] = iw * p.qbox[0];
    p.qoff[q][1] = ih * p.qbox[1];
    p.qoff[q][2] = in * p.qbox[2];
  }
  return true;
}

static int pick_bn(int n) {
  if (n % 256 == 0 || n > 512) return 256;
  if (n % 128 == 0 || n > 128) return 128;
  return 64;
}

// Pixel box (bw, bh, bn) with bw | W, bh | H and bw*bh*bn == target (128 or 64); prefers the
// widest spatial footprint so that neighbouring taps re-hit the same L2 lines.
static void pick_box(int W, int H, int target, int* bw, int* bh, int* bn) {
  int best_w = 1, best_h = 1;
  for (int w = 1; w <= W && w <= target; ++w) {
    if (W % w) continue;
    for (int h = 1; h <= H && w * h <= target; ++h) {
      if (H % h) continue;
      if (target % (w * h)) continue;
      if (w * h > best_w * best_h) { best_w = w; best_h = h; }
    }
  }
  *bw = best_w;
  *bh = best_h;
  *bn = target / (best_w * best_h);
}

// Deterministic split-K: every split writes its partial [M, ld_s] tile set to a scratch slab and this
// kernel adds the slabs in split order (out = (out +) sum_s part[s]).
__global__ void splitk_fold_kernel(const float* __restrict__ part, float* __restrict__ out, int M, int N,
                                   int ld_s, long long ldc, int splits, long long split_stride,
                                   int accumulate) {
  const long long total = (long long)M * N;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int m = (int)(i / N), n = (int)(i - (long long)m * N);
    const float* src = part + (long long)m * ld_s + n;
    float acc = accumulate ? out[m * ldc + n] : 0.f;
    for (int sidx = 0; sidx < splits; ++sidx) acc += src[sidx * split_stride];
    out[m * ldc + n] = acc;
  }
}

// Stream-ordered scratch for the deterministic split-K path (allocated and freed on the launch
// stream, so concurrent streams never share it and a graph capture records both nodes).
static int det_split_begin(GemmParams& p, int M, int N, cudaStream_t st, float** part, int* ld_s) {
  *ld_s = (N + 3) & ~3;
  const size_t bytes = sizeof(float) * (size_t)p.splits * M * (*ld_s);
  cudaError_t e = cudaMallocAsync(reinterpret_cast<void**>(part), bytes, st);
  if (e != cudaSuccess) return (int)e;
  p.C = *part; p.sW = *ld_s; p.split_stride = (long long)M * (*ld_s); p.atomic = 0;
  return 0;
}
static int det_split_end(const GemmParams& p, float* part, int ld_s, float* out, int M, int N,
                         long long ldc, int accumulate, cudaStream_t st) {
  const long long total = (long long)M * N;
  int blocks = (int)((total + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  splitk_fold_kernel<<<blocks, 256, 0, st>>>(part, out, M, N, ld_s, ldc, p.splits, p.split_stride,
                                             accumulate);
  cudaError_t e = cudaGetLastError();
  cudaError_t f = cudaFreeAsync(part, st);
  return (int)(e != cudaSuccess ? e : f);
}

// Halo mode applies to the 3x3 stride-1 64 -> 64 convolutions (ResNet layer1): B200DT_HALO=0 switches it off.
static bool halo_shape(int R, int stride, int Cin, int Cout) {
  static int enabled = -1;
  if (enabled < 0) {
    const char* e = getenv("B200DT_HALO");
    enabled = (e != nullptr && e[0] == '0') ? 0 : 1;
  }
  return enabled && R == 3 && stride == 1 && Cin == 64 && Cout == 64;
}
static void halo_tiling(GemmParams& p, int Wo, int Ho, int NB) {
  p.halo = 1;
  p.bw = 8; p.bh = 16; p.bn = 1;                 // 8-pixel row groups: one UMMA core-matrix group per output row
  p.tiles_w = (Wo + 7) / 8;
  p.tiles_h = (Ho + 15) / 16;                    // edge tiles hang over the image: zero fill in, clipped stores out
  p.tiles_m = p.tiles_w * p.tiles_h * NB;
}

}  // namespace b200

using namespace b200;

// C[M,N] (bf16 or fp32) = A[M,K] * B[N,K]^T (+ bias[N]) (ReLU)     -- all row-major, bf16 inputs.
static int clear_bn_sums(float* bn_sums, int C, cudaStream_t st) {
  if (bn_sums == nullptr || sums_prezeroed()) return 0;
  return (int)cudaMemsetAsync(bn_sums, 0, sizeof(float) * ((size_t)2 * C * bn_replicas_used(C) + 1), st);
}

extern "C" int b200dt_gemm_tn_bf16(const void* A, const void* B, void* C, int M, int N, int K,
                                   int lda, int ldb, int ldc, const float* bias, int relu,
                                   int out_fp32, float* bn_sums, void* stream) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  if ((lda % 8) || (ldb % 8) || (K % 8)) return -3;
  GemmParams p{};
  p.bn_sums = bn_sums; p.bn_C = N;
  p.kind = 0;
  p.BN = pick_bn(N);
  p.tiles_m = (M + kBM - 1) / kBM;
  p.tiles_n = (N + p.BN - 1) / p.BN;
  p.splits = 1;
  p.kb_total = (K + kBK - 1) / kBK;
  p.kb_per_split = p.kb_total;
  p.bw = 128; p.bh = 1; p.bn = 1;
  p.tiles_w = p.tiles_m; p.tiles_h = 1;
  p.W = M; p.H = 1; p.NB = 1;
  p.cin_blocks = p.kb_total;
  p.ntaps = 1;
  p.taps[0] = Tap{0, 0, 0, 0};
  p.C = C; p.sN = 0; p.sH = 0; p.sW = ldc;
  p.n_valid = N; p.m_valid = M;
  p.out_fp32 = out_fp32; p.relu = relu; p.atomic = 0; p.bias = bias;
  CUtensorMap ma[kMaxMaps], mb[kMaxMaps];
  long long dims[4] = {K, M, 1, 1};
  long long str[3] = {lda, (long long)lda * M, (long long)lda * M};
  int box[4] = {kBK, 128, 1, 1};
  int r = make_map4(&ma[0], A, dims, str, box);
  if (r) return r;
  ma[1] = ma[2] = ma[3] = ma[0];
  r = make_map2(&mb[0], B, K, N, ldb, kBK, p.BN);
  if (r) return r;
  mb[1] = mb[2] = mb[3] = mb[0];
  CUtensorMap mc;
  long long dimsC[4] = {N, M, 1, 1};
  long long strC[3] = {ldc, (long long)ldc * M, (long long)ldc * M};
  const bool tma_out = out_map_ok(C, strC, out_fp32 != 0) && quarter_box(p);
  if (tma_out) {
    int qb[4] = {0, p.qbox[0], p.qbox[1], p.qbox[2]};
    r = make_map_out(&mc, C, dimsC, strC, qb, out_fp32 != 0);
    if (r) return r;
  }
  if (bn_sums != nullptr) {
    if (!tma_out || out_fp32) return -3;        // fused statistics live in the TMA-store epilogue
    r = clear_bn_sums(bn_sums, N, (cudaStream_t)stream);
    if (r) return r;
  }
  return launch(ma, mb, tma_out ? &mc : nullptr, p, (cudaStream_t)stream);
}

// C[M,N] fp32 (+)= A[Kr,M]^T * B[Kr,N]   (reduction over the Kr rows; both operands row-major
// with the reduction index outermost, i.e. MN-major).  Used for weight gradients:
// dW[Cout,Cin] = dY[rows,Cout]^T * X[rows,Cin].  C must be zero-filled by the caller when
// `splits` > 1 (the kernel accumulates with red.add); pass splits = 0 to let the library choose,
// in which case C is cleared here.
extern "C" int b200dt_gemm_nt_reduce_bf16(const void* A, const void* B, float* C, int M, int N,
                                          int Kr, int lda, int ldb, int ldc, int splits,
                                          void* stream) {
  if (M <= 0 || N <= 0 || Kr <= 0) return 0;
  if ((lda % 8) || (ldb % 8)) return -3;
  GemmParams p{};
  p.kind = 1;
  p.BN = N >= 256 ? 256 : (N > 64 ? 128 : 64);
  p.tiles_m = (M + kBM - 1) / kBM;
  p.tiles_n = (N + p.BN - 1) / p.BN;
  p.kb_total = (Kr + 63) / 64;
  const int out_tiles = p.tiles_m * p.tiles_n;
  bool clear = false;
  if (splits <= 0) {
    splits = (2 * num_sms() + out_tiles - 1) / out_tiles;
    if (splits > p.kb_total) splits = p.kb_total;
    if (splits < 1) splits = 1;
    clear = splits > 1;
  }
  p.kb_per_split = (p.kb_total + splits - 1) / splits;
  p.splits = (p.kb_total + p.kb_per_split - 1) / p.kb_per_split;
  p.bw = 64; p.bh = 1; p.bn = 1;
  p.tiles_w = p.kb_total; p.tiles_h = 1;
  p.W = Kr; p.H = 1; p.NB = 1;
  p.cin_blocks = (N + 63) / 64;
  p.ntaps = 1;
  p.taps[0] = Tap{0, 0, 0, 0};
  p.C = C; p.sW = ldc; p.n_valid = N; p.m_valid = M;
  p.out_fp32 = 1; p.relu = 0; p.atomic = p.splits > 1; p.bias = nullptr;
  const bool det_split = deterministic() && p.splits > 1;
  float* part = nullptr;
  int ld_s = 0;
  if (det_split) {
    int e = det_split_begin(p, M, N, (cudaStream_t)stream, &part, &ld_s);
    if (e) return e;
  } else if (clear) {
    cudaError_t e = cudaMemsetAsync(C, 0, sizeof(float) * (size_t)ldc * M, (cudaStream_t)stream);
    if (e != cudaSuccess) return (int)e;
  }
  CUtensorMap ma[kMaxMaps], mb[kMaxMaps];
  long long dimsA[4] = {M, Kr, 1, 1};
  long long strA[3] = {lda, (long long)lda * Kr, (long long)lda * Kr};
  int box[4] = {64, 64, 1, 1};
  int r = make_map4(&ma[0], A, dimsA, strA, box);
  if (r) return r;
  ma[1] = ma[2] = ma[3] = ma[0];
  long long dimsB[4] = {N, Kr, 1, 1};
  long long strB[3] = {ldb, (long long)ldb * Kr, (long long)ldb * Kr};
  r = make_map4(&mb[0], B, dimsB, strB, box);
  if (r) return r;
  mb[1] = mb[2] = mb[3] = mb[0];
  CUtensorMap mc;
  const long long ldo = det_split ? ld_s : ldc;
  long long dimsC[4] = {N, M, det_split ? p.splits : 1, 1};
  long long strC[3] = {ldo, ldo * M, ldo * M * (det_split ? p.splits : 1)};
  int boxC[4] = {32, 32, 1, 1};
  const bool tma_out = out_map_ok(p.C, strC, true);
  if (tma_out) {
    r = make_map_out(&mc, p.C, dimsC, strC, boxC, true);
    if (r) return r;
  }
  r = launch(ma, mb, tma_out ? &mc : nullptr, p, (cudaStream_t)stream);
  if (det_split) {
    // caller-chosen splits accumulate into C (documented contract); library-chosen ones overwrite
    const int r2 = det_split_end(p, part, ld_s, C, M, N, ldc, clear ? 0 : 1, (cudaStream_t)stream);
    if (r == 0) r = r2;
  }
  return r;
}

// NHWC conv2d forward as implicit GEMM.  x: [NB, H, W, Cin] bf16, w: [Cout, R, S, Cin] bf16,
// y: [NB, Ho, Wo, Cout] bf16.  Square filters R = S in {1, 3}, pad = R/2, stride in {1, 2},
// Cin % 64 == 0.  For stride 2, H and W must be even.
extern "C" int b200dt_conv2d_fprop_bf16(const void* x, const void* w, void* y, int NB, int H, int W,
                                        int Cin, int Cout, int R, int stride, int relu,
                                        float* bn_sums, void* stream) {
  if (!(R == 1 || R == 3) || !(stride == 1 || stride == 2) || (Cin % 64) || (Cout % 8)) return -3;
  if (stride == 2 && ((H | W) & 1)) return -3;
  const int Ho = H / stride, Wo = W / stride;
  GemmParams p{};
  p.kind = 0;
  p.BN = pick_bn(Cout);
  pick_box(Wo, Ho, 128, &p.bw, &p.bh, &p.bn);
  p.tiles_w = Wo / p.bw;
  p.tiles_h = Ho / p.bh;
  const int tiles_nb = (NB + p.bn - 1) / p.bn;
  p.tiles_m = p.tiles_w * p.tiles_h * tiles_nb;
  const bool halo = halo_shape(R, stride, Cin, Cout);
  if (halo) halo_tiling(p, Wo, Ho, NB);
  p.tiles_n = (Cout + p.BN - 1) / p.BN;
  p.splits = 1;
  p.cin_blocks = Cin / 64;
  p.ntaps = R * R;
  p.kb_total = p.ntaps * p.cin_blocks;
  p.kb_per_split = p.kb_total;
  p.W = Wo; p.H = Ho; p.NB = NB;
  p.C = y;
  p.sW = Cout; p.sH = (long long)Wo * Cout; p.sN = (long long)Ho * Wo * Cout;
  p.n_valid = Cout; p.m_valid = 0;
  p.out_fp32 = 0; p.relu = relu; p.atomic = 0; p.bias = nullptr;
  CUtensorMap ma[kMaxMaps], mb[kMaxMaps];
  const __nv_bfloat16* xb = reinterpret_cast<const __nv_bfloat16*>(x);
  int box[4] = {64, p.bw, p.bh, p.bn};
  int r;
  if (stride == 1) {
    long long dims[4] = {Cin, W, H, NB};
    long long str[3] = {Cin, (long long)W * Cin, (long long)H * W * Cin};
    r = make_map4(&ma[0], xb, dims, str, box);
    if (r) return r;
    ma[1] = ma[2] = ma[3] = ma[0];
    if (halo) {
      int hbox[4] = {64, p.bw + 2, p.bh + 2, 1};
      r = make_map4(&ma[1], xb, dims, str, hbox);
      if (r) return r;
    }
    const int pad = R / 2;
    for (int rr = 0; rr < R; ++rr)
      for (int ss = 0; ss < R; ++ss)
        p.taps[rr * R + ss] = Tap{0, (int8_t)(ss - pad), (int8_t)(rr - pad), 0};
  } else {
    // Input pixel (2h + dy, 2w + dx) lives in phase view (dy & 1, dx & 1) at (h + dy>>1, w + dx>>1).
    for (int py = 0; py < 2; ++py)
      for (int px = 0; px < 2; ++px) {
        long long dims[4] = {Cin, W / 2, H / 2, NB};
        long long str[3] = {2LL * Cin, 2LL * W * Cin, (long long)H * W * Cin};
        r = make_map4(&ma[py * 2 + px], xb + ((long long)py * W + px) * Cin, dims, str, box);
        if (r) return r;
      }
    const int pad = R / 2;
    for (int rr = 0; rr < R; ++rr)
      for (int ss = 0; ss < R; ++ss) {
        const int dy = rr - pad, dx = ss - pad;
        const int py = dy & 1, px = dx & 1;
        p.taps[rr * R + ss] =
            Tap{(int8_t)(py * 2 + px), (int8_t)((dx - px) / 2), (int8_t)((dy - py) / 2), 0};
      }
  }
  r = make_map2(&mb[0], w, (long long)p.ntaps * Cin, Cout, (long long)p.ntaps * Cin, kBK, p.BN);
  if (r) return r;
  mb[1] = mb[2] = mb[3] = mb[0];
  CUtensorMap mc;
  long long dimsC[4] = {Cout, Wo, Ho, NB};
  long long strC[3] = {p.sW, p.sH, p.sN};
  const bool tma_out = out_map_ok(y, strC, false) && quarter_box(p);
  if (tma_out) {
    int qb[4] = {0, p.qbox[0], p.qbox[1], p.qbox[2]};
    r = make_map_out(&mc, y, dimsC, strC, qb, false);
    if (r) return r;
  }
  if (bn_sums != nullptr) {
    if (!tma_out || relu) return -3;            // statistics are those of the raw conv output
    p.bn_sums = bn_sums; p.bn_C = Cout;
    r = clear_bn_sums(bn_sums, Cout, (cudaStream_t)stream);
    if (r) return r;
  }
  return launch(ma, mb, tma_out ? &mc : nullptr, p, (cudaStream_t)stream);
}

// NHWC conv2d weight gradient.  x: [NB,H,W,Cin], dy: [NB,Ho,Wo,Cout] (bf16), dw: [Cout,R,S,Cin] fp32
// (overwritten).  Same shape constraints as the forward; additionally Cout % 64 == 0.
extern "C" int b200dt_conv2d_wgrad_bf16(const void* x, const void* dy, float* dw, int NB, int H,
                                        int W, int Cin, int Cout, int R, int stride,
                                        void* stream) {
  if (!(R == 1 || R == 3) || !(stride == 1 || stride == 2) || (Cin % 64) || (Cout % 64)) return -3;
  if (stride == 2 && ((H | W) & 1)) return -3;
  const int Ho = H / stride, Wo = W / stride;
  GemmParams p{};
  p.kind = 1;
  const int n_total = R * R * Cin;
  p.BN = n_total >= 256 ? 256 : (n_total >= 128 ? 128 : 64);
  pick_box(Wo, Ho, 64, &p.bw, &p.bh, &p.bn);
  p.tiles_w = Wo / p.bw;
  p.tiles_h = Ho / p.bh;
  const int tiles_nb = (NB + p.bn - 1) / p.bn;
  p.kb_total = p.tiles_w * p.tiles_h * tiles_nb;
  p.ntaps = R * R;
  p.cin_blocks = Cin / 64;
  p.tiles_m = (Cout + kBM - 1) / kBM;
  p.tiles_n = (n_total + p.BN - 1) / p.BN;
  const int out_tiles = p.tiles_m * p.tiles_n;
  int splits = (2 * num_sms() + out_tiles - 1) / out_tiles;
  if (splits > p.kb_total) splits = p.kb_total;
  if (splits < 1) splits = 1;
  p.kb_per_split = (p.kb_total + splits - 1) / splits;
  p.splits = (p.kb_total + p.kb_per_split - 1) / p.kb_per_split;
  p.W = Wo; p.H = Ho; p.NB = NB;
  p.C = dw; p.sW = (long long)p.ntaps * Cin; p.n_valid = p.ntaps * Cin; p.m_valid = Cout;
  p.out_fp32 = 1; p.relu = 0; p.atomic = p.splits > 1; p.bias = nullptr;
  const bool det_split = deterministic() && p.splits > 1;
  float* part = nullptr;
  int ld_s = 0;
  if (det_split) {
    int e = det_split_begin(p, Cout, n_total, (cudaStream_t)stream, &part, &ld_s);
    if (e) return e;
  } else if (p.atomic) {
    cudaError_t e = cudaMemsetAsync(dw, 0, sizeof(float) * (size_t)Cout * p.ntaps * Cin,
                                    (cudaStream_t)stream);
    if (e != cudaSuccess) return (int)e;
  }
  CUtensorMap ma[kMaxMaps], mb[kMaxMaps];
  int box[4] = {64, p.bw, p.bh, p.bn};
  long long dimsA[4] = {Cout, Wo, Ho, NB};
  long long strA[3] = {Cout, (long long)Wo * Cout, (long long)Ho * Wo * Cout};
  int r = make_map4(&ma[0], dy, dimsA, strA, box);
  if (r) return r;
  ma[1] = ma[2] = ma[3] = ma[0];
  const __nv_bfloat16* xb = reinterpret_cast<const __nv_bfloat16*>(x);
  const int pad = R / 2;
  if (stride == 1) {
    long long dims[4] = {Cin, W, H, NB};
    long long str[3] = {Cin, (long long)W * Cin, (long long)H * W * Cin};
    r = make_map4(&mb[0], xb, dims, str, box);
    if (r) return r;
    mb[1] = mb[2] = mb[3] = mb[0];
    for (int rr = 0; rr < R; ++rr)
      for (int ss = 0; ss < R; ++ss)
        p.taps[rr * R + ss] = Tap{0, (int8_t)(ss - pad), (int8_t)(rr - pad), 0};
  } else {
    for (int py = 0; py < 2; ++py)
      for (int px = 0; px < 2; ++px) {
        long long dims[4] = {Cin, W / 2, H / 2, NB};
        long long str[3] = {2LL * Cin, 2LL * W * Cin, (long long)H * W * Cin};
        r = make_map4(&mb[py * 2 + px], xb + ((long long)py * W + px) * Cin, dims, str, box);
        if (r) return r;
      }
    for (int rr = 0; rr < R; ++rr)
      for (int ss = 0; ss < R; ++ss) {
        const int dy_ = rr - pad, dx_ = ss - pad;
        const int py = dy_ & 1, px = dx_ & 1;
        p.taps[rr * R + ss] =
            Tap{(int8_t)(py * 2 + px), (int8_t)((dx_ - px) / 2), (int8_t)((dy_ - py) / 2), 0};
      }
  }
  CUtensorMap mc;
  long long dimsC[4] = {(long long)p.ntaps * Cin, Cout, det_split ? p.splits : 1, 1};
  long long strC[3] = {p.sW, p.sW * Cout, p.sW * Cout * (det_split ? p.splits : 1)};
  int boxC[4] = {32, 32, 1, 1};
  const bool tma_out = out_map_ok(p.C, strC, true);
  if (tma_out) {
    r = make_map_out(&mc, p.C, dimsC, strC, boxC, true);
    if (r) return r;
  }
  r = launch(ma, mb, tma_out ? &mc : nullptr, p, (cudaStream_t)stream);
  if (det_split) {
    const int r2 = det_split_end(p, part, ld_s, dw, Cout, n_total, n_total, 0, (cudaStream_t)stream);
    if (r == 0) r = r2;
  }
  return r;
}

// Zero one phase view of a NHWC tensor (used when no filter tap reaches that phase).
__global__ void zero_phase_kernel(__nv_bfloat16* base, int NB, int Ho, int Wo, int C, long long sN,
                                  long long sH, long long sW) {
  const long long total = (long long)NB * Ho * Wo * (C / 8);
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int c8 = (int)(i % (C / 8));
    long long pix = i / (C / 8);
    const int w = (int)(pix % Wo); pix /= Wo;
    const int h = (int)(pix % Ho);
    const int n = (int)(pix / Ho);
    *reinterpret_cast<uint4*>(base + n * sN + h * sH + w * sW + c8 * 8) = make_uint4(0, 0, 0, 0);
  }
}

// NHWC conv2d data gradient.  dy: [NB,Ho,Wo,Cout] bf16, w: the forward filter [Cout,R,R,Cin] bf16
// (read in place as an MN-major B operand: no transposed copy), dx: [NB,H,W,Cin] bf16, fully
// overwritten.  For every output phase (py, px) in [0,stride)^2 the taps (rr, ss) with
// (py + pad - rr) % stride == 0 and (px + pad - ss) % stride == 0 contribute
//   dx(s*i + py, s*j + px) += dy(i + (py + pad - rr)/s, j + (px + pad - ss)/s) * w[., rr, ss, .]
// which is again a tiled-pixel implicit GEMM whose output rows land in a phase view of dx.
// accumulate != 0: dx already holds a gradient of the same tensor (the residual branch) and the
// epilogue adds to it with TMA reduce-add instead of storing, so no separate add kernel runs.
static int conv2d_dgrad_impl(const void* dy, const void* w, void* dx, int NB, int H, int W, int Cin,
                             int Cout, int R, int stride, int accumulate, void* stream) {
  if (!(R == 1 || R == 3) || !(stride == 1 || stride == 2) || (Cout % 8) || (Cin % 8)) return -3;
  if (stride == 2 && ((H | W) & 1)) return -3;
  const int Ho = H / stride, Wo = W / stride;
  const int pad = R / 2;
  for (int py = 0; py < stride; ++py)
    for (int px = 0; px < stride; ++px) {
      GemmParams p{};
      p.kind = 0;
      p.b_mn = 1;
      p.BN = pick_bn(Cin);
      pick_box(Wo, Ho, 128, &p.bw, &p.bh, &p.bn);
      p.tiles_w = Wo / p.bw; p.tiles_h = Ho / p.bh;
      p.tiles_m = p.tiles_w * p.tiles_h * ((NB + p.bn - 1) / p.bn);
      const bool halo = halo_shape(R, stride, Cin, Cout);
      if (halo) halo_tiling(p, Wo, Ho, NB);
      p.tiles_n = (Cin + p.BN - 1) / p.BN;
      p.splits = 1;
      p.cin_blocks = (Cout + 63) / 64;
      int nt = 0;
      for (int rr = 0; rr < R; ++rr)
        for (int ss = 0; ss < R; ++ss) {
          const int ey = py + pad - rr, ex = px + pad - ss;
          if ((ey % stride) || (ex % stride)) continue;
          p.taps[nt++] = Tap{0, (int8_t)(ex / stride), (int8_t)(ey / stride), (int8_t)(rr * R + ss)};
        }
      p.ntaps = nt;
      __nv_bfloat16* dxp = reinterpret_cast<__nv_bfloat16*>(dx) + ((long long)py * W + px) * Cin;
      p.W = Wo; p.H = Ho; p.NB = NB;
      p.C = dxp;
      p.sW = (long long)stride * Cin; p.sH = (long long)stride * W * Cin;
      p.sN = (long long)H * W * Cin;
      p.n_valid = Cin; p.out_fp32 = 0; p.relu = 0; p.atomic = accumulate ? 1 : 0; p.bias = nullptr;
      if (nt == 0) {
        if (accumulate) continue;          // nothing to add to this phase
        const long long total = (long long)NB * Ho * Wo * (Cin / 8);
        int blocks = (int)((total + 255) / 256);
        if (blocks > 148 * 8) blocks = 148 * 8;
        zero_phase_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(dxp, NB, Ho, Wo, Cin, p.sN, p.sH,
                                                                    p.sW);
        continue;
      }
      p.kb_total = nt * p.cin_blocks; p.kb_per_split = p.kb_total;
      CUtensorMap ma[kMaxMaps], mb[kMaxMaps];
      int box[4] = {64, p.bw, p.bh, p.bn};
      long long dims[4] = {Cout, Wo, Ho, NB};
      long long str[3] = {Cout, (long long)Wo * Cout, (long long)Ho * Wo * Cout};
      int r = make_map4(&ma[0], dy, dims, str, box);
      if (r) return r;
      ma[1] = ma[2] = ma[3] = ma[0];
      if (halo) {
        int hbox[4] = {64, p.bw + 2, p.bh + 2, 1};
        r = make_map4(&ma[1], dy, dims, str, hbox);
        if (r) return r;
      }
      // B = w viewed as [Cout rows][R*R*Cin columns]; a chunk is 64 columns x 64 rows
      r = make_map2(&mb[0], w, (long long)R * R * Cin, Cout, (long long)R * R * Cin, 64, 64);
      if (r) return r;
      mb[1] = mb[2] = mb[3] = mb[0];
      CUtensorMap mc;
      long long dimsC[4] = {Cin, Wo, Ho, NB};
      long long strC[3] = {p.sW, p.sH, p.sN};
      const bool tma_out = out_map_ok(dxp, strC, false) && quarter_box(p);
      if (accumulate && !tma_out) return -4;    // caller falls back to dgrad + add
      if (tma_out) {
        int qb[4] = {0, p.qbox[0], p.qbox[1], p.qbox[2]};
        r = make_map_out(&mc, dxp, dimsC, strC, qb, false);
        if (r) return r;
      }
      r = launch(ma, mb, tma_out ? &mc : nullptr, p, (cudaStream_t)stream);
      if (r) return r;
    }
  return (int)cudaGetLastError();
}

extern "C" int b200dt_conv2d_dgrad_bf16(const void* dy, const void* w, void* dx, int NB, int H,
                                        int W, int Cin, int Cout, int R, int stride,
                                        void* stream) {
  return conv2d_dgrad_impl(dy, w, dx, NB, H, W, Cin, Cout, R, stride, 0, stream);
}

extern "C" int b200dt_conv2d_dgrad_acc_bf16(const void* dy, const void* w, void* dx, int NB, int H,
                                            int W, int Cin, int Cout, int R, int stride,
                                            void* stream) {
  return conv2d_dgrad_impl(dy, w, dx, NB, H, W, Cin, Cout, R, stride, 1, stream);
}

// C[M,N] (bf16 or fp32) = A[M,K] * B[K,N]; A and B row-major bf16 (B is consumed as an MN-major
// operand).  Linear dgrad: dX[M,in] = dY[M,out] * W[out,in] without transposing W.
extern "C" int b200dt_gemm_nn_bf16(const void* A, const void* B, void* C, int M, int N, int K,
                                   int lda, int ldb, int ldc, int out_fp32, void* stream) {
  if (M <= 0 || N <= 0 || K <= 0) return 0;
  if ((lda % 8) || (ldb % 8)) return -3;
  GemmParams p{};
  p.kind = 0;
  p.b_mn = 1;
  p.BN = pick_bn(N);
  p.tiles_m = (M + kBM - 1) / kBM;
  p.tiles_n = (N + p.BN - 1) / p.BN;
  p.splits = 1;
  p.kb_total = (K + kBK - 1) / kBK;
  p.kb_per_split = p.kb_total;
  p.bw = 128; p.bh = 1; p.bn = 1;
  p.tiles_w = p.tiles_m; p.tiles_h = 1;
  p.W = M; p.H = 1; p.NB = 1;
  p.cin_blocks = p.kb_total;
  p.ntaps = 1;
  p.taps[0] = Tap{0, 0, 0, 0};
  p.C = C; p.sN = 0; p.sH = 0; p.sW = ldc;
  p.n_valid = N; p.m_valid = M;
  p.out_fp32 = out_fp32; p.relu = 0; p.atomic = 0; p.bias = nullptr;
  CUtensorMap ma[kMaxMaps], mb[kMaxMaps];
  long long dims[4] = {K, M, 1, 1};
  long long str[3] = {lda, (long long)lda * M, (long long)lda * M};
  int box[4] = {kBK, 128, 1, 1};
  int r = make_map4(&ma[0], A, dims, str, box);
  if (r) return r;
  ma[1] = ma[2] = ma[3] = ma[0];
  r = make_map2(&mb[0], B, N, K, ldb, 64, 64);
  if (r) return r;
  mb[1] = mb[2] = mb[3] = mb[0];
  CUtensorMap mc;
  long long dimsC[4] = {N, M, 1, 1};
  long long strC[3] = {ldc, (long long)ldc * M, (long long)ldc * M};
  const bool tma_out = out_map_ok(C, strC, out_fp32 != 0) && quarter_box(p);
  if (tma_out) {
    int qb[4] = {0, p.qbox[0], p.qbox[1], p.qbox[2]};
    r = make_map_out(&mc, C, dimsC, strC, qb, out_fp32 != 0);
    if (r) return r;
  }
  return launch(ma, mb, tma_out ? &mc : nullptr, p, (cudaStream_t)stream);
}
