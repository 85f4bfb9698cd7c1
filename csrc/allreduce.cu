// Fused DDP bucket collectives over NVLink 5 / NVSwitch peer memory.  sm_100a only.
//
// One kernel per bucket does what the reference's stack does in four steps
// (/root/reference/ddp_gpus.py:32,38 -> torch Reducer: copy grads into a flat bucket scaled by
// 1/world, ncclAllReduce, wait, copy the bucket back into .grad):
//
//   P0  gather  : read the scattered gradient tensors (pointer table), scale by 1/world, cast to
//                 the wire dtype (fp32 or bf16) and write them into this rank's SYMMETRIC staging
//                 buffer (peer-mapped on every GPU of the node);
//   B1  barrier : per-CTA flags in the symmetric signal pads (release/acquire at system scope);
//   P1  reduce  : one-shot  - every rank loads all peers' staging and sums (latency-optimal);
//                 two-shot  - rank r sums sub-slice r of every chunk over all peers and stores the
//                             result into every peer's staging (bandwidth-optimal, P2P only);
//                 NVLS      - multimem.ld_reduce on sub-slice r (the switch adds), multimem.st of
//                             the result to all ranks (the switch fans out);
//   B2  barrier ; P2 scatter: staging -> the .grad tensors (fp32), i.e. no separate copy-back.
//
// CTA b of every rank owns the same chunks in every phase, so only CTA-local flags are needed and
// the kernel never requires all CTAs to be co-resident.  Flags are consumed (1 -> 0) by the waiter,
// so a slot can be reused by the next barrier without epochs.  Waits are bounded in TIME
// (%globaltimer): a rank that waits longer than the limit writes {code, CTA, peer} into the status
// words (pinned host memory, readable by the host at any moment), raises the abort word in every peer's
// signal pad so that they stop waiting too, and TRAPS: the averaged gradients are never written from
// partial data, the process dies loudly on its next CUDA call and the launcher tears the job down --
// what ProcessGroupNCCL's watchdog does for the reference's all-reduce (ddp_gpus.py:38).
#include <atomic>
#include <stdlib.h>
#include "common.cuh"
#include "ptx.cuh"
#include "../include/b200dt.h"

namespace b200 {

enum CollMode { kOneShot = 0, kTwoShot = 1, kNvls = 2, kBcast = 3 };

struct CollTensor {
  void* ptr;
  long long numel;
};
struct CollChunk {
  int tensor;        // index into the tensor table
  int start;         // first element inside the tensor
  int len;           // elements in this chunk
  int pad_;
  long long stage_off;   // element offset inside the staging region (multiple of 8)
};

struct CollArgs {
  void* bufs[B200DT_MAX_WORLD];
  uint32_t* sigs[B200DT_MAX_WORLD];
  void* mc;
  const CollTensor* tensors;
  const CollChunk* chunks;
  int n_chunks;
  int rank, world, root;
  long long stage_base;     // byte offset of the staging region inside each symmetric buffer
  int tensor_bf16;          // dtype of the user tensors: 0 fp32, 1 bf16
  int wire_bf16;            // dtype on the wire / in staging
  float scale;
  int sig_base;             // first signal word this launch may use
  int* status;              // {code, cta << 8 | peer, elapsed ms}; host-mapped when the arena provides it
  unsigned long long timeout_ns;
  int in_place;             // tensors live (or are copied unscaled) in the staging region; the result stays there
};

constexpr int kCollThreads = 512;

__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// A wait exceeded the limit (code 1: a peer never consumed my previous flag; 2: a peer's flag never
// arrived) or a peer raised the abort word (code 3).  Never returns.
__device__ __noinline__ void barrier_abort(const CollArgs& a, int code, int peer, unsigned long long waited_ns) {
  if (a.status != nullptr) {
    // plain stores (the words may live in host memory; whichever waiter gets here last wins)
    volatile int* st = a.status;
    st[1] = (int)((blockIdx.x << 8) | (unsigned)peer);
    st[2] = (int)(waited_ns / 1000000ull);
    __threadfence_system();
    st[0] = code;
  }
  for (int r = 0; r < a.world; ++r) st_relaxed_sys(a.sigs[r] + B200DT_SIGNAL_WORDS - 1, 1u);
  __threadfence_system();
  __trap();
}

__device__ __forceinline__ void block_barrier(const CollArgs& a) {
  __syncthreads();
  const int t = threadIdx.x;
  if (t < a.world) {
    const int slot = a.sig_base + blockIdx.x * a.world;
    uint32_t* remote = a.sigs[t] + slot + a.rank;
    const uint32_t* abort_word = a.sigs[a.rank] + B200DT_SIGNAL_WORDS - 1;
    unsigned long long t0 = 0;
    uint32_t spins = 0;
    while (cas_release_sys(remote, 0u, 1u) != 0u) {
      if ((++spins & 255u) == 0) {
        const unsigned long long now = global_ns();
        if (t0 == 0) t0 = now;
        if (ld_relaxed_sys(abort_word) != 0u) barrier_abort(a, 3, t, now - t0);
        if (now - t0 > a.timeout_ns) barrier_abort(a, 1, t, now - t0);
      }
    }
    uint32_t* local = a.sigs[a.rank] + slot + t;
    spins = 0; t0 = 0;
    while (cas_acquire_sys(local, 1u, 0u) != 1u) {
      if ((++spins & 255u) == 0) {
        const unsigned long long now = global_ns();
        if (t0 == 0) t0 = now;
        if (ld_relaxed_sys(abort_word) != 0u) barrier_abort(a, 3, t, now - t0);
        if (now - t0 > a.timeout_ns) barrier_abort(a, 2, t, now - t0);
      }
    }
  }
  __syncthreads();
}

// ---- 8-element group accessors -------------------------------------------------------------
__device__ __forceinline__ void load8_tensor(const void* base, long long idx, int bf16, int n,
                                             bool aligned, float (&f)[8]) {
  if (bf16) {
    const __nv_bfloat16* p = reinterpret_cast<const __nv_bfloat16*>(base) + idx;
    if (n == 8 && aligned) {
      unpack8(ld_v4(p), f);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] = j < n ? __bfloat162float(p[j]) : 0.f;
    }
  } else {
    const float* p = reinterpret_cast<const float*>(base) + idx;
    if (n == 8 && aligned) {
      const uint4 a = ld_v4(p), b = ld_v4(p + 4);
      f[0] = __uint_as_float(a.x); f[1] = __uint_as_float(a.y);
      f[2] = __uint_as_float(a.z); f[3] = __uint_as_float(a.w);
      f[4] = __uint_as_float(b.x); f[5] = __uint_as_float(b.y);
      f[6] = __uint_as_float(b.z); f[7] = __uint_as_float(b.w);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) f[j] = j < n ? p[j] : 0.f;
    }
  }
}
__device__ __forceinline__ void store8_tensor(void* base, long long idx, int bf16, int n,
                                              bool aligned, const float (&f)[8]) {
  if (bf16) {
    __nv_bfloat16* p = reinterpret_cast<__nv_bfloat16*>(base) + idx;
    if (n == 8 && aligned) {
      st_v4(p, pack8(f));
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (j < n) p[j] = __float2bfloat16(f[j]);
    }
  } else {
    float* p = reinterpret_cast<float*>(base) + idx;
    if (n == 8 && aligned) {
      st_v4(p, make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]),
                          __float_as_uint(f[3])));
      st_v4(p + 4, make_uint4(__float_as_uint(f[4]), __float_as_uint(f[5]), __float_as_uint(f[6]),
                              __float_as_uint(f[7])));
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (j < n) p[j] = f[j];
    }
  }
}
// Staging is always 16-byte aligned and padded to 8 elements.
__device__ __forceinline__ void load8_stage(const void* stage, long long e, int bf16, float (&f)[8]) {
  if (bf16) {
    unpack8(ld_cg_v4(reinterpret_cast<const __nv_bfloat16*>(stage) + e), f);
  } else {
    const float* p = reinterpret_cast<const float*>(stage) + e;
    const uint4 a = ld_cg_v4(p), b = ld_cg_v4(p + 4);
    f[0] = __uint_as_float(a.x); f[1] = __uint_as_float(a.y);
    f[2] = __uint_as_float(a.z); f[3] = __uint_as_float(a.w);
    f[4] = __uint_as_float(b.x); f[5] = __uint_as_float(b.y);
    f[6] = __uint_as_float(b.z); f[7] = __uint_as_float(b.w);
  }
}
__device__ __forceinline__ void store8_stage(void* stage, long long e, int bf16, const float (&f)[8]) {
  if (bf16) {
    st_v4(reinterpret_cast<__nv_bfloat16*>(stage) + e, pack8(f));
  } else {
    float* p = reinterpret_cast<float*>(stage) + e;
    st_v4(p, make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]),
                        __float_as_uint(f[3])));
    st_v4(p + 4, make_uint4(__float_as_uint(f[4]), __float_as_uint(f[5]), __float_as_uint(f[6]),
                            __float_as_uint(f[7])));
  }
}

// Sum one 8-element group over all ranks' staging buffers.  Every peer load is issued before the
// first add (up to 16 independent 16-byte NVLink loads in flight per thread); the summation order
// is rank 0..world-1 on every rank, so all ranks produce bit-identical results.
__device__ __forceinline__ void reduce_group(const CollArgs& a, long long e, float (&acc)[8]) {
  uint4 lo[B200DT_MAX_WORLD], hi[B200DT_MAX_WORLD];
  if (a.wire_bf16) {
#pragma unroll
    for (int r = 0; r < B200DT_MAX_WORLD; ++r)
      if (r < a.world)
        lo[r] = ld_cg_v4(reinterpret_cast<const __nv_bfloat16*>(
                              reinterpret_cast<const char*>(a.bufs[r]) + a.stage_base) + e);
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
#pragma unroll
    for (int r = 0; r < B200DT_MAX_WORLD; ++r)
      if (r < a.world) {
        float f[8];
        unpack8(lo[r], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
      }
  } else {
#pragma unroll
    for (int r = 0; r < B200DT_MAX_WORLD; ++r)
      if (r < a.world) {
        const float* p = reinterpret_cast<const float*>(
                             reinterpret_cast<const char*>(a.bufs[r]) + a.stage_base) + e;
        lo[r] = ld_cg_v4(p);
        hi[r] = ld_cg_v4(p + 4);
      }
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
#pragma unroll
    for (int r = 0; r < B200DT_MAX_WORLD; ++r)
      if (r < a.world) {
        acc[0] += __uint_as_float(lo[r].x); acc[1] += __uint_as_float(lo[r].y);
        acc[2] += __uint_as_float(lo[r].z); acc[3] += __uint_as_float(lo[r].w);
        acc[4] += __uint_as_float(hi[r].x); acc[5] += __uint_as_float(hi[r].y);
        acc[6] += __uint_as_float(hi[r].z); acc[7] += __uint_as_float(hi[r].w);
      }
  }
}

// Two-shot reduce + push for one chunk with a compile-time world size.  Works on 16-byte vectors so
// that every warp-level load / store touches one contiguous 512-byte run (no sector is requested
// twice over NVLink), and keeps 16 independent peer loads in flight per thread regardless of W.
template <int W>
__device__ __forceinline__ void twoshot_chunk(const CollArgs& a, const CollChunk& ch) {
  constexpr int VU = 16 / W;
  const int esz = a.wire_bf16 ? 2 : 4;
  const long long nvec = (long long)((ch.len + 7) / 8) * 8 * esz / 16;
  const long long per = (nvec + W - 1) / W;
  const long long v_beg = a.rank * per;
  const long long v_end = min(nvec, v_beg + per);
  const long long byte0 = a.stage_base + ch.stage_off * esz;
  const float post = a.in_place ? a.scale : 1.f;      // in-place mode scales the SUM (inputs are raw)
  for (long long v0 = v_beg + threadIdx.x; v0 < v_end; v0 += (long long)VU * blockDim.x) {
    uint4 in[VU][W];
#pragma unroll
    for (int u = 0; u < VU; ++u) {
      const long long v = v0 + (long long)u * blockDim.x;
      if (v < v_end) {
#pragma unroll
        for (int r = 0; r < W; ++r)
          in[u][r] = ld_cg_v4(reinterpret_cast<const char*>(a.bufs[r]) + byte0 + v * 16);
      }
    }
#pragma unroll
    for (int u = 0; u < VU; ++u) {
      const long long v = v0 + (long long)u * blockDim.x;
      if (v >= v_end) continue;
      uint4 out;
      if (a.wire_bf16) {
        float acc[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = 0.f;
#pragma unroll
        for (int r = 0; r < W; ++r) {          // rank order: identical result on every rank
          float f[8];
          unpack8(in[u][r], f);
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[j] += f[j];
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] *= post;
        out = pack8(acc);
      } else {
        float x = 0.f, y = 0.f, z = 0.f, w = 0.f;
#pragma unroll
        for (int r = 0; r < W; ++r) {
          x += __uint_as_float(in[u][r].x); y += __uint_as_float(in[u][r].y);
          z += __uint_as_float(in[u][r].z); w += __uint_as_float(in[u][r].w);
        }
        x *= post; y *= post; z *= post; w *= post;
        out = make_uint4(__float_as_uint(x), __float_as_uint(y), __float_as_uint(z), __float_as_uint(w));
      }
#pragma unroll
      for (int r = 0; r < W; ++r)
        st_v4(reinterpret_cast<char*>(a.bufs[r]) + byte0 + v * 16, out);
    }
  }
}

template <int kMode, int W>     // W: compile-time world size for the two-shot fast path, 0 = generic
__global__ void __launch_bounds__(kCollThreads)
coll_kernel(const __grid_constant__ CollArgs a) {
  const int esz_t = a.tensor_bf16 ? 2 : 4;
  char* my_stage = reinterpret_cast<char*>(a.bufs[a.rank]) + a.stage_base;

  // ---------------------------------------------------------------- P0: gather + scale + cast
  if (kMode != kBcast || a.rank == a.root) {
    for (int c = blockIdx.x; c < a.n_chunks; c += gridDim.x) {
      const CollChunk ch = a.chunks[c];
      const CollTensor tt = a.tensors[ch.tensor];
      const bool aligned = ((reinterpret_cast<uintptr_t>(tt.ptr) + (size_t)ch.start * esz_t) & 15) == 0;
      const int groups = (ch.len + 7) / 8;
      // in-place mode: a tensor that already lives at its slot of the staging region is not
      // touched at all; a stray one (small parameters) is copied unscaled, the sum is scaled in P1
      if (a.in_place && reinterpret_cast<const char*>(tt.ptr) + (size_t)ch.start * esz_t ==
                            my_stage + (size_t)ch.stage_off * esz_t)
        continue;
      const float pre = a.in_place ? 1.f : a.scale;
      if (kMode != kBcast) {
        // four groups per thread: all loads are issued before the first dependent store
        for (int g0 = threadIdx.x; g0 < groups; g0 += 4 * blockDim.x) {
          float f[4][8];
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int g = g0 + u * blockDim.x;
            if (g < groups)
              load8_tensor(tt.ptr, (long long)ch.start + g * 8, a.tensor_bf16, min(8, ch.len - g * 8),
                           aligned, f[u]);
          }
#pragma unroll
          for (int u = 0; u < 4; ++u) {
            const int g = g0 + u * blockDim.x;
            if (g < groups) {
              if (pre != 1.f) {
#pragma unroll
                for (int j = 0; j < 8; ++j) f[u][j] *= pre;
              }
              store8_stage(my_stage, ch.stage_off + g * 8, a.wire_bf16, f[u]);
            }
          }
        }
        continue;
      }
      for (int g = threadIdx.x; g < groups; g += blockDim.x) {
        float f[8];
        const int n = min(8, ch.len - g * 8);
        load8_tensor(tt.ptr, (long long)ch.start + g * 8, a.tensor_bf16, n, aligned, f);
        if (a.scale != 1.f) {       // scale == 1 keeps the words bit-exact (raw broadcast of any dtype)
#pragma unroll
          for (int j = 0; j < 8; ++j) f[j] *= a.scale;
        }
        if (kMode == kBcast) {
          // root pushes straight into every rank's staging (its own included)
          if (a.mc != nullptr) {
            // one multicast store fans out in the switch
            char* dst = reinterpret_cast<char*>(a.mc) + a.stage_base;
            if (a.wire_bf16) {
              multimem_st_v4(reinterpret_cast<__nv_bfloat16*>(dst) + ch.stage_off + g * 8, pack8(f));
            } else {
              float* p = reinterpret_cast<float*>(dst) + ch.stage_off + g * 8;
              multimem_st_v4(p, make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]),
                                           __float_as_uint(f[2]), __float_as_uint(f[3])));
              multimem_st_v4(p + 4, make_uint4(__float_as_uint(f[4]), __float_as_uint(f[5]),
                                               __float_as_uint(f[6]), __float_as_uint(f[7])));
            }
          } else {
            for (int r = 0; r < a.world; ++r)
              store8_stage(reinterpret_cast<char*>(a.bufs[r]) + a.stage_base, ch.stage_off + g * 8,
                           a.wire_bf16, f);
          }
        } else {
          store8_stage(my_stage, ch.stage_off + g * 8, a.wire_bf16, f);
        }
      }
    }
  }
  if (a.world > 1) block_barrier(a);

  // ---------------------------------------------------------------- P1: reduce
  if (kMode == kOneShot) {
    // Reduce every chunk from all peers and write the result straight into the tensors.
    for (int c = blockIdx.x; c < a.n_chunks; c += gridDim.x) {
      const CollChunk ch = a.chunks[c];
      const CollTensor tt = a.tensors[ch.tensor];
      const bool aligned = ((reinterpret_cast<uintptr_t>(tt.ptr) + (size_t)ch.start * esz_t) & 15) == 0;
      const int groups = (ch.len + 7) / 8;
      for (int g = threadIdx.x; g < groups; g += blockDim.x) {
        float acc[8];
        reduce_group(a, ch.stage_off + g * 8, acc);
        const int n = min(8, ch.len - g * 8);
        store8_tensor(tt.ptr, (long long)ch.start + g * 8, a.tensor_bf16, n, aligned, acc);
      }
    }
    // Peers may still be reading this rank's staging: hold the next writer back.
    if (a.world > 1) block_barrier(a);
    return;
  }

  if (kMode == kTwoShot || kMode == kNvls) {
    for (int c = blockIdx.x; c < a.n_chunks; c += gridDim.x) {
      const CollChunk ch = a.chunks[c];
      if (kMode == kTwoShot && W > 0) {
        twoshot_chunk<(W > 0 ? W : 2)>(a, ch);
        continue;
      }
      const int groups = (ch.len + 7) / 8;
      const int per = (groups + a.world - 1) / a.world;      // groups owned by each rank
      const int g0 = a.rank * per;
      const int g1 = min(groups, g0 + per);
      if (kMode == kNvls) {
        // The switch adds (multimem.ld_reduce) and fans out (multimem.st).  A load-reduce is a round trip
        // through the NVSwitch, so several of them are kept in flight per thread before the first store.
        constexpr int U = 4;
        char* mcs = reinterpret_cast<char*>(a.mc) + a.stage_base;
        const float post = a.in_place ? a.scale : 1.f;
        for (int gb = g0 + threadIdx.x; gb < g1; gb += U * blockDim.x) {
          if (a.wire_bf16) {
            uint4 v[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
              const int g = gb + u * blockDim.x;
              if (g < g1) v[u] = multimem_ld_reduce_bf16x8(reinterpret_cast<__nv_bfloat16*>(mcs) + ch.stage_off + (long long)g * 8);
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
              const int g = gb + u * blockDim.x;
              if (g >= g1) continue;
              if (post != 1.f) {
                float f[8];
                unpack8(v[u], f);
#pragma unroll
                for (int j = 0; j < 8; ++j) f[j] *= post;
                v[u] = pack8(f);
              }
              multimem_st_v4(reinterpret_cast<__nv_bfloat16*>(mcs) + ch.stage_off + (long long)g * 8, v[u]);
            }
          } else {
            uint4 v0[U], v1[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
              const int g = gb + u * blockDim.x;
              if (g < g1) {
                float* p = reinterpret_cast<float*>(mcs) + ch.stage_off + (long long)g * 8;
                v0[u] = multimem_ld_reduce_f32x4(p);
                v1[u] = multimem_ld_reduce_f32x4(p + 4);
              }
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
              const int g = gb + u * blockDim.x;
              if (g >= g1) continue;
              float* p = reinterpret_cast<float*>(mcs) + ch.stage_off + (long long)g * 8;
              if (post != 1.f) {
                v0[u] = make_uint4(__float_as_uint(__uint_as_float(v0[u].x) * post), __float_as_uint(__uint_as_float(v0[u].y) * post),
                                   __float_as_uint(__uint_as_float(v0[u].z) * post), __float_as_uint(__uint_as_float(v0[u].w) * post));
                v1[u] = make_uint4(__float_as_uint(__uint_as_float(v1[u].x) * post), __float_as_uint(__uint_as_float(v1[u].y) * post),
                                   __float_as_uint(__uint_as_float(v1[u].z) * post), __float_as_uint(__uint_as_float(v1[u].w) * post));
              }
              multimem_st_v4(p, v0[u]);
              multimem_st_v4(p + 4, v1[u]);
            }
          }
        }
        continue;
      }
      for (int g = g0 + threadIdx.x; g < g1; g += blockDim.x) {
        const long long e = ch.stage_off + (long long)g * 8;
        float acc[8];
        reduce_group(a, e, acc);
        if (a.in_place) {
#pragma unroll
          for (int j = 0; j < 8; ++j) acc[j] *= a.scale;
        }
#pragma unroll
        for (int r = 0; r < B200DT_MAX_WORLD; ++r)
          if (r < a.world)
            store8_stage(reinterpret_cast<char*>(a.bufs[r]) + a.stage_base, e, a.wire_bf16, acc);
      }
    }
    if (a.world > 1) block_barrier(a);
  }
  if (a.in_place && kMode != kBcast) return;     // the staging region IS the result (gradient views)

  // ---------------------------------------------------------------- P2: staging -> tensors
  for (int c = blockIdx.x; c < a.n_chunks; c += gridDim.x) {
    const CollChunk ch = a.chunks[c];
    const CollTensor tt = a.tensors[ch.tensor];
    const bool aligned = ((reinterpret_cast<uintptr_t>(tt.ptr) + (size_t)ch.start * esz_t) & 15) == 0;
    const int groups = (ch.len + 7) / 8;
    for (int g0 = threadIdx.x; g0 < groups; g0 += 4 * blockDim.x) {
      float f[4][8];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int g = g0 + u * blockDim.x;
        if (g < groups) load8_stage(my_stage, ch.stage_off + g * 8, a.wire_bf16, f[u]);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int g = g0 + u * blockDim.x;
        if (g < groups)
          store8_tensor(tt.ptr, (long long)ch.start + g * 8, a.tensor_bf16, min(8, ch.len - g * 8), aligned, f[u]);
      }
    }
  }
  if (kMode == kBcast && a.world > 1) {
    // root must not overwrite its peers' staging with the next broadcast before they drained it
    block_barrier(a);
  }
}

// Single-GPU emulation of the reduce arithmetic (tests without peers): out = scale * sum_r bufs[r].
__global__ void emulate_kernel(const CollArgs a, void* out, long long n) {
  for (long long g = blockIdx.x * (long long)blockDim.x + threadIdx.x; g * 8 < n;
       g += (long long)gridDim.x * blockDim.x) {
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    const int cnt = (int)min(8LL, n - g * 8);
    for (int r = 0; r < a.world; ++r) {
      float f[8];
      load8_tensor(a.bufs[r], g * 8, a.wire_bf16, cnt, true, f);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += f[j] * a.scale;
    }
    store8_tensor(out, g * 8, a.wire_bf16, cnt, true, acc);
  }
}

__global__ void scale_cast_kernel(const float* __restrict__ src, void* __restrict__ dst,
                                  long long n, float scale, int dst_bf16) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    const float v = src[i] * scale;
    if (dst_bf16) reinterpret_cast<__nv_bfloat16*>(dst)[i] = __float2bfloat16(v);
    else reinterpret_cast<float*>(dst)[i] = v;
  }
}
__global__ void cast_to_f32_kernel(const void* __restrict__ src, float* __restrict__ dst,
                                   long long n, int src_bf16) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x)
    dst[i] = src_bf16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(src)[i])
                      : reinterpret_cast<const float*>(src)[i];
}

// Wait limit of the flag barriers: B200DT_COLL_TIMEOUT_S (seconds, fractional allowed), default 30 s.
static std::atomic<long long> g_timeout_ns{-1};
static unsigned long long coll_timeout_ns() {
  long long v = g_timeout_ns.load(std::memory_order_relaxed);
  if (v < 0) {
    const char* e = getenv("B200DT_COLL_TIMEOUT_S");
    double sec = e != nullptr ? atof(e) : 30.0;
    if (!(sec > 0.0)) sec = 30.0;
    v = (long long)(sec * 1e9);
    g_timeout_ns.store(v, std::memory_order_relaxed);
  }
  return (unsigned long long)v;
}

}  // namespace b200

using namespace b200;

extern "C" void b200dt_set_coll_timeout_ms(long long ms) {
  g_timeout_ns.store(ms > 0 ? ms * 1000000ll : -1, std::memory_order_relaxed);
}

extern "C" int b200dt_coll_chunk_bytes(void) { return (int)sizeof(CollChunk); }

// mode: 0 one-shot all-reduce, 1 two-shot, 2 NVLS, 3 broadcast from `root`.
// bufs / sigs: HOST arrays of `world` peer-mapped device pointers (symmetric buffer, signal pad).
// tensors: device array of {ptr, numel}; chunks: device array of CollChunk (see
// pytorch_distributed_training_tutorials_b200/ops/collectives.py for the builder).
extern "C" int b200dt_coll_launch(int mode, const uint64_t* bufs, const uint64_t* sigs, uint64_t mc,
                                  int rank, int world, int root, const void* tensors,
                                  const void* chunks, int n_chunks, long long stage_base_bytes,
                                  int tensor_bf16, int wire_bf16, float scale, int sig_base,
                                  int grid, int* status, int in_place, void* stream) {
  if (world < 1 || world > B200DT_MAX_WORLD) return -3;
  if (n_chunks <= 0) return 0;
  if (mode == kNvls && mc == 0) return -4;
  // in place: every rank overwrites the slice it owns, which only two-shot / NVLS make race-free,
  // and the tensors must already have the wire dtype
  if (in_place && (mode == kOneShot || mode == kBcast || (wire_bf16 != 0) != (tensor_bf16 != 0))) return -3;
  if (grid < 1) grid = 1;
  if (sig_base < 0 || sig_base + grid * world > B200DT_SIGNAL_WORDS - 1) return -5;   // last word = abort flag
  CollArgs a{};
  for (int r = 0; r < world; ++r) {
    a.bufs[r] = reinterpret_cast<void*>(bufs[r]);
    a.sigs[r] = reinterpret_cast<uint32_t*>(sigs[r]);
  }
  a.mc = reinterpret_cast<void*>(mc);
  a.tensors = reinterpret_cast<const CollTensor*>(tensors);
  a.chunks = reinterpret_cast<const CollChunk*>(chunks);
  a.n_chunks = n_chunks;
  a.rank = rank; a.world = world; a.root = root;
  a.stage_base = stage_base_bytes;
  a.tensor_bf16 = tensor_bf16; a.wire_bf16 = wire_bf16 || tensor_bf16;
  a.scale = scale; a.sig_base = sig_base; a.status = status;
  a.timeout_ns = coll_timeout_ns();
  a.in_place = in_place;
  cudaStream_t st = (cudaStream_t)stream;
  switch (mode) {
    case kOneShot: coll_kernel<kOneShot, 0><<<grid, kCollThreads, 0, st>>>(a); break;
    case kTwoShot:
      if (world == 8) coll_kernel<kTwoShot, 8><<<grid, kCollThreads, 0, st>>>(a);
      else if (world == 4) coll_kernel<kTwoShot, 4><<<grid, kCollThreads, 0, st>>>(a);
      else if (world == 2) coll_kernel<kTwoShot, 2><<<grid, kCollThreads, 0, st>>>(a);
      else coll_kernel<kTwoShot, 0><<<grid, kCollThreads, 0, st>>>(a);
      break;
    case kNvls: coll_kernel<kNvls, 0><<<grid, kCollThreads, 0, st>>>(a); break;
    case kBcast: coll_kernel<kBcast, 0><<<grid, kCollThreads, 0, st>>>(a); break;
    default: return -3;
  }
  return (int)cudaGetLastError();
}

extern "C" int b200dt_allreduce_local_emulate(const uint64_t* bufs, int world, void* out,
                                              long long n, int dtype, float scale, void* stream) {
  if (world < 1 || world > B200DT_MAX_WORLD) return -3;
  CollArgs a{};
  for (int r = 0; r < world; ++r) a.bufs[r] = reinterpret_cast<void*>(bufs[r]);
  a.world = world; a.wire_bf16 = dtype; a.scale = scale;
  long long blocks = (n / 8 + 255) / 256 + 1;
  if (blocks > 148 * 8) blocks = 148 * 8;
  emulate_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(a, out, n);
  return (int)cudaGetLastError();
}

extern "C" int b200dt_scale_cast(const float* src, void* dst, long long n, float scale,
                                 int dst_bf16, void* stream) {
  if (n <= 0) return 0;
  long long blocks = (n + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  scale_cast_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(src, dst, n, scale, dst_bf16);
  return (int)cudaGetLastError();
}

extern "C" int b200dt_cast_to_f32(const void* src, float* dst, long long n, int src_bf16,
                                  void* stream) {
  if (n <= 0) return 0;
  long long blocks = (n + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  cast_to_f32_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(src, dst, n, src_bf16);
  return (int)cudaGetLastError();
}
