// Probe: does tcgen05.mma read a 128-byte-swizzled K-major operand correctly when the descriptor's start
// address is an arbitrary 128-byte row of a larger staged patch and the 8-row groups are 1280 bytes apart
// (SBO = (8 + 2) * 128)?  That is what halo reuse for 3x3 convolutions needs: ONE staged (bh+2) x (bw+2) pixel
// patch, nine shifted operand views.  The patch is written with the swizzle TMA applies (chunk ^= row & 7 on the
// absolute shared address).  Build: nvcc -gencode arch=compute_100a,code=sm_100a -I csrc -I include ... ; prints
// one line per case: OK / MISMATCH (+ the first wrong element).
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "../../csrc/ptx.cuh"
#include "../../csrc/common.cuh"

using namespace b200;

constexpr int kPatchRows = 256;

struct Case { int r0, sbo, use_base_offset; };

__global__ void __launch_bounds__(128, 1)
probe_kernel(const __nv_bfloat16* __restrict__ A, const __nv_bfloat16* __restrict__ B, float* __restrict__ D,
             int r0, int sbo, int use_base_offset) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* sA = smem;                                   // kPatchRows x 128 B
  uint8_t* sB = smem + kPatchRows * 128;                // 64 x 128 B
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  if ((smem_u32(smem) & 1023u) != 0) __trap();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_mbar_init(); }
  if (warp == 0) tmem_alloc<64>(&tmem_base_s);
  for (int i = threadIdx.x; i < kPatchRows * 8; i += 128) {
    const int r = i >> 3, c = i & 7;
    *reinterpret_cast<uint4*>(sA + r * 128 + ((c ^ (r & 7)) << 4)) = *reinterpret_cast<const uint4*>(A + r * 64 + c * 8);
  }
  for (int i = threadIdx.x; i < 64 * 8; i += 128) {
    const int r = i >> 3, c = i & 7;
    *reinterpret_cast<uint4*>(sB + r * 128 + ((c ^ (r & 7)) << 4)) = *reinterpret_cast<const uint4*>(B + r * 64 + c * 8);
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_base_s;
  if (threadIdx.x == 0) {
    const uint32_t idesc = umma_idesc_bf16(128, 64, false, false);
    const uint32_t a0 = smem_u32(sA) + r0 * 128, b0 = smem_u32(sB);
    for (int ks = 0; ks < 4; ++ks) {
      uint64_t da = umma_desc_sw128(a0 + ks * 32, 16, sbo);
      if (use_base_offset) da |= static_cast<uint64_t>((a0 >> 7) & 7) << 49;
      umma_f16(tmem_base, da, umma_desc_sw128(b0 + ks * 32, 16, 1024), idesc, ks ? 1u : 0u);
    }
    umma_commit(&bar);
  }
  mbar_wait(&bar, 0);
  tc_fence_after();
  uint32_t v0[32], v1[32];
  const uint32_t taddr = tmem_base + (static_cast<uint32_t>(warp * 32) << 16);
  tmem_ld_32x32(taddr, v0);
  tmem_ld_32x32(taddr + 32, v1);
  tmem_ld_wait();
  float* d = D + (warp * 32 + lane) * 64;
  for (int j = 0; j < 32; ++j) { d[j] = __uint_as_float(v0[j]); d[32 + j] = __uint_as_float(v1[j]); }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc<64>(tmem_base); }
}

int main() {
  std::vector<__nv_bfloat16> hA(kPatchRows * 64), hB(64 * 64);
  std::vector<float> fA(kPatchRows * 64), fB(64 * 64);
  for (int r = 0; r < kPatchRows; ++r)
    for (int k = 0; k < 64; ++k) {
      const float v = (float)((r * 7 + k * 3 + (r * k) % 5) % 9 - 4);
      fA[r * 64 + k] = v; hA[r * 64 + k] = __float2bfloat16(v);
    }
  for (int n = 0; n < 64; ++n)
    for (int k = 0; k < 64; ++k) {
      const float v = (float)((n * 5 + k * 2 + (n + k) % 3) % 7 - 3);
      fB[n * 64 + k] = v; hB[n * 64 + k] = __float2bfloat16(v);
    }
  __nv_bfloat16 *dA, *dB; float* dD;
  cudaMalloc(&dA, hA.size() * 2); cudaMalloc(&dB, hB.size() * 2); cudaMalloc(&dD, 128 * 64 * 4);
  cudaMemcpy(dA, hA.data(), hA.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, hB.data(), hB.size() * 2, cudaMemcpyHostToDevice);
  const int smem = kPatchRows * 128 + 64 * 128;
  cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  const Case cases[] = {{0, 1024, 0}, {8, 1024, 0}, {3, 1024, 0}, {3, 1024, 1}, {0, 1280, 0}, {11, 1280, 0},
                        {11, 1280, 1}, {1, 1280, 0}, {21, 1280, 0}, {0, 2304, 0}, {19, 2304, 0}, {5, 640, 0}};
  for (const Case& c : cases) {
    cudaMemset(dD, 0, 128 * 64 * 4);
    probe_kernel<<<1, 128, smem>>>(dA, dB, dD, c.r0, c.sbo, c.use_base_offset);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("r0=%d sbo=%d bo=%d: CUDA error %s\n", c.r0, c.sbo, c.use_base_offset, cudaGetErrorString(e)); return 1; }
    std::vector<float> hD(128 * 64);
    cudaMemcpy(hD.data(), dD, hD.size() * 4, cudaMemcpyDeviceToHost);
    int bad = 0, fm = -1, fn = -1; float got = 0, want = 0;
    for (int m = 0; m < 128; ++m) {
      const int row = c.r0 + (m / 8) * (c.sbo / 128) + m % 8;      // "absolute address" hypothesis
      for (int n = 0; n < 64; ++n) {
        float ref = 0;
        for (int k = 0; k < 64; ++k) ref += fA[row * 64 + k] * fB[n * 64 + k];
        if (hD[m * 64 + n] != ref) { if (!bad) { fm = m; fn = n; got = hD[m * 64 + n]; want = ref; } ++bad; }
      }
    }
    printf("r0=%2d sbo=%4d base_offset=%d : %s", c.r0, c.sbo, c.use_base_offset, bad ? "MISMATCH" : "OK");
    if (bad) printf("  (%d wrong, first at m=%d n=%d got %.0f want %.0f)", bad, fm, fn, got, want);
    printf("\n");
  }
  return 0;
}
