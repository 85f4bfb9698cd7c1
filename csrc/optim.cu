// Multi-tensor SGD and Adam: one launch updates every parameter of the model, optionally
// refreshes the bf16 shadow copy that the tcgen05 kernels read, and optionally zeroes the
// gradient in the same pass (fused zero_grad).
// Replaces torch.optim.SGD (/root/reference/ddp_gpus.py:82, 03.model_parallel.ipynb:534,977),
// torch.optim.Adam (/root/reference/01.data_parallel.ipynb:287) and optimizer.zero_grad()
// (/root/reference/ddp_gpus.py:35).
#include "common.cuh"
#include "ptx.cuh"
#include "../include/b200dt.h"

namespace b200 {

struct TensorMeta {
  float* param;
  float* grad;
  __nv_bfloat16* shadow;
  long long numel;
  float* exp_avg;
  float* exp_avg_sq;
  long long r0, r1;
};
static_assert(sizeof(TensorMeta) == 64, "meta rows are 8 x int64");

template <bool kAdam>
__global__ void __launch_bounds__(256)
multi_opt_kernel(const TensorMeta* __restrict__ meta, const int2* __restrict__ chunks, float lr,
                 float beta1, float beta2, float eps, float wd, float bc1, float bc2_rsqrt,
                 float gscale, int zero_grad, const float* __restrict__ bc_dev) {
  if (kAdam && bc_dev != nullptr) {       // capturable Adam: the bias corrections follow a device-side step count
    bc1 = bc_dev[0];
    bc2_rsqrt = bc_dev[1];
  }
  const int2 ch = chunks[blockIdx.x];
  const TensorMeta m = meta[ch.x];
  const long long begin = ch.y;
  long long end = begin + B200DT_OPT_CHUNK;
  if (end > m.numel) end = m.numel;
  const bool vec = ((reinterpret_cast<uintptr_t>(m.param) | reinterpret_cast<uintptr_t>(m.grad) |
                     (kAdam ? (reinterpret_cast<uintptr_t>(m.exp_avg) |
                               reinterpret_cast<uintptr_t>(m.exp_avg_sq)) : 0)) & 15) == 0 &&
                   (m.shadow == nullptr || (reinterpret_cast<uintptr_t>(m.shadow) & 7) == 0);
  // Vector body over [begin, end_vec), scalar loop over the (< 4 element) tail or, when the
  // pointers are not 16-byte aligned, over the whole chunk.
  const long long end_vec = vec ? begin + ((end - begin) / 4) * 4 : begin;
  for (long long i = begin + threadIdx.x * 4; i < end_vec; i += 256 * 4) {
      float4 p = *reinterpret_cast<float4*>(m.param + i);
      float4 g = *reinterpret_cast<float4*>(m.grad + i);
      float pv[4] = {p.x, p.y, p.z, p.w};
      float gv[4] = {g.x * gscale, g.y * gscale, g.z * gscale, g.w * gscale};
      if (kAdam) {
        float4 a = *reinterpret_cast<float4*>(m.exp_avg + i);
        float4 b = *reinterpret_cast<float4*>(m.exp_avg_sq + i);
        float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float gj = gv[j] + wd * pv[j];
          av[j] = beta1 * av[j] + (1.f - beta1) * gj;
          bv[j] = beta2 * bv[j] + (1.f - beta2) * gj * gj;
          const float denom = sqrtf(bv[j]) * bc2_rsqrt + eps;
          pv[j] -= (lr / bc1) * (av[j] / denom);
        }
        *reinterpret_cast<float4*>(m.exp_avg + i) = make_float4(av[0], av[1], av[2], av[3]);
        *reinterpret_cast<float4*>(m.exp_avg_sq + i) = make_float4(bv[0], bv[1], bv[2], bv[3]);
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) pv[j] -= lr * (gv[j] + wd * pv[j]);
      }
      *reinterpret_cast<float4*>(m.param + i) = make_float4(pv[0], pv[1], pv[2], pv[3]);
      if (m.shadow != nullptr) {
        uint2 u;
        u.x = pack_bf16x2(pv[0], pv[1]);
        u.y = pack_bf16x2(pv[2], pv[3]);
        *reinterpret_cast<uint2*>(m.shadow + i) = u;
      }
      if (zero_grad) *reinterpret_cast<float4*>(m.grad + i) = make_float4(0.f, 0.f, 0.f, 0.f);
  }
  for (long long i = end_vec + threadIdx.x; i < end; i += 256) {
    float p = m.param[i];
    float g = m.grad[i] * gscale;
    if (kAdam) {
      g += wd * p;
      const float a = beta1 * m.exp_avg[i] + (1.f - beta1) * g;
      const float b = beta2 * m.exp_avg_sq[i] + (1.f - beta2) * g * g;
      m.exp_avg[i] = a;
      m.exp_avg_sq[i] = b;
      p -= (lr / bc1) * (a / (sqrtf(b) * bc2_rsqrt + eps));
    } else {
      p -= lr * (g + wd * p);
    }
    m.param[i] = p;
    if (m.shadow != nullptr) m.shadow[i] = __float2bfloat16(p);
    if (zero_grad) m.grad[i] = 0.f;
  }
}

// step += 1 and the two bias corrections derived from it, all on the device: what lets Adam live in a CUDA graph
// (the host-side step count of the plain entry point would be frozen into the captured launch).
__global__ void adam_bump_kernel(float* step, float* bc, float beta1, float beta2) {
  const double t = (double)step[0] + 1.0;
  step[0] = (float)t;
  bc[0] = (float)(1.0 - pow((double)beta1, t));
  bc[1] = (float)(1.0 / sqrt(1.0 - pow((double)beta2, t)));
}

}  // namespace b200

using namespace b200;

extern "C" int b200dt_multi_sgd(const long long* meta, const int* chunks, int n_chunks, float lr,
                                float weight_decay, float grad_scale, int zero_grad, void* stream) {
  if (n_chunks <= 0) return 0;
  multi_opt_kernel<false><<<n_chunks, 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const TensorMeta*>(meta), reinterpret_cast<const int2*>(chunks), lr, 0.f, 0.f,
      0.f, weight_decay, 1.f, 1.f, grad_scale, zero_grad, nullptr);
  return (int)cudaGetLastError();
}

extern "C" int b200dt_multi_adam(const long long* meta, const int* chunks, int n_chunks, float lr,
                                 float beta1, float beta2, float eps, float weight_decay, int step,
                                 float grad_scale, int zero_grad, void* stream) {
  if (n_chunks <= 0) return 0;
  const double bc1 = 1.0 - pow((double)beta1, (double)step);
  const double bc2 = 1.0 - pow((double)beta2, (double)step);
  multi_opt_kernel<true><<<n_chunks, 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const TensorMeta*>(meta), reinterpret_cast<const int2*>(chunks), lr, beta1,
      beta2, eps, weight_decay, (float)bc1, (float)(1.0 / sqrt(bc2)), grad_scale, zero_grad, nullptr);
  return (int)cudaGetLastError();
}

// Adam whose step count lives on the device: step_state = {step, bc1, 1/sqrt(bc2)} (3 floats; step is
// incremented by the call).  Two launches, both capturable in a CUDA graph.
extern "C" int b200dt_multi_adam_capturable(const long long* meta, const int* chunks, int n_chunks, float lr,
                                            float beta1, float beta2, float eps, float weight_decay,
                                            float* step_state, float grad_scale, int zero_grad, void* stream) {
  if (step_state == nullptr) return -3;
  adam_bump_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(step_state, step_state + 1, beta1, beta2);
  if (n_chunks <= 0) return (int)cudaGetLastError();
  multi_opt_kernel<true><<<n_chunks, 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const TensorMeta*>(meta), reinterpret_cast<const int2*>(chunks), lr, beta1,
      beta2, eps, weight_decay, 1.f, 1.f, grad_scale, zero_grad, step_state + 1);
  return (int)cudaGetLastError();
}
