// Linear (y = x W^T + b, optional ReLU) as a C++ autograd node over this library's tcgen05 GEMMs.
//
// Why native: nn.DataParallel-style training runs the replicas' backward passes on one autograd thread per
// device.  With a Python autograd.Function every node of every replica takes the GIL, so eight replicas of a toy MLP
// serialise on the host (13 ms per step against torch's 7.7 ms, profiles/toy_bench_dp_8.json); a C++ node runs them
// truly in parallel.  Replaces at::linear / addmm behind nn.Linear (/root/reference/ddp_gpus.py:81,
// 01.data_parallel.ipynb:171).  The kernels are reached through their C ABI by address (include/b200dt.h), like the
// reducer reaches b200dt_coll_launch: this extension does not link against libb200dt.so.
#include <atomic>

#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGuard.h>
#include <torch/extension.h>

#include "linear.h"

namespace {

using GemmTnFn = int (*)(const void*, const void*, void*, int, int, int, int, int, int, const float*, int, int,
                         float*, void*);
using GemmNnFn = int (*)(const void*, const void*, void*, int, int, int, int, int, int, int, void*);
using GemmNtFn = int (*)(const void*, const void*, float*, int, int, int, int, int, int, int, void*);
using ReluBwdFn = int (*)(const void*, const void*, void*, float*, long long, int, void*);

GemmTnFn g_tn = nullptr;
GemmNnFn g_nn = nullptr;
GemmNtFn g_nt = nullptr;
ReluBwdFn g_relu = nullptr;
std::atomic<long long> g_launches{0};

void* cur_stream(const at::Tensor& t) {
  return static_cast<void*>(at::cuda::getCurrentCUDAStream(t.device().index()).stream());
}

struct LinearFn : public torch::autograd::Function<LinearFn> {
  // weight: the fp32 master (only an edge of the graph: its gradient is produced here); wb: its bf16 operand copy
  static at::Tensor forward(torch::autograd::AutogradContext* ctx, const at::Tensor& x, const at::Tensor& weight,
                            const at::Tensor& wb, const c10::optional<at::Tensor>& bias, bool relu, bool out_fp32) {
    TORCH_CHECK(g_tn != nullptr, "b200 linear: kernel addresses were not registered");
    const int64_t N = wb.size(0), K = wb.size(1);
    c10::cuda::CUDAGuard guard(x.device());
    at::Tensor x2 = x.reshape({-1, K});
    if (x2.scalar_type() != at::kBFloat16) x2 = x2.to(at::kBFloat16);
    x2 = x2.contiguous();
    const int64_t M = x2.size(0);
    at::Tensor y = at::empty({M, N}, x2.options().dtype(out_fp32 ? at::kFloat : at::kBFloat16));
    const float* bptr = nullptr;
    at::Tensor bias_c;
    if (bias.has_value() && bias->defined()) {
      bias_c = bias->contiguous();
      TORCH_CHECK(bias_c.scalar_type() == at::kFloat, "b200 linear: bias must be fp32");
      bptr = bias_c.data_ptr<float>();
    }
    const int rc = g_tn(x2.data_ptr(), wb.data_ptr(), y.data_ptr(), (int)M, (int)N, (int)K, (int)K, (int)K,
                        (int)y.stride(0), bptr, relu ? 1 : 0, out_fp32 ? 1 : 0, nullptr, cur_stream(x2));
    TORCH_CHECK(rc == 0, "b200dt_gemm_tn_bf16 failed: rc ", rc);
    g_launches.fetch_add(1, std::memory_order_relaxed);
    ctx->save_for_backward({x2, wb, relu ? y : at::Tensor()});
    ctx->saved_data["relu"] = relu;
    ctx->saved_data["has_bias"] = bptr != nullptr;
    ctx->saved_data["x_shape"] = x.sizes().vec();
    ctx->saved_data["x_dtype"] = static_cast<int64_t>(x.scalar_type());
    std::vector<int64_t> out_shape(x.sizes().begin(), x.sizes().end() - 1);
    out_shape.push_back(N);
    return y.view(out_shape);
  }

  static torch::autograd::variable_list backward(torch::autograd::AutogradContext* ctx,
                                                 torch::autograd::variable_list grads) {
    const auto saved = ctx->get_saved_variables();
    const at::Tensor& x2 = saved[0];
    const at::Tensor& wb = saved[1];
    const at::Tensor& y = saved[2];
    const bool relu = ctx->saved_data["relu"].toBool();
    const bool has_bias = ctx->saved_data["has_bias"].toBool();
    const auto x_shape = ctx->saved_data["x_shape"].toIntVector();
    const auto x_dtype = static_cast<at::ScalarType>(ctx->saved_data["x_dtype"].toInt());
    const int64_t N = wb.size(0), K = wb.size(1);
    c10::cuda::CUDAGuard guard(x2.device());
    void* st = cur_stream(x2);
    at::Tensor dy2 = grads[0].reshape({-1, N});
    const int64_t M = dy2.size(0);
    bool want_db = has_bias && ctx->needs_input_grad(3);
    at::Tensor dyb, db;
    if ((relu || want_db) && (!relu || y.scalar_type() == at::kBFloat16)) {
      // ReLU mask and bias gradient in one pass (bf16 in, bf16 out, fp32 column sums)
      at::Tensor src = dy2.scalar_type() == at::kBFloat16 ? dy2.contiguous() : dy2.to(at::kBFloat16).contiguous();
      dyb = relu ? at::empty_like(src) : src;
      if (want_db) db = at::empty({N}, src.options().dtype(at::kFloat));
      const int rc = g_relu(src.data_ptr(), relu ? y.data_ptr() : nullptr, relu ? dyb.data_ptr() : nullptr,
                            want_db ? db.data_ptr<float>() : nullptr, (long long)M, (int)N, st);
      TORCH_CHECK(rc == 0, "b200dt_relu_bwd_bias failed: rc ", rc);
      g_launches.fetch_add(1, std::memory_order_relaxed);
      want_db = false;
    } else {
      if (relu) dy2 = dy2 * (y.reshape({-1, N}) > 0);
      dyb = dy2.to(at::kBFloat16).contiguous();
    }
    at::Tensor dx, dw;
    if (ctx->needs_input_grad(0)) {
      // W [out, in] is consumed in place as the MN-major B operand: no transposed copy
      at::Tensor d = at::empty({M, K}, dyb.options());
      const int rc = g_nn(dyb.data_ptr(), wb.data_ptr(), d.data_ptr(), (int)M, (int)K, (int)N, (int)N, (int)K,
                          (int)d.stride(0), 0, st);
      TORCH_CHECK(rc == 0, "b200dt_gemm_nn_bf16 failed: rc ", rc);
      g_launches.fetch_add(1, std::memory_order_relaxed);
      dx = d.view(x_shape);
      if (x_dtype != at::kBFloat16) dx = dx.to(x_dtype);
    }
    if (ctx->needs_input_grad(1)) {
      dw = at::empty({N, K}, dyb.options().dtype(at::kFloat));
      const int rc = g_nt(dyb.data_ptr(), x2.data_ptr(), dw.data_ptr<float>(), (int)N, (int)K, (int)M, (int)N, (int)K,
                          (int)dw.stride(0), 0, st);
      TORCH_CHECK(rc == 0, "b200dt_gemm_nt_reduce_bf16 failed: rc ", rc);
      g_launches.fetch_add(1, std::memory_order_relaxed);
    }
    if (want_db) db = dy2.to(at::kFloat).sum(0);
    return {dx, dw, at::Tensor(), db, at::Tensor(), at::Tensor()};
  }
};

}  // namespace

void b200_linear_set_fns(uintptr_t tn, uintptr_t nn, uintptr_t nt, uintptr_t relu_bwd) {
  g_tn = reinterpret_cast<GemmTnFn>(tn);
  g_nn = reinterpret_cast<GemmNnFn>(nn);
  g_nt = reinterpret_cast<GemmNtFn>(nt);
  g_relu = reinterpret_cast<ReluBwdFn>(relu_bwd);
}

at::Tensor b200_linear(const at::Tensor& x, const at::Tensor& weight, const at::Tensor& wb,
                       const c10::optional<at::Tensor>& bias, bool relu, bool out_fp32) {
  return LinearFn::apply(x, weight, wb, bias, relu, out_fp32);
}

long long b200_linear_launches() { return g_launches.load(std::memory_order_relaxed); }
