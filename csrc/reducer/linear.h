// C++ autograd Linear over the library's GEMM entry points (csrc/reducer/linear.cpp).
#pragma once
#include <cstdint>

#include <torch/extension.h>

void b200_linear_set_fns(uintptr_t tn, uintptr_t nn, uintptr_t nt, uintptr_t relu_bwd);
at::Tensor b200_linear(const at::Tensor& x, const at::Tensor& weight, const at::Tensor& wb,
                       const c10::optional<at::Tensor>& bias, bool relu, bool out_fp32);
long long b200_linear_launches();
