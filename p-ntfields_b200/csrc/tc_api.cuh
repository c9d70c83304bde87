// Tensor-core (tcgen05) precision mode entry points used by capi.cu.
#pragma once
#include <cstddef>
#include <cstdint>

namespace pntf {

size_t tc_workspace_bytes(int op, int E, int64_t N, int d, size_t budget_bytes);

int tc_forward(int mode, int grad_flavour, const float* params, const float* B, const float* coords, int E, int64_t N,
               int d, float* tau, float* dtau, float* ltau, void* workspace, size_t workspace_bytes, void* stream);

// with_grad = 0: Model.Loss forward; 1: forward + parameter gradients
int tc_loss(int with_grad, int variant, const float* params, const float* B, const float* coords, const float* speed,
            int E, int64_t N, int d, float alpha, float beta, float gamma, float* diff, float* loss_out,
            float* grad_params, void* workspace, size_t workspace_bytes, void* stream);

}  // namespace pntf
