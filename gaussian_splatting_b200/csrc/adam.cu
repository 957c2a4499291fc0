// adam.cu — fused Adam step over one parameter tensor of the Gaussian model, restricted to the
// Gaussians that were visible in the step.
//
// The fork probes for `SparseGaussianAdam` (train.py:41-45, scene/gaussian_model.py:25-28,246-251) and calls
// `optimizer.step(visible, N)` (train.py:240-247) but ships no implementation.  Semantics here:
// exactly torch.optim.Adam's update (bias-corrected, no weight decay / amsgrad), which is what the
// fork runs by default (gaussian_model.py:245), applied to the rows whose Gaussian is visible; rows
// of invisible Gaussians (zero gradient) keep parameter and moments untouched.  With an all-true
// mask it reproduces torch.optim.Adam (tests/test_adam_gpu.py).
// One pass: 16 B read + 12 B written per element (torch's foreach Adam makes ~4 passes).
#include <algorithm>
#include "common.cuh"

namespace gsr {

template <bool MASKED>
__global__ void __launch_bounds__(256)
adam_update_kernel(float* __restrict__ param, const float* __restrict__ grad, float* __restrict__ exp_avg,
                   float* __restrict__ exp_avg_sq, const uint8_t* __restrict__ visible, float lr_over_bc1,
                   float one_minus_b1, float b2, float one_minus_b2, float eps, float inv_bc2_sqrt, long long total,
                   int M) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride) {
        if (MASKED && !visible[i / M]) continue;
        const float g = grad[i];
        float m = exp_avg[i], v = exp_avg_sq[i];
        m = m + one_minus_b1 * (g - m);                 // lerp_(grad, 1 - beta1)
        v = v * b2 + one_minus_b2 * g * g;              // mul_(beta2).addcmul_(grad, grad, 1 - beta2)
        const float denom = sqrtf(v) * inv_bc2_sqrt + eps;
        param[i] = param[i] - lr_over_bc1 * (m / denom);
        exp_avg[i] = m;
        exp_avg_sq[i] = v;
    }
}

}  // namespace gsr

extern "C" int gsr_adam_update(float* param, const float* grad, float* exp_avg, float* exp_avg_sq,
                               const uint8_t* visible, double lr, double beta1, double beta2, double eps,
                               double bias_correction1, double bias_correction2_sqrt, int64_t N, int32_t M,
                               void* stream) {
    using namespace gsr;
    if (N < 0 || M <= 0 || (N > 0 && (!param || !grad || !exp_avg || !exp_avg_sq))) {
        set_error("gsr_adam_update: bad arguments");
        return GSR_ERR_INVALID_ARG;
    }
    const long long total = (long long)N * M;
    if (total == 0) return GSR_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const int blocks = (int)std::min<long long>((total + 255) / 256, 148LL * 16);
    // hyper-parameters are combined in double (as torch does in Python) and rounded to float once
    const float lr_over_bc1 = (float)(lr / bias_correction1), inv_bc2_sqrt = (float)(1.0 / bias_correction2_sqrt);
    const float omb1 = (float)(1.0 - beta1), omb2 = (float)(1.0 - beta2), b2f = (float)beta2, epsf = (float)eps;
    if (visible)
        GSR_LAUNCH(adam_update_kernel<true>, blocks, 256, 0, st, param, grad, exp_avg, exp_avg_sq, visible, lr_over_bc1,
                                                        omb1, b2f, omb2, epsf, inv_bc2_sqrt, total, (int)M);
    else
        GSR_LAUNCH(adam_update_kernel<false>, blocks, 256, 0, st, param, grad, exp_avg, exp_avg_sq, visible, lr_over_bc1,
                                                         omb1, b2f, omb2, epsf, inv_bc2_sqrt, total, (int)M);
    return check_stage("adam update", false, st);
}
