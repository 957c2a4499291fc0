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

__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, float lr_over_bc1, float one_minus_b1,
                                         float b2, float one_minus_b2, float eps, float inv_bc2_sqrt) {
    m = m + one_minus_b1 * (g - m);                 // lerp_(grad, 1 - beta1)
    v = v * b2 + one_minus_b2 * g * g;              // mul_(beta2).addcmul_(grad, grad, 1 - beta2)
    const float denom = sqrtf(v) * inv_bc2_sqrt + eps;
    p = p - lr_over_bc1 * (m / denom);
}

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
        float p = param[i], m = exp_avg[i], v = exp_avg_sq[i];
        adam_one(p, g, m, v, lr_over_bc1, one_minus_b1, b2, one_minus_b2, eps, inv_bc2_sqrt);
        param[i] = p;
        exp_avg[i] = m;
        exp_avg_sq[i] = v;
    }
}

// 16-byte path (total % 4 == 0, 16 B aligned tensors, total < 2^31): one thread owns 4 consecutive floats; the
// visibility bytes of the (up to 4) Gaussians they belong to are fetched together with the data, so a thread
// has five independent loads in flight instead of a dependent chain, and one division serves four elements.
template <bool MASKED>
__global__ void __launch_bounds__(256)
adam_update_vec4_kernel(float4* __restrict__ param, const float4* __restrict__ grad, float4* __restrict__ exp_avg,
                        float4* __restrict__ exp_avg_sq, const uint8_t* __restrict__ visible, float lr_over_bc1,
                        float one_minus_b1, float b2, float one_minus_b2, float eps, float inv_bc2_sqrt, int total4,
                        int M) {
    const int stride = gridDim.x * blockDim.x;
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < total4; t += stride) {
        bool vis[4] = {true, true, true, true};
        if (MASKED) {
            int q = (4 * t) / M, r = 4 * t - q * M;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                vis[j] = visible[q] != 0;
                if (++r >= M) { r = 0; q++; }
            }
            if (!(vis[0] || vis[1] || vis[2] || vis[3])) continue;
        }
        const float4 g4 = grad[t];
        float4 p4 = param[t], m4 = exp_avg[t], v4 = exp_avg_sq[t];
        float* p = reinterpret_cast<float*>(&p4);
        float* m = reinterpret_cast<float*>(&m4);
        float* v = reinterpret_cast<float*>(&v4);
        const float* g = reinterpret_cast<const float*>(&g4);
#pragma unroll
        for (int j = 0; j < 4; j++)
            if (vis[j]) adam_one(p[j], g[j], m[j], v[j], lr_over_bc1, one_minus_b1, b2, one_minus_b2, eps, inv_bc2_sqrt);
        param[t] = p4;           // invisible elements are written back unchanged
        exp_avg[t] = m4;
        exp_avg_sq[t] = v4;
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
    const auto misaligned = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) != 0; };
    if ((total & 3) == 0 && total < (1LL << 31) && !misaligned(param) && !misaligned(grad) && !misaligned(exp_avg) &&
        !misaligned(exp_avg_sq)) {
        const int total4 = (int)(total >> 2);
        const int blocks4 = (int)std::min<long long>((total4 + 255) / 256, 148LL * 16);
#define ADAM4(MASK) GSR_LAUNCH(adam_update_vec4_kernel<MASK>, blocks4, 256, 0, st, reinterpret_cast<float4*>(param), \
                               reinterpret_cast<const float4*>(grad), reinterpret_cast<float4*>(exp_avg), \
                               reinterpret_cast<float4*>(exp_avg_sq), visible, lr_over_bc1, omb1, b2f, omb2, epsf, \
                               inv_bc2_sqrt, total4, (int)M)
        if (visible) ADAM4(true);
        else ADAM4(false);
#undef ADAM4
        return check_stage("adam update", false, st);
    }
    if (visible)
        GSR_LAUNCH(adam_update_kernel<true>, blocks, 256, 0, st, param, grad, exp_avg, exp_avg_sq, visible, lr_over_bc1,
                                                        omb1, b2f, omb2, epsf, inv_bc2_sqrt, total, (int)M);
    else
        GSR_LAUNCH(adam_update_kernel<false>, blocks, 256, 0, st, param, grad, exp_avg, exp_avg_sq, visible, lr_over_bc1,
                                                         omb1, b2f, omb2, epsf, inv_bc2_sqrt, total, (int)M);
    return check_stage("adam update", false, st);
}
