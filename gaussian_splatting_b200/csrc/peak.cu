// peak.cu — measurement hook: the FP32 FMA rate this GPU reaches at its current clocks, for the roofline of the
// issue-bound blend kernels ("query at run time").  No counterpart in the reference.
#include "common.cuh"

namespace gsr {

constexpr int PEAK_THREADS = 256;
constexpr int PEAK_CHAINS = 8;

// 8 independent scalar FFMA chains per thread; nothing else in the loop.  The result is stored only when it equals
// an impossible value, which keeps the compiler from dropping the chains.
__global__ void __launch_bounds__(PEAK_THREADS)
fp32_peak_kernel(int iters, float seed, float* __restrict__ sink) {
    float a[PEAK_CHAINS];
#pragma unroll
    for (int c = 0; c < PEAK_CHAINS; c++) a[c] = seed + (float)(threadIdx.x + c);
    const float m = 0.999999f, b = 1e-7f;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 4; u++)
#pragma unroll
            for (int c = 0; c < PEAK_CHAINS; c++) a[c] = fmaf(a[c], m, b);
    }
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < PEAK_CHAINS; c++) s += a[c];
    if (s == -123456.789f) sink[0] = s;
}

}  // namespace gsr

using namespace gsr;

extern "C" int gsr_debug_fp32_peak(int32_t blocks, int32_t iters, float* sink, double* flops, void* stream) {
    if (blocks <= 0 || iters <= 0 || !sink) { set_error("gsr_debug_fp32_peak: bad arguments"); return GSR_ERR_INVALID_ARG; }
    cudaStream_t st = (cudaStream_t)stream;
    GSR_LAUNCH(fp32_peak_kernel, blocks, PEAK_THREADS, 0, st, iters, 1.0f, sink);
    if (flops) *flops = 2.0 * (double)blocks * PEAK_THREADS * PEAK_CHAINS * 4.0 * (double)iters;
    return check_stage("fp32 peak", false, st);
}
