// ssim.cu — fused SSIM forward / backward.
//
// Replaces fusedssim / fusedssim_backward of the reference's fused-ssim submodule
// (submodules/fused-ssim/ssim.cu:187-281 forward kernel, :283-365 backward kernel, :367-444 hosts),
// i.e. the D-SSIM term of the training loss (train.py:155-162, utils/loss_utils.py:56-86):
//   mu_k = G * img_k,  sigma_k^2 = G * img_k^2 - mu_k^2,  sigma_12 = G * (img_1 img_2) - mu_1 mu_2,
//   ssim = (2 mu_1 mu_2 + C1)(2 sigma_12 + C2) / ((mu_1^2 + mu_2^2 + C1)(sigma_1^2 + sigma_2^2 + C2)),
// with G the separable 11x11 Gaussian window (sigma 1.5, fp32 taps of create_window,
// loss_utils.py:46-56) and zero padding ("same").
//
// Design (B200): the stencil is bound by shared-memory issue, not HBM (52 B per pixel-channel over
// both passes, ~130 FMA), so the kernel minimises shared-memory instructions per output:
//  * one CTA = one 32x32 tile of one (batch, channel) plane; the 42x42 input window of both
//    images is staged ONCE and all five moments come from the same registers (the reference reloads
//    and re-synchronises per moment: 19 block.sync() per channel, ssim.cu:213-262);
//  * horizontal pass: each thread owns 4 consecutive outputs of one row — 16 inputs per image
//    arrive as 4 LDS.128, 220 FMAs, 5 STS.128;
//  * vertical pass: each thread owns 4 consecutive rows of one column — 14 conflict-free LDS per
//    moment for 44 FMAs;
//  * the taps are accumulated in the reference's order (left to right, fused multiply-add).
// Planes go to blockIdx.z, so the grid is (W/32) x (H/32) x (B*CH) CTAs of 256 threads.
// The mean over the map (what the loss uses) can be produced in the same launch as deterministic
// per-CTA partial sums + a one-block fp64 finish, so the map never has to be written.
#include <algorithm>
#include <limits>
#include <mutex>
#include "common.cuh"

namespace gsr {

constexpr int SS_T = 32;                 // tile edge
constexpr int SS_R = 5;                  // window radius
constexpr int SS_IN = SS_T + 2 * SS_R;   // 42 staged rows / columns
constexpr int SS_PITCH = 44;             // floats per staged row (16 B aligned rows, 2 pad columns)
constexpr int SS_THREADS = 256;
#ifndef SS_MIN_BLOCKS
#define SS_MIN_BLOCKS 5                   // 5 x 41.7 KB of shared memory, <= 51 registers
#endif

// fp32 taps of gaussian(11, 1.5) / sum (loss_utils.py:42-44)
#define SS_G(k)                                                                                          \
    ((k) == 0 || (k) == 10 ? 0.001028380123898387f                                                      \
     : (k) == 1 || (k) == 9 ? 0.0075987582094967365f                                                    \
     : (k) == 2 || (k) == 8 ? 0.036000773310661316f                                                     \
     : (k) == 3 || (k) == 7 ? 0.10936068743467331f                                                      \
     : (k) == 4 || (k) == 6 ? 0.21300552785396576f                                                      \
                            : 0.26601171493530273f)

// 4 consecutive outputs of an 11-tap correlation over 14 inputs v[0..13]
__device__ __forceinline__ float4 taps4(const float (&v)[16]) {
    float o[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        float acc = 0.f;
#pragma unroll
        for (int k = 0; k < 11; k++) acc = fmaf(SS_G(k), v[j + k], acc);
        o[j] = acc;
    }
    return make_float4(o[0], o[1], o[2], o[3]);
}

__device__ __forceinline__ void load16(float (&v)[16], const float* row) {
    const float4* p = reinterpret_cast<const float4*>(row);
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const float4 t = p[i];
        v[4 * i] = t.x; v[4 * i + 1] = t.y; v[4 * i + 2] = t.z; v[4 * i + 3] = t.w;
    }
}

// vertical pass for one moment: 4 consecutive rows of column `col`
__device__ __forceinline__ void vtaps4(const float (*sx)[SS_T], int row0, int col, float (&o)[4]) {
    float v[14];
#pragma unroll
    for (int i = 0; i < 14; i++) v[i] = sx[row0 + i][col];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        float acc = 0.f;
#pragma unroll
        for (int k = 0; k < 11; k++) acc = fmaf(SS_G(k), v[j + k], acc);
        o[j] = acc;
    }
}

// 1/x to ~1 ulp without the slow path of IEEE division (x is never denormal here: A >= C1, B >= C2)
__device__ __forceinline__ float rcp_newton(float x) {
    float r;
#ifndef GSR_HOST_EMUL
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
#else
    r = 1.0f / x;
#endif
    return fmaf(r, fmaf(-x, r, 1.0f), r);
}

// TRAIN: write the three partial-derivative maps.  `map`, `partial` and `l1_partial` are optional (nullptr);
// l1_partial: per-tile sums of |img1 - img2| (the L1 term of the training loss, from the staged pixels).
template <bool TRAIN>
__device__ __forceinline__ void
ssim_fwd_body(const float* __restrict__ img1, const float* __restrict__ img2, float* __restrict__ map,
              float* __restrict__ dm_dmu1, float* __restrict__ dm_dsigma1_sq, float* __restrict__ dm_dsigma12,
              float* __restrict__ partial, float* __restrict__ l1_partial, int H, int W, float C1, float C2, int crop) {
    __shared__ __align__(16) float s_in[2][SS_IN][SS_PITCH];
    __shared__ __align__(16) float s_x[5][SS_IN][SS_T];
    __shared__ float s_red[SS_THREADS / 32];
    const int tid = threadIdx.x;
    const int x0 = blockIdx.x * SS_T, y0 = blockIdx.y * SS_T;
    const size_t pbase = (size_t)blockIdx.z * H * W;
    const float* p1 = img1 + pbase;
    const float* p2 = img2 + pbase;

    // staging: warp w takes rows w, w+8, ..; lane l takes columns l and l+32 (< 44).  Column validity
    // is per thread, row validity per warp, so an element costs one predicated load and one store.
    {
        const int lane = tid & 31, warp = tid >> 5;
        const int xa = x0 + lane - SS_R, xb = xa + 32;
        const bool ca = xa >= 0 && xa < W, cb = lane < SS_PITCH - 32 && xb >= 0 && xb < W;
#pragma unroll
        for (int r = 0; r < (SS_IN + 7) / 8; r++) {
            const int ly = warp + 8 * r;
            if (ly < SS_IN) {
                const int y = y0 + ly - SS_R;
                const bool rv = y >= 0 && y < H;
                const size_t ro = (size_t)y * W;
                s_in[0][ly][lane] = (rv && ca) ? __ldg(p1 + ro + xa) : 0.f;
                s_in[1][ly][lane] = (rv && ca) ? __ldg(p2 + ro + xa) : 0.f;
                if (lane < SS_PITCH - 32) {
                    s_in[0][ly][lane + 32] = (rv && cb) ? __ldg(p1 + ro + xb) : 0.f;
                    s_in[1][ly][lane + 32] = (rv && cb) ? __ldg(p2 + ro + xb) : 0.f;
                }
            }
        }
    }
    __syncthreads();

    // horizontal pass: 42 rows x 8 runs of 4 outputs
    for (int item = tid; item < SS_IN * (SS_T / 4); item += SS_THREADS) {
        const int row = item >> 3, run = item & 7;
        float a[16], b[16], t[16];
        load16(a, &s_in[0][row][run * 4]);
        load16(b, &s_in[1][row][run * 4]);
        *reinterpret_cast<float4*>(&s_x[0][row][run * 4]) = taps4(a);
        *reinterpret_cast<float4*>(&s_x[1][row][run * 4]) = taps4(b);
#pragma unroll
        for (int i = 0; i < 14; i++) t[i] = __fmul_rn(a[i], a[i]);
        t[14] = t[15] = 0.f;
        *reinterpret_cast<float4*>(&s_x[2][row][run * 4]) = taps4(t);
#pragma unroll
        for (int i = 0; i < 14; i++) t[i] = __fmul_rn(b[i], b[i]);
        *reinterpret_cast<float4*>(&s_x[3][row][run * 4]) = taps4(t);
#pragma unroll
        for (int i = 0; i < 14; i++) t[i] = __fmul_rn(a[i], b[i]);
        *reinterpret_cast<float4*>(&s_x[4][row][run * 4]) = taps4(t);
    }
    __syncthreads();

    // vertical pass: thread = (column, 4 rows)
    const int col = tid & 31, row0 = (tid >> 5) * 4;
    float mu1[4], mu2[4], e11[4], e22[4], e12[4];
    vtaps4(s_x[0], row0, col, mu1);
    vtaps4(s_x[1], row0, col, mu2);
    vtaps4(s_x[2], row0, col, e11);
    vtaps4(s_x[3], row0, col, e22);
    vtaps4(s_x[4], row0, col, e12);

    const int x = x0 + col;
    float local = 0.f, l1 = 0.f;
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const int y = y0 + row0 + j;
        if (x >= W || y >= H) continue;
        if (l1_partial) l1 += fabsf(s_in[0][row0 + j + SS_R][col + SS_R] - s_in[1][row0 + j + SS_R][col + SS_R]);
        const float m1 = mu1[j], m2 = mu2[j];
        const float mu1_sq = m1 * m1, mu2_sq = m2 * m2, mu1_mu2 = m1 * m2;
        const float sigma1_sq = e11[j] - mu1_sq, sigma2_sq = e22[j] - mu2_sq, sigma12 = e12[j] - mu1_mu2;
        const float C = 2.0f * mu1_mu2 + C1;
        const float D = 2.0f * sigma12 + C2;
        const float A = mu1_sq + mu2_sq + C1;
        const float B = sigma1_sq + sigma2_sq + C2;
        const float rA = rcp_newton(A), rB = rcp_newton(B), rAB = rA * rB;
        const float m = (C * D) * rAB;
        const size_t g = pbase + (size_t)y * W + x;
        if (map) map[g] = m;
        if (x >= crop && x < W - crop && y >= crop && y < H - crop) local += m;
        if (TRAIN) {
            // ssim.cu:268-277
            dm_dmu1[g] = 2.0f * rAB * (m2 * (D - C) + m1 * (C * D) * (rB - rA));
            dm_dsigma1_sq[g] = -(C * D) * rAB * rB;
            dm_dsigma12[g] = 2.0f * C * rAB;
        }
    }
    if (partial) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
        if ((tid & 31) == 0) s_red[tid >> 5] = local;
        __syncthreads();
        if (tid == 0) {
            float s = 0.f;
#pragma unroll
            for (int w = 0; w < SS_THREADS / 32; w++) s += s_red[w];
            partial[((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x] = s;
        }
        __syncthreads();
    }
    if (l1_partial) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) l1 += __shfl_xor_sync(0xffffffffu, l1, o);
        if ((tid & 31) == 0) s_red[tid >> 5] = l1;
        __syncthreads();
        if (tid == 0) {
            float s = 0.f;
#pragma unroll
            for (int w = 0; w < SS_THREADS / 32; w++) s += s_red[w];
            l1_partial[((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x] = s;
        }
    }
}

template <bool TRAIN>
__global__ void __launch_bounds__(SS_THREADS, SS_MIN_BLOCKS)
ssim_fwd_kernel(const float* __restrict__ img1, const float* __restrict__ img2, float* __restrict__ map,
                float* __restrict__ dm_dmu1, float* __restrict__ dm_dsigma1_sq, float* __restrict__ dm_dsigma12,
                float* __restrict__ partial, int H, int W, float C1, float C2, int crop) {
    ssim_fwd_body<TRAIN>(img1, img2, map, dm_dmu1, dm_dsigma1_sq, dm_dsigma12, partial, nullptr, H, W, C1, C2, crop);
}

// ---- the training loss of train.py:155-174 in one forward and one backward launch ------------------------------
//   loss = (1 - lambda) * mean|image - gt| + lambda * (1 - mean SSIM(image, gt)) + w_d * mean|(invdepth - mono) * mask|
// Planes 0..CH-1 of the grid are the SSIM tiles above (which also sum |image - gt| from the pixels they staged);
// plane CH holds the 32x32 tiles of the single-channel inverse-depth term.
struct LossArgs {
    const float *image, *gt, *invdepth, *mono, *mask;
    int CH, H, W;
    float C1, C2;
};

// |(invdepth - mono) * mask| summed over one tile -> depth_partial[tile]
__device__ __forceinline__ void loss_depth_fwd_tile(const LossArgs& a, float* __restrict__ depth_partial) {
    __shared__ float s_redd[SS_THREADS / 32];
    const int tid = threadIdx.x, col = tid & 31, row0 = (tid >> 5) * 4;
    const int x = blockIdx.x * SS_T + col;
    float sum = 0.f;
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const int y = blockIdx.y * SS_T + row0 + j;
        if (x >= a.W || y >= a.H) continue;
        const size_t g = (size_t)y * a.W + x;
        float v = __ldg(a.invdepth + g) - __ldg(a.mono + g);
        if (a.mask) v *= __ldg(a.mask + g);
        sum += fabsf(v);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if ((tid & 31) == 0) s_redd[tid >> 5] = sum;
    __syncthreads();
    if (tid == 0) {
        float s = 0.f;
#pragma unroll
        for (int w = 0; w < SS_THREADS / 32; w++) s += s_redd[w];
        depth_partial[blockIdx.y * gridDim.x + blockIdx.x] = s;
    }
}

__global__ void __launch_bounds__(SS_THREADS, SS_MIN_BLOCKS)
loss_fwd_kernel(LossArgs a, float* __restrict__ dm_dmu1, float* __restrict__ dm_dsigma1_sq, float* __restrict__ dm_dsigma12,
                float* __restrict__ ssim_partial, float* __restrict__ l1_partial, float* __restrict__ depth_partial) {
    if ((int)blockIdx.z == a.CH) {       // uniform per CTA
        loss_depth_fwd_tile(a, depth_partial);
        return;
    }
    ssim_fwd_body<true>(a.image, a.gt, nullptr, dm_dmu1, dm_dsigma1_sq, dm_dsigma12, ssim_partial, l1_partial, a.H, a.W,
                        a.C1, a.C2, 0);
}

// out = {loss, L1, mean SSIM, depth L1}: fixed-order fp64 sums of the three partial arrays (deterministic)
__global__ void __launch_bounds__(1024)
loss_finish_kernel(const float* __restrict__ ssim_partial, const float* __restrict__ l1_partial, long long n_img,
                   const float* __restrict__ depth_partial, long long n_depth, double inv_img, double inv_depth,
                   float lambda_dssim, float depth_weight, float* __restrict__ out) {
    __shared__ double s_red[3][32];
    double s[3] = {0.0, 0.0, 0.0};
    for (long long i = threadIdx.x; i < n_img; i += 1024) { s[0] += (double)ssim_partial[i]; s[1] += (double)l1_partial[i]; }
    for (long long i = threadIdx.x; i < n_depth; i += 1024) s[2] += (double)depth_partial[i];
#pragma unroll
    for (int q = 0; q < 3; q++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s[q] += __shfl_xor_sync(0xffffffffu, s[q], o);
        if ((threadIdx.x & 31) == 0) s_red[q][threadIdx.x >> 5] = s[q];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double t[3] = {0.0, 0.0, 0.0};
        for (int q = 0; q < 3; q++)
            for (int w = 0; w < 32; w++) t[q] += s_red[q][w];
        const float ssim = (float)(t[0] * inv_img), l1 = (float)(t[1] * inv_img), dl1 = (float)(t[2] * inv_depth);
        // the reference's expression order (train.py:162,173-174)
        float loss = (1.0f - lambda_dssim) * l1 + lambda_dssim * (1.0f - ssim);
        if (n_depth > 0) loss += depth_weight * dl1;
        out[0] = loss; out[1] = l1; out[2] = ssim; out[3] = dl1;
    }
}

// mean = sum(partials) / count, summed in a fixed order in fp64 (deterministic)
__global__ void __launch_bounds__(1024) ssim_mean_kernel(const float* __restrict__ partial, long long n, double inv_count,
                                                         float* __restrict__ mean_out) {
    __shared__ double s_red[32];
    double s = 0.0;
    for (long long i = threadIdx.x; i < n; i += 1024) s += (double)partial[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < 32; w++) t += s_red[w];
        *mean_out = (float)(t * inv_count);
    }
}

// dL_dimg1 = G * (dL_dmap dm_dmu1) + 2 img1 . G * (dL_dmap dm_dsigma1_sq) + img2 . G * (dL_dmap dm_dsigma12)
// (ssim.cu:309-358).  SCALAR: dL_dmap is the constant  *dL_dmean * inv_count  inside the crop window
// and zero outside — the gradient of map[crop].mean() — so no dL_dmap tensor is read.
// l1_coef != 0 (training loss): the result also gets  l1_coef * *dL_dmean * sign(img1 - img2), the gradient of the
// L1 term; inv_count then carries the SSIM term's factor -lambda / count.
template <bool SCALAR>
__device__ __forceinline__ void
ssim_bwd_body(const float* __restrict__ img1, const float* __restrict__ img2, const float* __restrict__ dL_dmap,
              const float* __restrict__ dL_dmean, float inv_count, const float* __restrict__ dm_dmu1,
              const float* __restrict__ dm_dsigma1_sq, const float* __restrict__ dm_dsigma12,
              float* __restrict__ dL_dimg1, int H, int W, int crop, float l1_coef) {
    __shared__ __align__(16) float s_in[3][SS_IN][SS_PITCH];
    __shared__ __align__(16) float s_x[3][SS_IN][SS_T];
    const int tid = threadIdx.x;
    const int x0 = blockIdx.x * SS_T, y0 = blockIdx.y * SS_T;
    const size_t pbase = (size_t)blockIdx.z * H * W;
    const float gscale = SCALAR ? __ldg(dL_dmean) * inv_count : 0.f;
    const float l1_scale = (SCALAR && l1_coef != 0.f) ? __ldg(dL_dmean) * l1_coef : 0.f;

    {
        const int lane = tid & 31, warp = tid >> 5;
#pragma unroll
        for (int r = 0; r < (SS_IN + 7) / 8; r++) {
            const int ly = warp + 8 * r;
            if (ly >= SS_IN) continue;
            const int y = y0 + ly - SS_R;
            const bool rv = y >= 0 && y < H;
#pragma unroll
            for (int h = 0; h < 2; h++) {
                const int lx = lane + 32 * h;
                if (lx >= SS_PITCH) continue;
                const int x = x0 + lx - SS_R;
                float a = 0.f, b = 0.f, c = 0.f;
                if (rv && lx < SS_IN && x >= 0 && x < W) {
                    const size_t g = pbase + (size_t)y * W + x;
                    float dl;
                    if (SCALAR) dl = (x >= crop && x < W - crop && y >= crop && y < H - crop) ? gscale : 0.f;
                    else dl = __ldg(dL_dmap + g);
                    a = __fmul_rn(__ldg(dm_dmu1 + g), dl);
                    b = __fmul_rn(__ldg(dm_dsigma1_sq + g), dl);
                    c = __fmul_rn(__ldg(dm_dsigma12 + g), dl);
                }
                s_in[0][ly][lx] = a;
                s_in[1][ly][lx] = b;
                s_in[2][ly][lx] = c;
            }
        }
    }
    __syncthreads();

    for (int item = tid; item < SS_IN * (SS_T / 4); item += SS_THREADS) {
        const int row = item >> 3, run = item & 7;
        float v[16];
#pragma unroll
        for (int q = 0; q < 3; q++) {
            load16(v, &s_in[q][row][run * 4]);
            *reinterpret_cast<float4*>(&s_x[q][row][run * 4]) = taps4(v);
        }
    }
    __syncthreads();

    const int col = tid & 31, row0 = (tid >> 5) * 4;
    float ga[4], gb[4], gc[4];
    vtaps4(s_x[0], row0, col, ga);
    vtaps4(s_x[1], row0, col, gb);
    vtaps4(s_x[2], row0, col, gc);
    const int x = x0 + col;
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const int y = y0 + row0 + j;
        if (x >= W || y >= H) continue;
        const size_t g = pbase + (size_t)y * W + x;
        const float pix1 = __ldg(img1 + g), pix2 = __ldg(img2 + g);
        float d = ga[j];
        d += pix1 * 2.0f * gb[j];
        d += pix2 * gc[j];
        if (l1_coef != 0.f) d += l1_scale * (float)((pix1 > pix2) - (pix1 < pix2));      // torch.abs backward: sign(x)
        dL_dimg1[g] = d;
    }
}

template <bool SCALAR>
__global__ void __launch_bounds__(SS_THREADS, SS_MIN_BLOCKS)
ssim_bwd_kernel(const float* __restrict__ img1, const float* __restrict__ img2, const float* __restrict__ dL_dmap,
                const float* __restrict__ dL_dmean, float inv_count, const float* __restrict__ dm_dmu1,
                const float* __restrict__ dm_dsigma1_sq, const float* __restrict__ dm_dsigma12,
                float* __restrict__ dL_dimg1, int H, int W, int crop) {
    ssim_bwd_body<SCALAR>(img1, img2, dL_dmap, dL_dmean, inv_count, dm_dmu1, dm_dsigma1_sq, dm_dsigma12, dL_dimg1, H, W, crop, 0.f);
}

// dL_dimage = dL_dloss * ((1 - lambda) / N * sign(image - gt) - lambda / N * dSSIM/dimage) on planes 0..CH-1,
// dL_dinvdepth = dL_dloss * w_d / (H W) * sign((invdepth - mono) * mask) * mask on plane CH.
__global__ void __launch_bounds__(SS_THREADS, SS_MIN_BLOCKS)
loss_bwd_kernel(LossArgs a, const float* __restrict__ dL_dloss, float ssim_coef, float l1_coef, float depth_coef,
                const float* __restrict__ dm_dmu1, const float* __restrict__ dm_dsigma1_sq,
                const float* __restrict__ dm_dsigma12, float* __restrict__ dL_dimage, float* __restrict__ dL_dinvdepth) {
    if ((int)blockIdx.z == a.CH) {       // uniform per CTA
        const int tid = threadIdx.x, col = tid & 31, row0 = (tid >> 5) * 4;
        const int x = blockIdx.x * SS_T + col;
        const float gd = __ldg(dL_dloss) * depth_coef;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int y = blockIdx.y * SS_T + row0 + j;
            if (x >= a.W || y >= a.H) continue;
            const size_t g = (size_t)y * a.W + x;
            const float m = a.mask ? __ldg(a.mask + g) : 1.0f;
            const float v = (__ldg(a.invdepth + g) - __ldg(a.mono + g)) * m;
            dL_dinvdepth[g] = gd * (float)((v > 0.f) - (v < 0.f)) * m;
        }
        return;
    }
    ssim_bwd_body<true>(a.image, a.gt, nullptr, dL_dloss, ssim_coef, dm_dmu1, dm_dsigma1_sq, dm_dsigma12, dL_dimage, a.H, a.W,
                        0, l1_coef);
}

// several CTAs per SM only fit with the large shared-memory carveout
template <typename K>
static void ssim_prefer_smem(K kernel) {
    cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
}
static void ssim_init_once() {
    static const bool done = [] {
        ssim_prefer_smem(ssim_fwd_kernel<true>);
        ssim_prefer_smem(ssim_fwd_kernel<false>);
        ssim_prefer_smem(ssim_bwd_kernel<true>);
        ssim_prefer_smem(ssim_bwd_kernel<false>);
        ssim_prefer_smem(loss_fwd_kernel);
        ssim_prefer_smem(loss_bwd_kernel);
        return true;
    }();
    (void)done;
}

static long long ssim_tiles(int B, int CH, int H, int W) {
    return (long long)B * CH * ((H + SS_T - 1) / SS_T) * ((W + SS_T - 1) / SS_T);
}

static int ssim_check(const GsrSsimParams* p, const char* who) {
    if (!p || p->B < 0 || p->CH < 0 || p->H < 0 || p->W < 0 || p->crop < 0 ||
        (long long)p->B * p->CH > 65535 || (long long)p->H * p->W >= (1LL << 31)) {
        set_error("%s: bad dimensions (B*CH <= 65535, H*W < 2^31)", who);
        return GSR_ERR_INVALID_ARG;
    }
    if ((long long)p->B * p->CH * p->H * p->W > 0 && (!p->img1 || !p->img2)) {
        set_error("%s: null image pointer", who);
        return GSR_ERR_INVALID_ARG;
    }
    return GSR_OK;
}

// ---- allocation-free variant (gsr_ssim_forward_noalloc) --------------------------------------------------
// The per-tile partial sums of the mean live in a file-scope device array instead of caller-provided scratch, so
// the call needs no cudaMalloc (CUDA-graph capture, serving loops, harnesses that forbid allocation).  Every GPU
// that loads the library holds these GSR_SSIM_NOALLOC_MAX_TILES floats (256 KB) until the process exits.
static __device__ __align__(256) float g_ssim_static_partials[GSR_SSIM_NOALLOC_MAX_TILES];

// In host code the array's name is a host-side stand-in, not a device address: the address a kernel can use comes
// from cudaGetSymbolAddress, looked up once per device with that device current.
static int ssim_static_partials(float** out) {
    static std::mutex mu;
    static float* ptr[64] = {};
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= 64) {
        set_error("gsr_ssim_forward_noalloc: cudaGetDevice failed or device index out of range: %s", cudaGetErrorString(e));
        return GSR_ERR_CUDA;
    }
    std::lock_guard<std::mutex> lock(mu);
    if (ptr[dev] == nullptr) {
        void* q = nullptr;
        e = cudaGetSymbolAddress(&q, g_ssim_static_partials);
        if (e != cudaSuccess || q == nullptr) {
            set_error("gsr_ssim_forward_noalloc: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
            return GSR_ERR_CUDA;
        }
        ptr[dev] = static_cast<float*>(q);
    }
    *out = ptr[dev];
    return GSR_OK;
}

}  // namespace gsr

extern "C" int64_t gsr_ssim_partials(int32_t B, int32_t CH, int32_t H, int32_t W) {
    return (int64_t)gsr::ssim_tiles(B, CH, H, W);
}

extern "C" int gsr_ssim_forward(const GsrSsimParams* p, float* ssim_map, float* dm_dmu1, float* dm_dsigma1_sq,
                                float* dm_dsigma12, float* mean_out, float* partials, void* stream) {
    using namespace gsr;
    if (int rc = ssim_check(p, "gsr_ssim_forward")) return rc;
    const bool train = dm_dmu1 || dm_dsigma1_sq || dm_dsigma12;
    if (train && !(dm_dmu1 && dm_dsigma1_sq && dm_dsigma12)) {
        set_error("gsr_ssim_forward: pass all three derivative maps or none");
        return GSR_ERR_INVALID_ARG;
    }
    if (mean_out && !partials) {
        set_error("gsr_ssim_forward: mean_out needs the partials scratch (gsr_ssim_partials floats)");
        return GSR_ERR_INVALID_ARG;
    }
    cudaStream_t st = (cudaStream_t)stream;
    ssim_init_once();
    const long long tiles = ssim_tiles(p->B, p->CH, p->H, p->W);
    if (tiles > 0) {
        const dim3 grid((p->W + SS_T - 1) / SS_T, (p->H + SS_T - 1) / SS_T, p->B * p->CH);
        float* part = mean_out ? partials : nullptr;
        if (train)
            GSR_LAUNCH(ssim_fwd_kernel<true>, grid, SS_THREADS, 0, st, p->img1, p->img2, ssim_map, dm_dmu1, dm_dsigma1_sq,
                       dm_dsigma12, part, p->H, p->W, p->C1, p->C2, p->crop);
        else
            GSR_LAUNCH(ssim_fwd_kernel<false>, grid, SS_THREADS, 0, st, p->img1, p->img2, ssim_map, dm_dmu1, dm_dsigma1_sq,
                       dm_dsigma12, part, p->H, p->W, p->C1, p->C2, p->crop);
    }
    if (mean_out) {
        const long long ch = std::max(0, p->H - 2 * p->crop), cw = std::max(0, p->W - 2 * p->crop);
        const double count = (double)p->B * p->CH * ch * cw;
        // an empty map has mean NaN, as torch's .mean() of an empty tensor
        const double inv = count > 0 ? 1.0 / count : std::numeric_limits<double>::quiet_NaN();
        GSR_LAUNCH(ssim_mean_kernel, 1, 1024, 0, st, (const float*)partials, tiles, inv, mean_out);
    }
    return check_stage("ssim forward", false, st);
}

extern "C" int gsr_ssim_backward(const GsrSsimParams* p, const float* dL_dmap, const float* dL_dmean,
                                 const float* dm_dmu1, const float* dm_dsigma1_sq, const float* dm_dsigma12,
                                 float* dL_dimg1, void* stream) {
    using namespace gsr;
    if (int rc = ssim_check(p, "gsr_ssim_backward")) return rc;
    const long long tiles = ssim_tiles(p->B, p->CH, p->H, p->W);
    if (tiles == 0) return GSR_OK;
    if (!dm_dmu1 || !dm_dsigma1_sq || !dm_dsigma12 || !dL_dimg1 || (!dL_dmap && !dL_dmean)) {
        set_error("gsr_ssim_backward: null pointer (derivative maps come from a train-mode forward)");
        return GSR_ERR_INVALID_ARG;
    }
    cudaStream_t st = (cudaStream_t)stream;
    ssim_init_once();
    const dim3 grid((p->W + SS_T - 1) / SS_T, (p->H + SS_T - 1) / SS_T, p->B * p->CH);
    if (dL_dmap) {
        GSR_LAUNCH(ssim_bwd_kernel<false>, grid, SS_THREADS, 0, st, p->img1, p->img2, dL_dmap, dL_dmean, 0.f, dm_dmu1,
                   dm_dsigma1_sq, dm_dsigma12, dL_dimg1, p->H, p->W, p->crop);
    } else {
        const long long ch = std::max(0, p->H - 2 * p->crop), cw = std::max(0, p->W - 2 * p->crop);
        const double count = (double)p->B * p->CH * ch * cw;
        const float inv = count > 0 ? (float)(1.0 / count) : 0.f;
        GSR_LAUNCH(ssim_bwd_kernel<true>, grid, SS_THREADS, 0, st, p->img1, p->img2, dL_dmap, dL_dmean, inv, dm_dmu1,
                   dm_dsigma1_sq, dm_dsigma12, dL_dimg1, p->H, p->W, p->crop);
    }
    return check_stage("ssim backward", false, st);
}

// Same work as gsr_ssim_forward with mean_out, but the partial sums go to the library's own static device array:
// no cudaMalloc / cudaFree / allocating call anywhere on this path.  Every tile's slot is rewritten by each call and
// only the first `tiles` slots are read, so nothing has to be cleared between calls; the array is shared by every
// call on the device, so two calls that overlap on the GPU (different streams, no ordering) race.
extern "C" int gsr_ssim_forward_noalloc(const GsrSsimParams* p, float* ssim_map, float* dm_dmu1, float* dm_dsigma1_sq,
                                        float* dm_dsigma12, float* mean_out, void* stream) {
    using namespace gsr;
    if (int rc = ssim_check(p, "gsr_ssim_forward_noalloc")) return rc;
    if (!mean_out) {
        set_error("gsr_ssim_forward_noalloc: mean_out is required (without it the forward needs no scratch: use gsr_ssim_forward)");
        return GSR_ERR_INVALID_ARG;
    }
    const bool train = dm_dmu1 || dm_dsigma1_sq || dm_dsigma12;
    if (train && !(dm_dmu1 && dm_dsigma1_sq && dm_dsigma12)) {
        set_error("gsr_ssim_forward_noalloc: pass all three derivative maps or none");
        return GSR_ERR_INVALID_ARG;
    }
    const long long tiles = ssim_tiles(p->B, p->CH, p->H, p->W);
    if (tiles > (long long)GSR_SSIM_NOALLOC_MAX_TILES) {
        set_error("gsr_ssim_forward_noalloc: %lld tiles of 32x32 exceed the static scratch of %d (GSR_SSIM_NOALLOC_MAX_TILES); "
                  "use gsr_ssim_forward with caller-provided partials", tiles, (int)GSR_SSIM_NOALLOC_MAX_TILES);
        return GSR_ERR_CAPACITY;
    }
    float* part = nullptr;
    if (int rc = ssim_static_partials(&part)) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    ssim_init_once();
    if (tiles > 0) {
        const dim3 grid((p->W + SS_T - 1) / SS_T, (p->H + SS_T - 1) / SS_T, p->B * p->CH);
        if (train)
            GSR_LAUNCH(ssim_fwd_kernel<true>, grid, SS_THREADS, 0, st, p->img1, p->img2, ssim_map, dm_dmu1, dm_dsigma1_sq,
                       dm_dsigma12, part, p->H, p->W, p->C1, p->C2, p->crop);
        else
            GSR_LAUNCH(ssim_fwd_kernel<false>, grid, SS_THREADS, 0, st, p->img1, p->img2, ssim_map, dm_dmu1, dm_dsigma1_sq,
                       dm_dsigma12, part, p->H, p->W, p->C1, p->C2, p->crop);
    }
    const long long ch = std::max(0, p->H - 2 * p->crop), cw = std::max(0, p->W - 2 * p->crop);
    const double count = (double)p->B * p->CH * ch * cw;
    const double inv = count > 0 ? 1.0 / count : std::numeric_limits<double>::quiet_NaN();
    GSR_LAUNCH(ssim_mean_kernel, 1, 1024, 0, st, (const float*)part, tiles, inv, mean_out);
    return check_stage("ssim forward (no-alloc)", false, st);
}


// ---------------------------------------------------------------------------------------------------------------
// Training loss (train.py:155-174): C-ABI.
static int loss_check(const GsrLossParams* p, const char* who) {
    if (!p || p->CH <= 0 || p->CH > 65534 || p->H <= 0 || p->W <= 0 || (long long)p->H * p->W >= (1LL << 31)) {
        gsr::set_error("%s: bad dimensions", who);
        return GSR_ERR_INVALID_ARG;
    }
    if (!p->image || !p->gt || ((p->invdepth == nullptr) != (p->mono_invdepth == nullptr)) || (p->depth_mask && !p->invdepth)) {
        gsr::set_error("%s: image and gt are required; invdepth and mono_invdepth come together", who);
        return GSR_ERR_INVALID_ARG;
    }
    return GSR_OK;
}

extern "C" int64_t gsr_loss_partials(int32_t CH, int32_t H, int32_t W) {
    return 2 * (int64_t)gsr::ssim_tiles(1, CH, H, W) + (int64_t)gsr::ssim_tiles(1, 1, H, W);
}

extern "C" int gsr_loss_forward(const GsrLossParams* p, float* dm_dmu1, float* dm_dsigma1_sq, float* dm_dsigma12,
                                float* out4, float* partials, void* stream) {
    using namespace gsr;
    if (int rc = loss_check(p, "gsr_loss_forward")) return rc;
    if (!dm_dmu1 || !dm_dsigma1_sq || !dm_dsigma12 || !out4 || !partials) {
        set_error("gsr_loss_forward: null output pointer");
        return GSR_ERR_INVALID_ARG;
    }
    cudaStream_t st = (cudaStream_t)stream;
    ssim_init_once();
    const long long n_img = ssim_tiles(1, p->CH, p->H, p->W), n_depth = p->invdepth ? ssim_tiles(1, 1, p->H, p->W) : 0;
    const LossArgs a{p->image, p->gt, p->invdepth, p->mono_invdepth, p->depth_mask, p->CH, p->H, p->W, p->C1, p->C2};
    const dim3 grid((p->W + SS_T - 1) / SS_T, (p->H + SS_T - 1) / SS_T, p->CH + (p->invdepth ? 1 : 0));
    float* l1p = partials + n_img;
    float* dp = partials + 2 * n_img;
    GSR_LAUNCH(loss_fwd_kernel, grid, SS_THREADS, 0, st, a, dm_dmu1, dm_dsigma1_sq, dm_dsigma12, partials, l1p, dp);
    const double inv_img = 1.0 / ((double)p->CH * p->H * p->W), inv_depth = 1.0 / ((double)p->H * p->W);
    GSR_LAUNCH(loss_finish_kernel, 1, 1024, 0, st, (const float*)partials, (const float*)l1p, n_img, (const float*)dp, n_depth,
               inv_img, inv_depth, p->lambda_dssim, p->depth_weight, out4);
    return check_stage("training loss forward", false, st);
}

extern "C" int gsr_loss_backward(const GsrLossParams* p, const float* dL_dloss, const float* dm_dmu1,
                                 const float* dm_dsigma1_sq, const float* dm_dsigma12, float* dL_dimage,
                                 float* dL_dinvdepth, void* stream) {
    using namespace gsr;
    if (int rc = loss_check(p, "gsr_loss_backward")) return rc;
    if (!dL_dloss || !dm_dmu1 || !dm_dsigma1_sq || !dm_dsigma12 || !dL_dimage || (p->invdepth && !dL_dinvdepth)) {
        set_error("gsr_loss_backward: null pointer");
        return GSR_ERR_INVALID_ARG;
    }
    cudaStream_t st = (cudaStream_t)stream;
    ssim_init_once();
    const LossArgs a{p->image, p->gt, p->invdepth, p->mono_invdepth, p->depth_mask, p->CH, p->H, p->W, p->C1, p->C2};
    const dim3 grid((p->W + SS_T - 1) / SS_T, (p->H + SS_T - 1) / SS_T, p->CH + (p->invdepth ? 1 : 0));
    const double inv_img = 1.0 / ((double)p->CH * p->H * p->W), inv_depth = 1.0 / ((double)p->H * p->W);
    GSR_LAUNCH(loss_bwd_kernel, grid, SS_THREADS, 0, st, a, dL_dloss, (float)(-(double)p->lambda_dssim * inv_img),
               (float)((1.0 - (double)p->lambda_dssim) * inv_img), (float)((double)p->depth_weight * inv_depth), dm_dmu1,
               dm_dsigma1_sq, dm_dsigma12, dL_dimage, dL_dinvdepth);
    return check_stage("training loss backward", false, st);
}
