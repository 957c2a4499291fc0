// preprocess_bwd.cu — kernel 4b of the path: fused per-Gaussian backward.
//
// One kernel replaces computeCov2DCUDA + preprocessCUDA backward + the ten torch::zeros
// fills (DGR/cuda_rasterizer/backward.cu:153-290, 296-365, 372-429; rasterize_points.cu:
// 186-204): it consumes the 12-float accumulator written by render_bwd.cu (colour / inverse-depth
// gradients and six moments, converted here to dL_dmean2D, dL_dconic, dL_dopacity) and writes every
// output gradient exactly once (zeros for culled Gaussians), so no zero-init pass is needed.
//
// The tail of computeCov2DCUDA that this fork elided (backward.cu:285-290: dL_dcov3D and the
// covariance / inverse-depth part of dL_dmeans) is implemented from the chain rule of
// cov2D = (J Rw) Sigma (J Rw)^T   and validated against fp64 autograd.
// Like the reference, d(scale) ignores scale_modifier and the 1.3*tanfov clamp only masks
// dL/dt.x, dL/dt.y (x_grad_mul / y_grad_mul).
#include "common.cuh"
#include "gs_math.cuh"
#include "async_copy.cuh"

namespace gsr {

constexpr int PB_THREADS = 128;
#ifndef PB_MIN_BLOCKS
#define PB_MIN_BLOCKS 5
#endif
constexpr int PB_WARPS = PB_THREADS / 32;
constexpr int PB_ROW = 48;
// Shared-memory rows per Gaussian.  Bulk path (3*M % 4 == 0, 16 B aligned pointers): a warp's 32 SH
// rows are contiguous in global memory, so they arrive with ONE cp.async.bulk (TMA engine, no
// register staging) issued before the covariance math and leave with ONE bulk store; rows are dense
// (stride 3*M floats, LDS.128 / STS.128 per lane).  Scalar path: 49-float padded rows.
constexpr int PB_STRIDE_SCALAR = PB_ROW + 1;

__device__ __forceinline__ float sq(float x) { return x * x; }

// Output of one gradient value.  MM = false: plain store.  MM = true: `p` is an NVLink multicast
// address; the value is ADDED to every rank's replica in the switch (multimem.red), zeros are
// skipped (the replicas start zeroed), so the all-reduce of the view-parallel step happens inside
// this kernel's stores.
template <bool MM>
__device__ __forceinline__ void out1(float* p, float v) {
    if (MM) {
#ifndef GSR_HOST_EMUL
        if (v != 0.f) asm volatile("multimem.red.relaxed.sys.global.add.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
#else
        *p += v;                         // CPU emulation: one rank, the "multicast" buffer is the local one
#endif
    } else {
        *p = v;
    }
}
template <bool MM>
__device__ __forceinline__ void out4(float* p, float4 v) {
    if (MM) {
#ifndef GSR_HOST_EMUL
        if (v.x != 0.f || v.y != 0.f || v.z != 0.f || v.w != 0.f)
            asm volatile("multimem.red.relaxed.sys.global.add.v4.f32 [%0], {%1, %2, %3, %4};"
                         ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
#else
        p[0] += v.x; p[1] += v.y; p[2] += v.z; p[3] += v.w;
#endif
    } else {
        *reinterpret_cast<float4*>(p) = v;
    }
}

// Plain 16 B store; MM = true: `p` is a multicast address, which only multimem.* instructions may access
// (PTX ISA: ld/st on a multimem address is undefined) -- multimem.st writes the value to every replica.
template <bool MM>
__device__ __forceinline__ void store4(float* p, float4 v) {
#ifndef GSR_HOST_EMUL
    if (MM) {
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                     ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
        return;
    }
#endif
    *reinterpret_cast<float4*>(p) = v;
}

// auxiliary.h:128-139
__device__ __forceinline__ float3 dnormvdv(float3 v, float3 dv) {
    float sum2 = v.x * v.x + v.y * v.y + v.z * v.z;
    float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
    float3 r;
    r.x = ((+sum2 - v.x * v.x) * dv.x - v.y * v.x * dv.y - v.z * v.x * dv.z) * invsum32;
    r.y = (-v.x * v.y * dv.x + (sum2 - v.y * v.y) * dv.y - v.z * v.y * dv.z) * invsum32;
    r.z = (-v.x * v.z * dv.x - v.y * v.z * dv.y + (sum2 - v.z * v.z) * dv.z) * invsum32;
    return r;
}

// SH backward for one Gaussian (backward.cu:12-146).  `row` is this lane's padded row in
// shared memory holding the coefficients on entry (first 3*(deg+1)^2 floats valid) and the
// coefficient gradients on exit (all `rowlen` floats written).  Returns dL/d(mean).
// STORE = false (factored exchange): only dL/d(mean) is needed, the row is left untouched.
template <bool STORE = true>
__device__ __forceinline__ float3 sh_backward(float* row, int rowlen, int deg, float3 dir_orig,
                                              const float dL_dRGB[3]) {
    const float len = sqrtf(dir_orig.x * dir_orig.x + dir_orig.y * dir_orig.y + dir_orig.z * dir_orig.z);
    const float x = dir_orig.x / len, y = dir_orig.y / len, z = dir_orig.z / len;
    float w[16];                       // dRGB/dsh_k
    float dx[3] = {0.f, 0.f, 0.f}, dy[3] = {0.f, 0.f, 0.f}, dz[3] = {0.f, 0.f, 0.f};
#pragma unroll
    for (int k = 0; k < 16; k++) w[k] = 0.f;
    w[0] = GSR_SH_C0;
#define SHC(k, c) row[(k) * 3 + (c)]
    if (deg > 0) {
        w[1] = -GSR_SH_C1 * y;
        w[2] = GSR_SH_C1 * z;
        w[3] = -GSR_SH_C1 * x;
#pragma unroll
        for (int c = 0; c < 3; c++) {
            dx[c] = -GSR_SH_C1 * SHC(3, c);
            dy[c] = -GSR_SH_C1 * SHC(1, c);
            dz[c] = GSR_SH_C1 * SHC(2, c);
        }
        if (deg > 1) {
            const float xx = x * x, yy = y * y, zz = z * z;
            const float xy = x * y, yz = y * z, xz = x * z;
            w[4] = GSR_SH_C2_0 * xy;
            w[5] = GSR_SH_C2_1 * yz;
            w[6] = GSR_SH_C2_2 * (2.f * zz - xx - yy);
            w[7] = GSR_SH_C2_3 * xz;
            w[8] = GSR_SH_C2_4 * (xx - yy);
#pragma unroll
            for (int c = 0; c < 3; c++) {
                dx[c] += GSR_SH_C2_0 * y * SHC(4, c) + GSR_SH_C2_2 * 2.f * -x * SHC(6, c) +
                         GSR_SH_C2_3 * z * SHC(7, c) + GSR_SH_C2_4 * 2.f * x * SHC(8, c);
                dy[c] += GSR_SH_C2_0 * x * SHC(4, c) + GSR_SH_C2_1 * z * SHC(5, c) +
                         GSR_SH_C2_2 * 2.f * -y * SHC(6, c) + GSR_SH_C2_4 * 2.f * -y * SHC(8, c);
                dz[c] += GSR_SH_C2_1 * y * SHC(5, c) + GSR_SH_C2_2 * 2.f * 2.f * z * SHC(6, c) +
                         GSR_SH_C2_3 * x * SHC(7, c);
            }
            if (deg > 2) {
                w[9] = GSR_SH_C3_0 * y * (3.f * xx - yy);
                w[10] = GSR_SH_C3_1 * xy * z;
                w[11] = GSR_SH_C3_2 * y * (4.f * zz - xx - yy);
                w[12] = GSR_SH_C3_3 * z * (2.f * zz - 3.f * xx - 3.f * yy);
                w[13] = GSR_SH_C3_4 * x * (4.f * zz - xx - yy);
                w[14] = GSR_SH_C3_5 * z * (xx - yy);
                w[15] = GSR_SH_C3_6 * x * (xx - 3.f * yy);
#pragma unroll
                for (int c = 0; c < 3; c++) {
                    dx[c] += (GSR_SH_C3_0 * SHC(9, c) * 3.f * 2.f * xy +
                              GSR_SH_C3_1 * SHC(10, c) * yz +
                              GSR_SH_C3_2 * SHC(11, c) * -2.f * xy +
                              GSR_SH_C3_3 * SHC(12, c) * -3.f * 2.f * xz +
                              GSR_SH_C3_4 * SHC(13, c) * (-3.f * xx + 4.f * zz - yy) +
                              GSR_SH_C3_5 * SHC(14, c) * 2.f * xz +
                              GSR_SH_C3_6 * SHC(15, c) * 3.f * (xx - yy));
                    dy[c] += (GSR_SH_C3_0 * SHC(9, c) * 3.f * (xx - yy) +
                              GSR_SH_C3_1 * SHC(10, c) * xz +
                              GSR_SH_C3_2 * SHC(11, c) * (-3.f * yy + 4.f * zz - xx) +
                              GSR_SH_C3_3 * SHC(12, c) * -3.f * 2.f * yz +
                              GSR_SH_C3_4 * SHC(13, c) * -2.f * xy +
                              GSR_SH_C3_5 * SHC(14, c) * -2.f * yz +
                              GSR_SH_C3_6 * SHC(15, c) * -3.f * 2.f * xy);
                    dz[c] += (GSR_SH_C3_1 * SHC(10, c) * xy +
                              GSR_SH_C3_2 * SHC(11, c) * 4.f * 2.f * yz +
                              GSR_SH_C3_3 * SHC(12, c) * 3.f * (2.f * zz - xx - yy) +
                              GSR_SH_C3_4 * SHC(13, c) * 4.f * 2.f * xz +
                              GSR_SH_C3_5 * SHC(14, c) * (xx - yy));
                }
            }
        }
    }
#undef SHC
    if (STORE) {
        const int ncoef = (deg + 1) * (deg + 1);
        const int nrow = rowlen / 3;
#pragma unroll
        for (int k = 0; k < 16; k++) {
            if (k < nrow) {
                const float wk = k < ncoef ? w[k] : 0.f;
                row[k * 3 + 0] = wk * dL_dRGB[0];
                row[k * 3 + 1] = wk * dL_dRGB[1];
                row[k * 3 + 2] = wk * dL_dRGB[2];
            }
        }
    }
    float3 dL_ddir = {dx[0] * dL_dRGB[0] + dx[1] * dL_dRGB[1] + dx[2] * dL_dRGB[2],
                      dy[0] * dL_dRGB[0] + dy[1] * dL_dRGB[1] + dy[2] * dL_dRGB[2],
                      dz[0] * dL_dRGB[0] + dz[1] * dL_dRGB[1] + dz[2] * dL_dRGB[2]};
    return dnormvdv(dir_orig, dL_ddir);
}

// Where a lane's coefficient row lives in the warp's shared-memory staging area.
//   kind 0: padded rows of PB_STRIDE_SCALAR floats (scalar path)
//   kind 1: dense rows of `rowlen` floats, whole float4s (vector path: LDS.128 / STS.128)
//   kind 2: split storage — 32 dc rows of 3 floats, then from SPLIT_REST_OFF 32 rest rows of rowlen-3 floats
//           (odd stride: conflict-free scalar accesses)
struct RowLayout { int kind; int rowlen; };
constexpr int SPLIT_REST_OFF = 96;

// SH backward for this lane's row, through registers.  STORE: write the coefficient gradients back in place.
template <bool STORE, int KIND>
__device__ __forceinline__ float3 sh_backward_smem(float* ws, int lane, RowLayout lay, int deg, float3 dir_orig,
                                                   const float dL_dRGB[3]) {
    if (KIND == 0) return sh_backward<STORE>(ws + lane * PB_STRIDE_SCALAR, lay.rowlen, deg, dir_orig, dL_dRGB);
    float sh[PB_ROW];
    const int rowlen = lay.rowlen, rest = rowlen - 3;
    float4* row4 = reinterpret_cast<float4*>(ws + lane * rowlen);
    if (KIND == 1) {
#pragma unroll
        for (int m = 0; m < PB_ROW / 4; m++) {
            const float4 v = row4[m];
            sh[4 * m] = v.x; sh[4 * m + 1] = v.y; sh[4 * m + 2] = v.z; sh[4 * m + 3] = v.w;
        }
    } else {
#pragma unroll
        for (int k = 0; k < 3; k++) sh[k] = ws[lane * 3 + k];
#pragma unroll
        for (int k = 3; k < PB_ROW; k++) sh[k] = k < rowlen ? ws[SPLIT_REST_OFF + lane * rest + (k - 3)] : 0.f;
    }
    const float3 r = sh_backward<STORE>(sh, rowlen, deg, dir_orig, dL_dRGB);
    if (STORE) {
        if (KIND == 1) {
            const int n4 = rowlen >> 2;
#pragma unroll
            for (int m = 0; m < PB_ROW / 4; m++)
                if (m < n4) row4[m] = make_float4(sh[4 * m], sh[4 * m + 1], sh[4 * m + 2], sh[4 * m + 3]);
        } else {
#pragma unroll
            for (int k = 0; k < 3; k++) ws[lane * 3 + k] = sh[k];
#pragma unroll
            for (int k = 3; k < PB_ROW; k++)
                if (k < rowlen) ws[SPLIT_REST_OFF + lane * rest + (k - 3)] = sh[k];
        }
    }
    return r;
}

// zero gradient row of a culled Gaussian
template <int KIND>
__device__ __forceinline__ void row_zero(float* ws, int lane, RowLayout lay) {
    const int rowlen = lay.rowlen;
    if (KIND == 0) {
        for (int k = 0; k < rowlen; k++) ws[lane * PB_STRIDE_SCALAR + k] = 0.f;
    } else if (KIND == 1) {
        for (int k = 0; k < (rowlen >> 2); k++) reinterpret_cast<float4*>(ws + lane * rowlen)[k] = make_float4(0.f, 0.f, 0.f, 0.f);
    } else {
        for (int k = 0; k < 3; k++) ws[lane * 3 + k] = 0.f;
        for (int k = 0; k < rowlen - 3; k++) ws[SPLIT_REST_OFF + lane * (rowlen - 3) + k] = 0.f;
    }
}

// cov3D backward (backward.cu:296-365) with GLM indexing m[c][r].
__device__ __forceinline__ void cov3D_backward(float3 scale, float mod, float4 rot, const float* dL_dcov3D,
                                               float3* dL_dscale, float4* dL_drot) {
    const float r = rot.x, x = rot.y, y = rot.z, z = rot.w;
    Mat3 R = quat_to_R(r, x, y, z);
    Mat3 S = mat3_cols(1.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 1.f);
    const float3 s = {mod * scale.x, mod * scale.y, mod * scale.z};
    S.m[0][0] = s.x; S.m[1][1] = s.y; S.m[2][2] = s.z;
    Mat3 M = mat3_mul(S, R);
    Mat3 dL_dSigma = mat3_cols(
        dL_dcov3D[0], 0.5f * dL_dcov3D[1], 0.5f * dL_dcov3D[2],
        0.5f * dL_dcov3D[1], dL_dcov3D[3], 0.5f * dL_dcov3D[4],
        0.5f * dL_dcov3D[2], 0.5f * dL_dcov3D[4], dL_dcov3D[5]);
    Mat3 dL_dM = mat3_mul(M, dL_dSigma);
#pragma unroll
    for (int c = 0; c < 3; c++)
#pragma unroll
        for (int rr = 0; rr < 3; rr++) dL_dM.m[c][rr] *= 2.0f;
    Mat3 Rt = mat3_transpose(R);
    Mat3 dL_dMt = mat3_transpose(dL_dM);
    dL_dscale->x = Rt.m[0][0] * dL_dMt.m[0][0] + Rt.m[0][1] * dL_dMt.m[0][1] + Rt.m[0][2] * dL_dMt.m[0][2];
    dL_dscale->y = Rt.m[1][0] * dL_dMt.m[1][0] + Rt.m[1][1] * dL_dMt.m[1][1] + Rt.m[1][2] * dL_dMt.m[1][2];
    dL_dscale->z = Rt.m[2][0] * dL_dMt.m[2][0] + Rt.m[2][1] * dL_dMt.m[2][1] + Rt.m[2][2] * dL_dMt.m[2][2];
#pragma unroll
    for (int rr = 0; rr < 3; rr++) {
        dL_dMt.m[0][rr] *= s.x;
        dL_dMt.m[1][rr] *= s.y;
        dL_dMt.m[2][rr] *= s.z;
    }
    const Mat3& D = dL_dMt;
    dL_drot->x = 2 * z * (D.m[0][1] - D.m[1][0]) + 2 * y * (D.m[2][0] - D.m[0][2]) + 2 * x * (D.m[1][2] - D.m[2][1]);
    dL_drot->y = 2 * y * (D.m[1][0] + D.m[0][1]) + 2 * z * (D.m[2][0] + D.m[0][2]) + 2 * r * (D.m[1][2] - D.m[2][1]) - 4 * x * (D.m[2][2] + D.m[1][1]);
    dL_drot->z = 2 * x * (D.m[1][0] + D.m[0][1]) + 2 * r * (D.m[2][0] - D.m[0][2]) + 2 * z * (D.m[1][2] + D.m[2][1]) - 4 * y * (D.m[2][2] + D.m[0][0]);
    dL_drot->w = 2 * r * (D.m[0][1] - D.m[1][0]) + 2 * x * (D.m[2][0] + D.m[0][2]) + 2 * y * (D.m[1][2] + D.m[2][1]) - 4 * z * (D.m[1][1] + D.m[0][0]);
}

// SPLIT: split SH storage (GsrParams.sh_dc).  A template parameter so that the dense instantiation carries none
// of the split layout's code (the runtime-flag version cost the dense path 16 %).
template <bool VEC, bool MM, bool SPLIT>
__global__ void __launch_bounds__(PB_THREADS, PB_MIN_BLOCKS)
preprocess_bwd_kernel(GsrParams p, float focal_x, float focal_y, const GsRecord* __restrict__ rec,
                      const int32_t* __restrict__ radii, const float* __restrict__ acc,
                      bool have_invdepth, GsrBackwardIO io, const uint32_t* __restrict__ band_list,
                      const uint32_t* __restrict__ band_ctl, int band) {
    if (MM && band_list != nullptr && (uint32_t)(blockIdx.x * PB_THREADS) >= band_ctl[BAND_CNT + band]) return;
    __shared__ __align__(128) float sh_smem[PB_WARPS][32 * PB_STRIDE_SCALAR];
    __shared__ __align__(8) uint64_t s_bar[PB_WARPS];
    __shared__ float s_view[16], s_proj[16], s_cam[3];
    if (VEC && (threadIdx.x & 31) == 0) {
        mbar_init(smem_u32(&s_bar[threadIdx.x >> 5]), 1);
        mbar_fence_init();
    }
    if (threadIdx.x < 16) {
        s_view[threadIdx.x] = p.viewmatrix[threadIdx.x];
        s_proj[threadIdx.x] = p.projmatrix[threadIdx.x];
    }
    if (threadIdx.x < 3) s_cam[threadIdx.x] = p.campos[threadIdx.x];
    __syncthreads();

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int warp_base = blockIdx.x * PB_THREADS + warp * 32;
    // List mode (multimem reduction, banded): this launch processes the visible Gaussians of one
    // tile-row band, whose indices were compacted into band_list by band_fill_kernel.
    const bool LIST = MM && band_list != nullptr;
    int idx = blockIdx.x * PB_THREADS + threadIdx.x;
    bool in_range = idx < p.P;
    bool visible;
    if (LIST) {
        const uint32_t count = band_ctl[BAND_CNT + band];
        in_range = (uint32_t)idx < count;
        idx = in_range ? (int)band_list[band_ctl[BAND_OFF + band] + (uint32_t)idx] : 0;
        visible = in_range;
    } else {
        visible = in_range && radii[idx] > 0;
    }
    const bool mine = in_range;
    const unsigned vis_mask = __ballot_sync(0xffffffffu, visible);
    const int sh_row = 3 * p.M;                                   // floats per Gaussian
    const int nvalid = max(0, min(32, p.P - warp_base));
    const uint32_t sh_bytes = (uint32_t)(nvalid * sh_row) * 4u;
    // Split SH storage (GsrParams.sh_dc): the warp's 32 dc rows (96 floats) and 32 rest rows are two contiguous
    // blocks; in shared memory the dc block comes first, the rest block follows at SPLIT_REST_OFF.
    constexpr bool split = SPLIT;
    const int rest_row = sh_row - 3;
    // bulk copies need 16 B multiples: a ragged last warp (nvalid % 4 != 0) of the split layout is copied by the lanes
    const bool bulk_ok = !split || (nvalid & 3) == 0;
    if (VEC && LIST) {                                            // gather: one bulk copy per Gaussian row
        const uint32_t bar = smem_u32(&s_bar[warp]);
        if (visible) bulk_g2s(smem_u32(sh_smem[warp] + lane * sh_row), p.shs + (size_t)idx * sh_row, (uint32_t)sh_row * 4u, bar);
        if (lane == 0 && vis_mask != 0) mbar_arrive_expect_tx(bar, (uint32_t)__popc(vis_mask) * (uint32_t)sh_row * 4u);
    } else if (VEC && vis_mask != 0 && lane == 0 && bulk_ok) {    // prefetch this warp's (contiguous) SH rows
        const uint32_t bar = smem_u32(&s_bar[warp]);
        mbar_arrive_expect_tx(bar, sh_bytes);
        if (split) {
            bulk_g2s(smem_u32(sh_smem[warp]), p.sh_dc + (size_t)warp_base * 3, (uint32_t)nvalid * 12u, bar);
            bulk_g2s(smem_u32(sh_smem[warp] + SPLIT_REST_OFF), p.shs + (size_t)warp_base * rest_row,
                     (uint32_t)(nvalid * rest_row) * 4u, bar);
        } else {
            bulk_g2s(smem_u32(sh_smem[warp]), p.shs + (size_t)warp_base * sh_row, sh_bytes, bar);
        }
    }
    float3 dL_dmean = {0.f, 0.f, 0.f};
    float2 dL_dmean2D = {0.f, 0.f};
    float dL_dopacity = 0.f;
    float dL_dcolor[3] = {0.f, 0.f, 0.f};
    float dL_dcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float3 dL_dscale = {0.f, 0.f, 0.f};
    float4 dL_drot = {0.f, 0.f, 0.f, 0.f};
    float3 mean = {0.f, 0.f, 0.f};
    int clamped_bits = 0;

    if (visible) {
        const float4* a4 = reinterpret_cast<const float4*>(acc + (size_t)idx * ACC_FLOATS);
        const float4 a0 = a4[0], a1 = a4[1], a2 = a4[2];
        dL_dcolor[0] = a0.x; dL_dcolor[1] = a0.y; dL_dcolor[2] = a0.z;
        const float dL_dinvdepth = a0.w;
        // moments of h = G * dL_dalpha about the centre (render_bwd.cu): S h, S h dx, S h dy, S h dx^2, S h dx dy, S h dy^2
        const float m0 = a1.x, mx = a1.y, my = a1.z, mxx = a1.w, mxy = a2.x, myy = a2.y;
        const float dL_dopacity_v = m0;                                   // backward.cu:587-609
        clamped_bits = __float_as_int(rec[idx].q3.x);

        mean = make_float3(p.means3D[3 * idx], p.means3D[3 * idx + 1], p.means3D[3 * idx + 2]);
        float cov3D[6];
        float3 sc = {0.f, 0.f, 0.f};
        float4 rot = {0.f, 0.f, 0.f, 0.f};
        float rot_inv_norm = 1.f;
        if (p.cov3D_precomp != nullptr) {
#pragma unroll
            for (int k = 0; k < 6; k++) cov3D[k] = p.cov3D_precomp[6 * idx + k];
        } else {
            sc = make_float3(p.scales[3 * idx], p.scales[3 * idx + 1], p.scales[3 * idx + 2]);
            rot = reinterpret_cast<const float4*>(p.rotations)[idx];
            if (p.raw_inputs & GSR_RAW_SCALES) sc = act_scale(sc);
            if (p.raw_inputs & GSR_RAW_ROTATIONS) rot = act_rotation(rot, &rot_inv_norm);
            compute_cov3D(sc, p.scale_modifier, rot, cov3D);
        }
        Cov2DCtx ctx;
        const float3 cov = compute_cov2D(mean, focal_x, focal_y, p.tanfovx, p.tanfovy, cov3D, s_view, &ctx);
        const float3 t = ctx.t;
        const float x_grad_mul = (ctx.txtz < -ctx.limx || ctx.txtz > ctx.limx) ? 0.f : 1.f;
        const float y_grad_mul = (ctx.tytz < -ctx.limy || ctx.tytz > ctx.limy) ? 0.f : 1.f;

        float c_xx = cov.x, c_xy = cov.y, c_yy = cov.z;
        float3 dL_dconic;
        constexpr float h_var = 0.3f;
        float d_inside_root = 0.f;
        dL_dopacity = dL_dopacity_v;
        const float opacity = (p.raw_inputs & GSR_RAW_OPACITIES) ? act_opacity(p.opacities[idx]) : p.opacities[idx];
        float o_blend = opacity;                     // what the blend multiplied G with (forward.cu:344)
        if (p.antialiasing) {
            const float det_cov = c_xx * c_yy - c_xy * c_xy;
            c_xx += h_var;
            c_yy += h_var;
            const float det_cov_plus_h_cov = c_xx * c_yy - c_xy * c_xy;
            const float h_convolution_scaling = sqrtf(fmaxf(0.000025f, det_cov / det_cov_plus_h_cov));
            const float d_h_convolution_scaling = dL_dopacity_v * opacity;
            dL_dopacity = dL_dopacity_v * h_convolution_scaling;
            o_blend = opacity * h_convolution_scaling;
            d_inside_root = (det_cov / det_cov_plus_h_cov) <= 0.000025f
                                ? 0.f
                                : d_h_convolution_scaling / (2 * h_convolution_scaling);
        } else {
            c_xx += h_var;
            c_yy += h_var;
        }
        float dL_dc_xx = 0, dL_dc_xy = 0, dL_dc_yy = 0;
        if (p.antialiasing) {
            const float x = c_xx, y = c_yy, z = c_xy, w = h_var;
            const float denom_f = d_inside_root / sq(w * w + w * (x + y) + x * y - z * z);
            dL_dc_xx = w * (w * y + y * y + z * z) * denom_f;
            dL_dc_yy = w * (w * x + x * x + z * z) * denom_f;
            dL_dc_xy = -2.f * w * z * (w + x + y) * denom_f;
        }
        const float denom = c_xx * c_yy - c_xy * c_xy;
        // The blend's per-Gaussian constants: conic = inverse of the filtered covariance (forward.cu:291-293),
        // blend opacity o = opacity * AA scaling, viewport scale 0.5 W / 0.5 H.  With dL_dG = o * dL_dalpha:
        //   dL_dmean2D = -0.5 (W, H) * o * conic * (S h dx, S h dy),  dL_dconic = -0.5 o (S h dx^2, S h dx dy, S h dy^2)
        // (backward.cu:569-586 summed over the pixels).
        {
            const float det_inv = 1.f / denom;
            const float ca = c_yy * det_inv, cb = -c_xy * det_inv, cc = c_xx * det_inv;
            const float o = o_blend;
            const float ddelx_dx = 0.5 * p.W, ddely_dy = 0.5 * p.H;
            dL_dmean2D.x = -ddelx_dx * o * (ca * mx + cb * my);
            dL_dmean2D.y = -ddely_dy * o * (cc * my + cb * mx);
            dL_dconic = make_float3(-0.5f * o * mxx, -0.5f * o * mxy, -0.5f * o * myy);
        }
        const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
        if (denom2inv != 0) {
            dL_dc_xx += denom2inv * (-c_yy * c_yy * dL_dconic.x + 2 * c_xy * c_yy * dL_dconic.y + (denom - c_xx * c_yy) * dL_dconic.z);
            dL_dc_yy += denom2inv * (-c_xx * c_xx * dL_dconic.z + 2 * c_xx * c_xy * dL_dconic.y + (denom - c_xx * c_yy) * dL_dconic.x);
            dL_dc_xy += denom2inv * 2 * (c_xy * c_yy * dL_dconic.x - (denom + 2 * c_xy * c_xy) * dL_dconic.y + c_xx * c_xy * dL_dconic.z);
        }

        // ---- restored tail (App. A.6): A = J*Rw, rows A0 = T.m[0][*], A1 = T.m[1][*] ----
        const float dL_da = dL_dc_xx, dL_db = dL_dc_xy, dL_dc = dL_dc_yy;
        const float A0[3] = {ctx.T.m[0][0], ctx.T.m[0][1], ctx.T.m[0][2]};
        const float A1[3] = {ctx.T.m[1][0], ctx.T.m[1][1], ctx.T.m[1][2]};
        dL_dcov[0] = A0[0] * A0[0] * dL_da + A0[0] * A1[0] * dL_db + A1[0] * A1[0] * dL_dc;
        dL_dcov[3] = A0[1] * A0[1] * dL_da + A0[1] * A1[1] * dL_db + A1[1] * A1[1] * dL_dc;
        dL_dcov[5] = A0[2] * A0[2] * dL_da + A0[2] * A1[2] * dL_db + A1[2] * A1[2] * dL_dc;
        dL_dcov[1] = 2 * A0[0] * A0[1] * dL_da + (A0[0] * A1[1] + A0[1] * A1[0]) * dL_db + 2 * A1[0] * A1[1] * dL_dc;
        dL_dcov[2] = 2 * A0[0] * A0[2] * dL_da + (A0[0] * A1[2] + A0[2] * A1[0]) * dL_db + 2 * A1[0] * A1[2] * dL_dc;
        dL_dcov[4] = 2 * A0[2] * A0[1] * dL_da + (A0[1] * A1[2] + A0[2] * A1[1]) * dL_db + 2 * A1[1] * A1[2] * dL_dc;

        const float V[3][3] = {{cov3D[0], cov3D[1], cov3D[2]},
                               {cov3D[1], cov3D[3], cov3D[4]},
                               {cov3D[2], cov3D[4], cov3D[5]}};
        float dL_dA0[3], dL_dA1[3];
#pragma unroll
        for (int k = 0; k < 3; k++) {
            const float VA0 = V[k][0] * A0[0] + V[k][1] * A0[1] + V[k][2] * A0[2];
            const float VA1 = V[k][0] * A1[0] + V[k][1] * A1[1] + V[k][2] * A1[2];
            dL_dA0[k] = 2 * VA0 * dL_da + VA1 * dL_db;
            dL_dA1[k] = 2 * VA1 * dL_dc + VA0 * dL_db;
        }
        // Rw(r,k) = view[4k + r]
        const float dL_dJ00 = s_view[0] * dL_dA0[0] + s_view[4] * dL_dA0[1] + s_view[8] * dL_dA0[2];
        const float dL_dJ02 = s_view[2] * dL_dA0[0] + s_view[6] * dL_dA0[1] + s_view[10] * dL_dA0[2];
        const float dL_dJ11 = s_view[1] * dL_dA1[0] + s_view[5] * dL_dA1[1] + s_view[9] * dL_dA1[2];
        const float dL_dJ12 = s_view[2] * dL_dA1[0] + s_view[6] * dL_dA1[1] + s_view[10] * dL_dA1[2];
        const float tz = 1.f / t.z;
        const float tz2 = tz * tz;
        const float tz3 = tz2 * tz;
        const float dL_dtx = x_grad_mul * -focal_x * tz2 * dL_dJ02;
        const float dL_dty = y_grad_mul * -focal_y * tz2 * dL_dJ12;
        float dL_dtz = -focal_x * tz2 * dL_dJ00 - focal_y * tz2 * dL_dJ11 +
                       (2 * focal_x * t.x) * tz3 * dL_dJ02 + (2 * focal_y * t.y) * tz3 * dL_dJ12;
        if (have_invdepth) dL_dtz -= dL_dinvdepth / (t.z * t.z);
        dL_dmean = transform_vec_4x3_transpose(make_float3(dL_dtx, dL_dty, dL_dtz), s_view);

        // ---- projection term (backward.cu:393-420) ----
        const float* proj = s_proj;
        const float3 m = mean;
        const float4 m_hom = transform_point_4x4(m, proj);
        const float m_w = 1.0f / (m_hom.w + 0.0000001f);
        const float mul1 = (proj[0] * m.x + proj[4] * m.y + proj[8] * m.z + proj[12]) * m_w * m_w;
        const float mul2 = (proj[1] * m.x + proj[5] * m.y + proj[9] * m.z + proj[13]) * m_w * m_w;
        dL_dmean.x += (proj[0] * m_w - proj[3] * mul1) * dL_dmean2D.x + (proj[1] * m_w - proj[3] * mul2) * dL_dmean2D.y;
        dL_dmean.y += (proj[4] * m_w - proj[7] * mul1) * dL_dmean2D.x + (proj[5] * m_w - proj[7] * mul2) * dL_dmean2D.y;
        dL_dmean.z += (proj[8] * m_w - proj[11] * mul1) * dL_dmean2D.x + (proj[9] * m_w - proj[11] * mul2) * dL_dmean2D.y;

        if (p.scales != nullptr) cov3D_backward(sc, p.scale_modifier, rot, dL_dcov, &dL_dscale, &dL_drot);
        // fused activations (GsrParams.raw_inputs): gradients w.r.t. the raw parameters, before any store
        if (p.raw_inputs & GSR_RAW_SCALES) dL_dscale = act_scale_bwd(sc, dL_dscale);
        if (p.raw_inputs & GSR_RAW_ROTATIONS) dL_drot = act_rotation_bwd(rot, rot_inv_norm, dL_drot);
        if (p.raw_inputs & GSR_RAW_OPACITIES) dL_dopacity = act_opacity_bwd(opacity, dL_dopacity);
    }

    // SH rows of the warp in shared memory.  Layouts (RowLayout): dense rows of `sh_row` floats (vector path),
    // dc block + rest block (split storage), or padded 49-float rows (scalar path).
    const bool has_sh = p.shs != nullptr;
    const bool fact = io.xchg_records != nullptr;
    const bool want_sh = fact || io.dL_dsh != nullptr;
    if (has_sh && !want_sh) {
        // dL_dsh not wanted (NULL): the view-direction term of dL_dmeans3D is still due.  Rare call shape
        // (a C-ABI caller; the autograd wrapper always asks for dL_dsh): coefficients straight from global memory.
        if (visible) {
            float row[PB_ROW];
            const int Lc = 3 * (p.D + 1) * (p.D + 1);
#pragma unroll
            for (int k = 0; k < PB_ROW; k++) {
                float v = 0.f;
                if (k < Lc) v = !SPLIT ? __ldg(p.shs + (size_t)idx * sh_row + k)
                                       : (k < 3 ? __ldg(p.sh_dc + (size_t)idx * 3 + k) : __ldg(p.shs + (size_t)idx * rest_row + (k - 3)));
                row[k] = v;
            }
            float dL_dRGB[3];
#pragma unroll
            for (int c = 0; c < 3; c++) dL_dRGB[c] = ((clamped_bits >> c) & 1) ? 0.f : dL_dcolor[c];
            const float3 dir_orig = {mean.x - s_cam[0], mean.y - s_cam[1], mean.z - s_cam[2]};
            const float3 dm = sh_backward<false>(row, sh_row, p.D, dir_orig, dL_dRGB);
            dL_dmean.x += dm.x; dL_dmean.y += dm.y; dL_dmean.z += dm.z;
        }
    }
    if (has_sh && want_sh) {
        float* ws = sh_smem[warp];
        constexpr int KIND = VEC ? (SPLIT ? 2 : 1) : 0;
        const RowLayout lay = {KIND, sh_row};
        float dL_dRGB[3];
#pragma unroll
        for (int c = 0; c < 3; c++) dL_dRGB[c] = ((clamped_bits >> c) & 1) ? 0.f : dL_dcolor[c];
        const float3 dir_orig = {mean.x - s_cam[0], mean.y - s_cam[1], mean.z - s_cam[2]};
        // ---- coefficients in ----
        if (VEC && (LIST || bulk_ok)) {
            if (vis_mask != 0) mbar_wait(smem_u32(&s_bar[warp]), 0);
        } else if (VEC) {                       // ragged last warp of the split layout
            for (int i = lane; i < nvalid * 3; i += 32) ws[i] = __ldg(p.sh_dc + (size_t)warp_base * 3 + i);
            for (int i = lane; i < nvalid * rest_row; i += 32) ws[SPLIT_REST_OFF + i] = __ldg(p.shs + (size_t)warp_base * rest_row + i);
            __syncwarp();
        } else {
            if (vis_mask != 0) {
                const int Lc = 3 * (p.D + 1) * (p.D + 1);
                for (int i = lane; i < 32 * Lc; i += 32) {
                    const int g = i / Lc, k = i - g * Lc;
                    if ((vis_mask >> g) & 1u) {
                        const size_t gi = (size_t)(warp_base + g);
                        ws[g * PB_STRIDE_SCALAR + k] = !split ? __ldg(p.shs + gi * sh_row + k)
                                                       : (k < 3 ? __ldg(p.sh_dc + gi * 3 + k) : __ldg(p.shs + gi * rest_row + (k - 3)));
                    }
                }
            }
            __syncwarp();
        }
        // ---- factored exchange: publish {mean, opacity, scale, rotation} gradients as one 48 B
        // record (summed across ranks) and the clamp-masked dL_dRGB in this rank's slot; dL_dsh is rebuilt by
        // every receiver (exchange.cu).  16 B multimem operations only; culled rows are not touched.
        if (fact) {
            if (visible) {
                const float3 dm = sh_backward_smem<false, KIND>(ws, lane, lay, p.D, dir_orig, dL_dRGB);
                dL_dmean.x += dm.x; dL_dmean.y += dm.y; dL_dmean.z += dm.z;
            }
            if (blockIdx.x == 0 && threadIdx.x == 0)     // this rank's camera centre, into its slot of every replica
                store4<MM>(io.xchg_campos, make_float4(s_cam[0], s_cam[1], s_cam[2], 1.f));
            if (!mine) return;
            if (io.dL_dmeans2D) {
                io.dL_dmeans2D[3 * idx] = dL_dmean2D.x; io.dL_dmeans2D[3 * idx + 1] = dL_dmean2D.y; io.dL_dmeans2D[3 * idx + 2] = 0.f;
            }
            if (visible) {
                float* r = io.xchg_records + 12 * (size_t)idx;
                out4<MM>(r, make_float4(dL_dmean.x, dL_dmean.y, dL_dmean.z, dL_dopacity));
                out4<MM>(r + 4, make_float4(dL_dscale.x, dL_dscale.y, dL_dscale.z, 0.f));
                out4<MM>(r + 8, dL_drot);
                out4<MM>(io.xchg_rgb + 4 * (size_t)idx, make_float4(dL_dRGB[0], dL_dRGB[1], dL_dRGB[2], 0.f));
            }
            return;
        }
        // ---- SH backward in place; dL_dsh rows leave coalesced ----
        if (visible) {
            const float3 dm = sh_backward_smem<true, KIND>(ws, lane, lay, p.D, dir_orig, dL_dRGB);
            dL_dmean.x += dm.x; dL_dmean.y += dm.y; dL_dmean.z += dm.z;
        } else {
            row_zero<KIND>(ws, lane, lay);
        }
        if (VEC && MM && LIST) {
            // 16 B multimem reductions, 12 consecutive lanes per (scattered) Gaussian row
            __syncwarp();
            const int row4 = sh_row >> 2;
            for (int f = lane; f < 32 * row4; f += 32) {
                const int gl = f / row4, k = f - gl * row4;
                const int gid = __shfl_sync(0xffffffffu, idx, gl);
                if ((vis_mask >> gl) & 1u)
                    out4<true>(io.dL_dsh + (size_t)gid * sh_row + 4 * k, reinterpret_cast<const float4*>(ws)[f]);
            }
        } else if (VEC && (MM || !bulk_ok)) {
            // the warp's rows are contiguous in global memory: coalesced 16 B (or, ragged split tail, 4 B) outputs
            __syncwarp();
            float* d0 = split ? io.dL_dsh_dc + (size_t)warp_base * 3 : io.dL_dsh + (size_t)warp_base * sh_row;
            const int n0 = split ? nvalid * 3 : nvalid * sh_row;
            float* d1 = io.dL_dsh + (size_t)warp_base * rest_row;
            const int n1 = split ? nvalid * rest_row : 0;
            if (bulk_ok) {
                for (int f = lane; f < (n0 >> 2); f += 32) out4<MM>(d0 + 4 * f, reinterpret_cast<const float4*>(ws)[f]);
                for (int f = lane; f < (n1 >> 2); f += 32) out4<MM>(d1 + 4 * f, reinterpret_cast<const float4*>(ws + SPLIT_REST_OFF)[f]);
            } else {
                for (int f = lane; f < n0; f += 32) out1<MM>(d0 + f, ws[f]);
                for (int f = lane; f < n1; f += 32) out1<MM>(d1 + f, ws[SPLIT_REST_OFF + f]);
            }
        } else if (VEC) {
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0 && sh_bytes != 0) {
                if (split) {
                    bulk_s2g(io.dL_dsh_dc + (size_t)warp_base * 3, smem_u32(ws), (uint32_t)nvalid * 12u);
                    bulk_s2g(io.dL_dsh + (size_t)warp_base * rest_row, smem_u32(ws + SPLIT_REST_OFF), (uint32_t)(nvalid * rest_row) * 4u);
                } else {
                    bulk_s2g(io.dL_dsh + (size_t)warp_base * sh_row, smem_u32(ws), sh_bytes);
                }
                bulk_commit();
            }
        } else {
            __syncwarp();
            for (int i = lane; i < nvalid * sh_row; i += 32) {
                const int g = i / sh_row, k = i - g * sh_row;
                const size_t gi = (size_t)(warp_base + g);
                float* dst = !split ? io.dL_dsh + gi * sh_row + k
                                    : (k < 3 ? io.dL_dsh_dc + gi * 3 + k : io.dL_dsh + gi * rest_row + (k - 3));
                out1<MM>(dst, ws[g * PB_STRIDE_SCALAR + k]);
            }
        }
    }

    if (!mine) return;
    if (io.dL_dmeans3D) {
        out1<MM>(io.dL_dmeans3D + 3 * idx, dL_dmean.x); out1<MM>(io.dL_dmeans3D + 3 * idx + 1, dL_dmean.y);
        out1<MM>(io.dL_dmeans3D + 3 * idx + 2, dL_dmean.z);
    }
    if (io.dL_dmeans2D) {       // per-view densification statistic: always local
        io.dL_dmeans2D[3 * idx] = dL_dmean2D.x; io.dL_dmeans2D[3 * idx + 1] = dL_dmean2D.y; io.dL_dmeans2D[3 * idx + 2] = 0.f;
    }
    if (io.dL_dopacity) out1<MM>(io.dL_dopacity + idx, dL_dopacity);
    if (io.dL_dcolors) {
#pragma unroll
        for (int k = 0; k < 3; k++) out1<MM>(io.dL_dcolors + 3 * idx + k, dL_dcolor[k]);
    }
    if (io.dL_dcov3D) {
#pragma unroll
        for (int k = 0; k < 6; k++) out1<MM>(io.dL_dcov3D + 6 * idx + k, dL_dcov[k]);
    }
    if (io.dL_dscales) {
        out1<MM>(io.dL_dscales + 3 * idx, dL_dscale.x); out1<MM>(io.dL_dscales + 3 * idx + 1, dL_dscale.y);
        out1<MM>(io.dL_dscales + 3 * idx + 2, dL_dscale.z);
    }
    if (io.dL_drotations) out4<MM>(io.dL_drotations + 4 * (size_t)idx, dL_drot);
    if (VEC && !MM && lane == 0) bulk_wait_read0();      // shared memory must outlive the bulk store's reads
}

// ---- band lists (multimem mode) ------------------------------------------------------------
struct BandLimits { int n; int hi[MAX_BANDS]; };

__device__ __forceinline__ int band_of(const GsRecord* r, int gy, const BandLimits& lim) {
    const float py = r->q0.y;
    const int rad = __float_as_int(r->q3.w);
    const int last_row = min(gy, max(0, (int)((py + rad + TILE_Y - 1) / TILE_Y))) - 1;   // get_rect, gs_math.cuh
    int k = 0;
    while (k < lim.n - 1 && last_row >= lim.hi[k]) k++;
    return k;
}

// Pass 1: per-band counts; culled Gaussians get their (never reduced) dL_dmeans2D zeroed here.
__global__ void __launch_bounds__(256)
band_count_kernel(int P, int gy, BandLimits lim, const GsRecord* __restrict__ rec, const int32_t* __restrict__ radii,
                  float* __restrict__ dL_dmeans2D, uint32_t* __restrict__ ctl) {
    __shared__ uint32_t s_cnt[MAX_BANDS];
    if (threadIdx.x < MAX_BANDS) s_cnt[threadIdx.x] = 0;
    __syncthreads();
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx < P) {
        if (radii[idx] > 0) {
            atomicAdd(&s_cnt[band_of(rec + idx, gy, lim)], 1u);
        } else if (dL_dmeans2D) {
            dL_dmeans2D[3 * idx] = 0.f; dL_dmeans2D[3 * idx + 1] = 0.f; dL_dmeans2D[3 * idx + 2] = 0.f;
        }
    }
    __syncthreads();
    if (threadIdx.x < lim.n && s_cnt[threadIdx.x]) atomicAdd(&ctl[BAND_CNT + threadIdx.x], s_cnt[threadIdx.x]);
}

__global__ void band_offsets_kernel(int n, uint32_t* __restrict__ ctl) {
    uint32_t run = 0;
    for (int k = 0; k < n; k++) {
        ctl[BAND_OFF + k] = run;
        ctl[BAND_CUR + k] = run;
        run += ctl[BAND_CNT + k];
    }
}

// Pass 2: compaction.  One atomic per (block, band); inside a block the ids stay in index order,
// so neighbouring list entries are mostly neighbours in memory.
__global__ void __launch_bounds__(256)
band_fill_kernel(int P, int gy, BandLimits lim, const GsRecord* __restrict__ rec, const int32_t* __restrict__ radii,
                 uint32_t* __restrict__ ctl, uint32_t* __restrict__ list) {
    __shared__ uint32_t s_cnt[MAX_BANDS], s_base[MAX_BANDS];
    if (threadIdx.x < MAX_BANDS) s_cnt[threadIdx.x] = 0;
    __syncthreads();
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const bool vis = idx < P && radii[idx] > 0;
    int k = 0;
    uint32_t slot = 0;
    if (vis) k = band_of(rec + idx, gy, lim);
    // rank inside the block in thread order: warps take turns (8 warps, few bands)
    for (int w = 0; w < 8; w++) {
        if ((threadIdx.x >> 5) == w) {
            for (int kb = 0; kb < lim.n; kb++) {
                const unsigned m = __ballot_sync(0xffffffffu, vis && k == kb);
                if (vis && k == kb) slot = s_cnt[kb] + __popc(m & ((1u << (threadIdx.x & 31)) - 1u));
                __syncwarp();
                if ((threadIdx.x & 31) == 0 && m) s_cnt[kb] += __popc(m);
                __syncwarp();
            }
        }
        __syncthreads();
    }
    if (threadIdx.x < lim.n && s_cnt[threadIdx.x]) s_base[threadIdx.x] = atomicAdd(&ctl[BAND_CUR + threadIdx.x], s_cnt[threadIdx.x]);
    __syncthreads();
    if (vis) list[s_base[k] + slot] = (uint32_t)idx;
}

int launch_band_lists(const GsrParams& p, GeomState g, Scratch s, const int32_t* radii, float* dL_dmeans2D,
                      int bands, const int* band_hi, cudaStream_t st) {
    BandLimits lim;
    lim.n = bands;
    for (int k = 0; k < MAX_BANDS; k++) lim.hi[k] = k < bands ? band_hi[k] : 0x7fffffff;
    const int gy = (p.H + TILE_Y - 1) / TILE_Y;
    const int blocks = (p.P + 255) / 256;
    StageTimer tm(GSR_STAGE_PREPROCESS_BWD, st);
    GSR_CUDA_TRY(cudaMemsetAsync(s.band_ctl, 0, BAND_CTL_WORDS * sizeof(uint32_t), st));
    GSR_LAUNCH(band_count_kernel, blocks, 256, 0, st, p.P, gy, lim, g.rec, radii, dL_dmeans2D, s.band_ctl);
    GSR_LAUNCH(band_offsets_kernel, 1, 1, 0, st, bands, s.band_ctl);
    GSR_LAUNCH(band_fill_kernel, blocks, 256, 0, st, p.P, gy, lim, g.rec, radii, s.band_ctl, s.band_list);
    return check_stage("band lists", p.debug, st);
}

int launch_preprocess_backward(const GsrParams& p, GeomState g, Scratch s, const int32_t* radii,
                               bool have_invdepth, const GsrBackwardIO& io, int band, cudaStream_t st) {
    const float focal_y = p.H / (2.0f * p.tanfovy);
    const float focal_x = p.W / (2.0f * p.tanfovx);
    const int blocks = (p.P + PB_THREADS - 1) / PB_THREADS;
    const uint32_t* blist = band >= 0 ? s.band_list : nullptr;
    StageTimer tm(GSR_STAGE_PREPROCESS_BWD, st);
    // vector path: SH rows are whole float4s (M in {4, 8, 12, 16}) and 16 B aligned
    const bool fact = io.xchg_records != nullptr;
    const auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
    // vector path: SH rows are whole float4s (M in {4, 8, 12, 16}) and 16 B aligned; split storage: any M, both
    // blocks (and both gradient blocks) 16 B aligned
    const bool split = p.sh_dc != nullptr;
    const bool vec = p.shs != nullptr && (fact || io.dL_dsh != nullptr) && al16(p.shs) &&
                     (split ? al16(p.sh_dc) && (fact || (al16(io.dL_dsh) && al16(io.dL_dsh_dc)))
                            : (p.M & 3) == 0 && (fact || al16(io.dL_dsh)));
    const bool mm = io.multimem_reduce != 0;
#define PB_LAUNCH(V, M, S) GSR_LAUNCH((preprocess_bwd_kernel<V, M, S>), blocks, PB_THREADS, 0, st, p, focal_x, focal_y, g.rec, \
                                      radii, s.acc, have_invdepth, io, blist, s.band_ctl, band)
#define PB_LAUNCH2(V, M) do { if (split) PB_LAUNCH(V, M, true); else PB_LAUNCH(V, M, false); } while (0)
    if (vec && mm) PB_LAUNCH2(true, true);
    else if (vec) PB_LAUNCH2(true, false);
    else if (mm) PB_LAUNCH2(false, true);
    else PB_LAUNCH2(false, false);
#undef PB_LAUNCH2
#undef PB_LAUNCH
    return check_stage("preprocess backward", p.debug, st);
}

}  // namespace gsr
