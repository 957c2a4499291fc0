// exchange.cu — receiver side of the factored gradient exchange of the view-parallel step
// (sender side: the xchg_* branch of preprocess_bwd.cu).
//
// The SH gradient of a Gaussian is an outer product, dL_dsh[k][c] = basis_k(mean - campos) * dL_dRGB[c]
// (computeColorFromSH backward, DGR/cuda_rasterizer/backward.cu:12-146), and the basis depends only on data
// every rank already holds (its copy of the means) plus the sender's camera centre.  So instead of
// all-reducing 3*M SH-gradient floats per Gaussian (192 of the 236 B at M = 16) a rank publishes 16 B
// {dL_dRGB, 0} into its own slot of a [world][P][4] buffer and 48 B of summed records; this kernel rebuilds
//   dL_dsh[g] = sum_r basis(means[g] - campos[r]) x rgb[r][g]
// in a fixed rank order (deterministic, unlike an atomic or ring reduction) and unpacks the records into
// the dense per-input gradient tensors autograd expects.  NVLink bytes per contributing Gaussian per rank:
// 64 instead of 236.  HBM-bound: reads 60 + 16*world B, writes 44 + 12*M B per Gaussian; dL_dsh rows leave
// through shared memory as one bulk store per warp (the rows of 32 consecutive Gaussians are contiguous).
#include "common.cuh"
#include "gs_math.cuh"
#include "async_copy.cuh"

namespace gsr {

constexpr int XE_THREADS = 128;
constexpr int XE_WARPS = XE_THREADS / 32;
constexpr int XE_ROW = 48;               // floats per dL_dsh row at M = 16
constexpr int XE_MAX_WORLD = 64;

// dRGB/dsh_k for the active degree (same polynomials as the forward, forward.cu:20-72); w[k >= (deg+1)^2] = 0
__device__ __forceinline__ void sh_basis(int deg, float3 dir_orig, float (&w)[16]) {
    const float len = sqrtf(dir_orig.x * dir_orig.x + dir_orig.y * dir_orig.y + dir_orig.z * dir_orig.z);
    const float x = dir_orig.x / len, y = dir_orig.y / len, z = dir_orig.z / len;
#pragma unroll
    for (int k = 0; k < 16; k++) w[k] = 0.f;
    w[0] = GSR_SH_C0;
    if (deg > 0) {
        w[1] = -GSR_SH_C1 * y;
        w[2] = GSR_SH_C1 * z;
        w[3] = -GSR_SH_C1 * x;
        if (deg > 1) {
            const float xx = x * x, yy = y * y, zz = z * z;
            const float xy = x * y, yz = y * z, xz = x * z;
            w[4] = GSR_SH_C2_0 * xy;
            w[5] = GSR_SH_C2_1 * yz;
            w[6] = GSR_SH_C2_2 * (2.f * zz - xx - yy);
            w[7] = GSR_SH_C2_3 * xz;
            w[8] = GSR_SH_C2_4 * (xx - yy);
            if (deg > 2) {
                w[9] = GSR_SH_C3_0 * y * (3.f * xx - yy);
                w[10] = GSR_SH_C3_1 * xy * z;
                w[11] = GSR_SH_C3_2 * y * (4.f * zz - xx - yy);
                w[12] = GSR_SH_C3_3 * z * (2.f * zz - 3.f * xx - 3.f * yy);
                w[13] = GSR_SH_C3_4 * x * (4.f * zz - xx - yy);
                w[14] = GSR_SH_C3_5 * z * (xx - yy);
                w[15] = GSR_SH_C3_6 * x * (xx - 3.f * yy);
            }
        }
    }
}

// BULK: 3*M == 48 floats and 16 B aligned dL_dsh -> rows staged in shared memory, one bulk store per warp.
template <bool BULK>
__global__ void __launch_bounds__(XE_THREADS)
exchange_expand_kernel(int P, int D, int M, int world, const float* __restrict__ means3D,
                       float4* __restrict__ records, float4* __restrict__ rgb,
                       const float4* __restrict__ campos, float* __restrict__ dL_dmeans3D,
                       float* __restrict__ dL_dopacity, float* __restrict__ dL_dsh, float* __restrict__ dL_dscales,
                       float* __restrict__ dL_drotations, float* __restrict__ dL_dsh_dc) {
    __shared__ __align__(128) float s_rows[BULK ? XE_WARPS * 32 * XE_ROW : 4];
    __shared__ float4 s_cam[XE_MAX_WORLD];
    if ((int)threadIdx.x < world) s_cam[threadIdx.x] = campos[threadIdx.x];
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int warp_base = blockIdx.x * XE_THREADS + warp * 32;
    const int idx = warp_base + lane;
    const bool mine = idx < P;

    float sh[XE_ROW];
#pragma unroll
    for (int k = 0; k < XE_ROW; k++) sh[k] = 0.f;
    if (mine) {
        const float4 r0 = records[3 * (size_t)idx], r1 = records[3 * (size_t)idx + 1], r2 = records[3 * (size_t)idx + 2];
        // Leave the exchange buffers clean for the next step: only rows that received something are rewritten,
        // so no per-step memset of the whole buffer is needed (the sender skips zeros, multimem.red adds).
        const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
        const auto nz = [](const float4& v) { return v.x != 0.f || v.y != 0.f || v.z != 0.f || v.w != 0.f; };
        if (nz(r0)) records[3 * (size_t)idx] = z4;
        if (nz(r1)) records[3 * (size_t)idx + 1] = z4;
        if (nz(r2)) records[3 * (size_t)idx + 2] = z4;
        if (dL_dmeans3D) {
            dL_dmeans3D[3 * idx] = r0.x; dL_dmeans3D[3 * idx + 1] = r0.y; dL_dmeans3D[3 * idx + 2] = r0.z;
        }
        if (dL_dopacity) dL_dopacity[idx] = r0.w;
        if (dL_dscales) {
            dL_dscales[3 * idx] = r1.x; dL_dscales[3 * idx + 1] = r1.y; dL_dscales[3 * idx + 2] = r1.z;
        }
        if (dL_drotations) reinterpret_cast<float4*>(dL_drotations)[idx] = r2;
        if (dL_dsh) {
            const float3 mean = {means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]};
            // the ranks' slots are fetched four at a time before any of them is used or cleared: with one load per
            // iteration (and the clearing store between two loads) a thread paid the memory latency `world` times
            // in a row — 0.37 ms at 8 ranks for 1.27 GB.  The order of the sum over ranks stays fixed.
            for (int rb = 0; rb < world; rb += 4) {
                float4 c[4];
#pragma unroll
                for (int j = 0; j < 4; j++) c[j] = rb + j < world ? rgb[(size_t)(rb + j) * P + idx] : z4;
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    if (c[j].x == 0.f && c[j].y == 0.f && c[j].z == 0.f) continue;      // (also the slots past `world`)
                    rgb[(size_t)(rb + j) * P + idx] = z4;
                    const float4 cam = s_cam[rb + j];
                    float w[16];
                    sh_basis(D, make_float3(mean.x - cam.x, mean.y - cam.y, mean.z - cam.z), w);
#pragma unroll
                    for (int k = 0; k < 16; k++) {
                        sh[3 * k] = fmaf(w[k], c[j].x, sh[3 * k]);
                        sh[3 * k + 1] = fmaf(w[k], c[j].y, sh[3 * k + 1]);
                        sh[3 * k + 2] = fmaf(w[k], c[j].z, sh[3 * k + 2]);
                    }
                }
            }
        }
    }
    if (!dL_dsh) return;
    const bool split = dL_dsh_dc != nullptr;       // split SH storage: dc rows [P][3] + rest rows [P][M-1][3]
    const int row = 3 * M, rest = row - 3;
    const int nvalid = max(0, min(32, P - warp_base));
    if (BULK && (!split || (nvalid & 3) == 0)) {
        float* ws = s_rows + warp * 32 * XE_ROW;
        if (!split) {
            float4* my = reinterpret_cast<float4*>(ws + lane * XE_ROW);
#pragma unroll
            for (int m = 0; m < XE_ROW / 4; m++) my[m] = make_float4(sh[4 * m], sh[4 * m + 1], sh[4 * m + 2], sh[4 * m + 3]);
        } else {
#pragma unroll
            for (int k = 0; k < 3; k++) ws[lane * 3 + k] = sh[k];
#pragma unroll
            for (int k = 3; k < XE_ROW; k++) ws[96 + lane * (XE_ROW - 3) + (k - 3)] = sh[k];
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0 && nvalid > 0) {
            if (!split) {
                bulk_s2g(dL_dsh + (size_t)warp_base * XE_ROW, smem_u32(ws), (uint32_t)nvalid * XE_ROW * 4u);
            } else {
                bulk_s2g(dL_dsh_dc + (size_t)warp_base * 3, smem_u32(ws), (uint32_t)nvalid * 12u);
                bulk_s2g(dL_dsh + (size_t)warp_base * (XE_ROW - 3), smem_u32(ws + 96), (uint32_t)nvalid * (XE_ROW - 3) * 4u);
            }
            bulk_commit();
            bulk_wait_read0();
        }
    } else if (mine) {
        float* dst = split ? dL_dsh + (size_t)idx * rest : dL_dsh + (size_t)idx * row;
        if (split) {
#pragma unroll
            for (int k = 0; k < 3; k++) dL_dsh_dc[3 * (size_t)idx + k] = sh[k];
        }
#pragma unroll
        for (int k = 0; k < XE_ROW; k++)
            if (k < row && (!split || k >= 3)) dst[split ? k - 3 : k] = sh[k];
    }
}

}  // namespace gsr

extern "C" int gsr_exchange_expand(const GsrParams* p, int32_t world, float* records, float* rgb,
                                   const float* campos, float* dL_dmeans3D, float* dL_dopacity, float* dL_dsh,
                                   float* dL_dscales, float* dL_drotations, float* dL_dsh_dc, void* stream) {
    using namespace gsr;
    if (!p || p->P < 0 || world < 1 || world > XE_MAX_WORLD || p->D < 0 || p->D > 3 || p->M < 0 || p->M > 16 ||
        (p->D + 1) * (p->D + 1) > (p->M > 0 ? p->M : 16)) {
        set_error("gsr_exchange_expand: bad arguments (1 <= world <= %d, D <= 3, (D+1)^2 <= M <= 16)", XE_MAX_WORLD);
        return GSR_ERR_INVALID_ARG;
    }
    if (p->P == 0) return GSR_OK;
    if (!records || !rgb || !campos || (dL_dsh && !p->means3D)) {
        set_error("gsr_exchange_expand: null buffer");
        return GSR_ERR_INVALID_ARG;
    }
    const auto misaligned = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) != 0; };
    if (misaligned(records) || misaligned(rgb) || misaligned(campos) || misaligned(dL_drotations)) {
        set_error("gsr_exchange_expand: buffers must be 16-byte aligned");
        return GSR_ERR_INVALID_ARG;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const int blocks = (p->P + XE_THREADS - 1) / XE_THREADS;
    if (dL_dsh_dc && (!dL_dsh || p->M < 2)) {
        set_error("gsr_exchange_expand: split SH storage needs dL_dsh (rest) and M >= 2");
        return GSR_ERR_INVALID_ARG;
    }
    const bool bulk = dL_dsh && p->M == 16 && !misaligned(dL_dsh) && !misaligned(dL_dsh_dc);
    StageTimer tm(GSR_STAGE_EXCHANGE, st);
#define XE_LAUNCH(B) GSR_LAUNCH(exchange_expand_kernel<B>, blocks, XE_THREADS, 0, st, p->P, p->D, p->M, (int)world, \
                                p->means3D, reinterpret_cast<float4*>(records), reinterpret_cast<float4*>(rgb), \
                                reinterpret_cast<const float4*>(campos), dL_dmeans3D, dL_dopacity, dL_dsh, dL_dscales, \
                                dL_drotations, dL_dsh_dc)
    if (bulk) XE_LAUNCH(true);
    else XE_LAUNCH(false);
#undef XE_LAUNCH
    return check_stage("exchange expand", false, st);
}
