// preprocess.cu — forward per-Gaussian stage (kernel 1 of the path) and markVisible.
//
// Replaces preprocessCUDA (DGR/cuda_rasterizer/forward.cu:222-351) and checkFrustum
// (rasterizer_impl.cu:56-73).  HBM-bound streaming kernel: 20 B read + 8 B written for
// every Gaussian, 224 B read + 72 B written for visible ones.
// SH coefficients (81 % of the bytes) are moved as coalesced float4 rows through a
// padded shared-memory transpose; only visible Gaussians' rows are fetched.
#include "common.cuh"
#include "gs_math.cuh"
#include "render_common.cuh"

namespace gsr {

constexpr int PRE_THREADS = 128;
constexpr int PRE_WARPS = PRE_THREADS / 32;
constexpr int SH_ROW = 48;            // floats per Gaussian at degree 3
constexpr int SH_ROW_PAD = SH_ROW + 1; // conflict-free per-lane row reads
constexpr int SPLIT_OFF = 96;            // split SH storage: the rest block follows the 32 x 3 dc block

// SH -> RGB for one colour channel; `sh` points at coefficient 0 of this Gaussian in
// shared memory with channel stride 3.  Summation order of forward.cu:36-66.
// `lo` points at coefficient 0 of the channel, `sh` is the base for coefficients 1.. (sh[3k] = coefficient k);
// dense rows: lo == sh.  Split storage: lo in the dc block, sh = rest row - 3.
__device__ __forceinline__ float sh_channel(const float* lo, const float* sh, int deg, float x, float y, float z) {
    float result = GSR_SH_C0 * lo[0];
    if (deg > 0) {
        result = result - GSR_SH_C1 * y * sh[3] + GSR_SH_C1 * z * sh[6] - GSR_SH_C1 * x * sh[9];
        if (deg > 1) {
            float xx = x * x, yy = y * y, zz = z * z;
            float xy = x * y, yz = y * z, xz = x * z;
            result = result +
                     GSR_SH_C2_0 * xy * sh[12] +
                     GSR_SH_C2_1 * yz * sh[15] +
                     GSR_SH_C2_2 * (2.0f * zz - xx - yy) * sh[18] +
                     GSR_SH_C2_3 * xz * sh[21] +
                     GSR_SH_C2_4 * (xx - yy) * sh[24];
            if (deg > 2) {
                result = result +
                         GSR_SH_C3_0 * y * (3.0f * xx - yy) * sh[27] +
                         GSR_SH_C3_1 * xy * z * sh[30] +
                         GSR_SH_C3_2 * y * (4.0f * zz - xx - yy) * sh[33] +
                         GSR_SH_C3_3 * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * sh[36] +
                         GSR_SH_C3_4 * x * (4.0f * zz - xx - yy) * sh[39] +
                         GSR_SH_C3_5 * z * (xx - yy) * sh[42] +
                         GSR_SH_C3_6 * x * (xx - 3.0f * yy) * sh[45];
            }
        }
    }
    return result + 0.5f;
}

__global__ void __launch_bounds__(PRE_THREADS)
preprocess_fwd_kernel(GsrParams p, float focal_x, float focal_y, int gx, int gy,
                      GsRecord* __restrict__ rec, uint2* __restrict__ rects,
                      uint64_t* __restrict__ depth_keys, int32_t* __restrict__ radii,
                      int32_t* __restrict__ counters) {
    __shared__ __align__(16) float sh_smem[PRE_WARPS][32 * SH_ROW_PAD];
    __shared__ float s_view[16], s_proj[16], s_cam[3];
    __shared__ uint32_t s_visible, s_tiles;
    if (threadIdx.x == 0) { s_visible = 0; s_tiles = 0; }
    if (threadIdx.x < 16) {
        s_view[threadIdx.x] = p.viewmatrix[threadIdx.x];
        s_proj[threadIdx.x] = p.projmatrix[threadIdx.x];
    }
    if (threadIdx.x < 3) s_cam[threadIdx.x] = p.campos[threadIdx.x];
    __syncthreads();

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int idx = blockIdx.x * PRE_THREADS + threadIdx.x;
    const int warp_base = blockIdx.x * PRE_THREADS + warp * 32;
    const bool in_range = idx < p.P;

    bool visible = false;
    float3 p_orig = {0.f, 0.f, 0.f};
    float depth = 0.f, my_radius = 0.f, opacity_aa = 0.f;
    float2 pix = {0.f, 0.f};
    float3 conic = {0.f, 0.f, 0.f};
    uint32_t minx = 0, miny = 0, maxx = 0, maxy = 0;

    if (in_range) {
        p_orig = make_float3(p.means3D[3 * idx], p.means3D[3 * idx + 1], p.means3D[3 * idx + 2]);
        // every per-Gaussian input is requested up front: behind the tests below the four loads were four dependent
        // round trips to HBM per warp (a culled Gaussian now costs 32 B more, a visible one the same)
        float3 sc = {0.f, 0.f, 0.f};
        float4 rot = {0.f, 0.f, 0.f, 0.f};
        if (p.cov3D_precomp == nullptr) {
            sc = make_float3(p.scales[3 * idx], p.scales[3 * idx + 1], p.scales[3 * idx + 2]);
            rot = reinterpret_cast<const float4*>(p.rotations)[idx];
        }
        const float o_raw = p.opacities[idx];
        // in_frustum, auxiliary.h:164-190 (only the z test culls in this fork)
        float3 p_view = transform_point_4x3(p_orig, s_view);
        if (p_view.z <= 0.2f) {
            if (p.prefiltered) {
                printf("Point is filtered although prefiltered is set. This shouldn't happen!");
                __trap();
            }
        } else {
            float4 p_hom = transform_point_4x4(p_orig, s_proj);
            float p_w = 1.0f / (p_hom.w + 0.0000001f);
            float3 p_proj = {p_hom.x * p_w, p_hom.y * p_w, p_hom.z * p_w};

            float cov3D[6];
            if (p.cov3D_precomp != nullptr) {
#pragma unroll
                for (int k = 0; k < 6; k++) cov3D[k] = p.cov3D_precomp[6 * idx + k];
            } else {
                if (p.raw_inputs & GSR_RAW_SCALES) sc = act_scale(sc);
                if (p.raw_inputs & GSR_RAW_ROTATIONS) rot = act_rotation(rot, nullptr);
                compute_cov3D(sc, p.scale_modifier, rot, cov3D);
            }
            float3 cov = compute_cov2D(p_orig, focal_x, focal_y, p.tanfovx, p.tanfovy, cov3D, s_view, nullptr);

            constexpr float h_var = 0.3f;
            const float det_cov = cov.x * cov.z - cov.y * cov.y;
            cov.x += h_var;
            cov.z += h_var;
            const float det_cov_plus_h_cov = cov.x * cov.z - cov.y * cov.y;
            float h_convolution_scaling = 1.0f;
            if (p.antialiasing)
                h_convolution_scaling = sqrtf(fmaxf(0.000025f, det_cov / det_cov_plus_h_cov));
            const float det = det_cov_plus_h_cov;
            if (det != 0.0f) {
                float det_inv = 1.f / det;
                conic = make_float3(cov.z * det_inv, -cov.y * det_inv, cov.x * det_inv);
                float mid = 0.5f * (cov.x + cov.z);
                float lambda1 = mid + sqrtf(fmaxf(0.1f, mid * mid - det));
                float lambda2 = mid - sqrtf(fmaxf(0.1f, mid * mid - det));
                my_radius = ceilf(3.f * sqrtf(fmaxf(lambda1, lambda2)));
                pix = make_float2(ndc2pix(p_proj.x, p.W), ndc2pix(p_proj.y, p.H));
                get_rect(pix.x, pix.y, (int)my_radius, gx, gy, minx, miny, maxx, maxy);
                if ((maxx - minx) * (maxy - miny) != 0) {
                    visible = true;
                    depth = p_view.z;
                    opacity_aa = ((p.raw_inputs & GSR_RAW_OPACITIES) ? act_opacity(o_raw) : o_raw) * h_convolution_scaling;
                }
            }
        }
    }

    // ---- colour: SH -> RGB through a coalesced shared-memory transpose ----
    float rgb[3] = {0.f, 0.f, 0.f};
    int clamped_bits = 0;
    if (p.colors_precomp != nullptr) {
        if (visible) {
            rgb[0] = p.colors_precomp[3 * idx];
            rgb[1] = p.colors_precomp[3 * idx + 1];
            rgb[2] = p.colors_precomp[3 * idx + 2];
        }
    } else {
        const unsigned vis_mask = __ballot_sync(0xffffffffu, visible);
        if (vis_mask != 0) {
            const int L = 3 * (p.D + 1) * (p.D + 1);     // floats used per Gaussian
            const int row = 3 * p.M;                      // floats stored per Gaussian
            float* ws = sh_smem[warp];
            const float* src = p.shs + (size_t)warp_base * row;
            const bool split = p.sh_dc != nullptr;
            const int rrow = row - 3;                     // split storage: floats per rest row
            if (split) {
                // split storage: coefficient 0 from sh_dc [P][3], coefficients 1.. from shs [P][M-1][3].  The two
                // blocks of the warp's 32 Gaussians are contiguous in global memory and are kept FLAT in shared
                // memory (dc block, then the rest block at SPLIT_OFF): a lane reads its rest row at stride
                // rrow (odd: conflict-free), so no transposition is needed.
                const int nvalid = min(32, p.P - warp_base);
                const float* dc = p.sh_dc + (size_t)warp_base * 3;
                const float* rest = p.shs + (size_t)warp_base * rrow;
                const bool al = ((reinterpret_cast<uintptr_t>(p.sh_dc) | reinterpret_cast<uintptr_t>(p.shs)) & 15) == 0 &&
                                (nvalid & 3) == 0;
                if (al && __popc(vis_mask) >= 12) {       // mostly visible: whole blocks, 16 B accesses
                    const int n0 = (nvalid * 3) >> 2, n1 = (nvalid * rrow) >> 2;
                    for (int i = lane; i < n0; i += 32)
                        reinterpret_cast<float4*>(ws)[i] = __ldg(reinterpret_cast<const float4*>(dc) + i);
                    for (int i = lane; i < n1; i += 32)
                        reinterpret_cast<float4*>(ws + SPLIT_OFF)[i] = __ldg(reinterpret_cast<const float4*>(rest) + i);
                } else {                                  // sparse: only the rows of visible Gaussians
                    for (int i = lane; i < 32 * 3; i += 32) {
                        const int g = i / 3;
                        if ((vis_mask >> g) & 1u) ws[i] = __ldg(dc + i);
                    }
                    const int Lr = L - 3;
                    for (unsigned m = vis_mask; m != 0; m &= m - 1) {
                        const int g = __ffs(m) - 1;
                        for (int k = lane; k < Lr; k += 32) ws[SPLIT_OFF + g * rrow + k] = __ldg(rest + (size_t)g * rrow + k);
                    }
                }
            } else if ((L & 3) == 0 && (row & 3) == 0 && (reinterpret_cast<uintptr_t>(p.shs) & 15) == 0) {
                const int L4 = L >> 2;
                for (int i = lane; i < 32 * L4; i += 32) {
                    int g = i / L4, k = i - g * L4;
                    if ((vis_mask >> g) & 1u) {
                        float4 v = __ldg(reinterpret_cast<const float4*>(src + (size_t)g * row) + k);
                        float* d = ws + g * SH_ROW_PAD + 4 * k;
                        d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
                    }
                }
            } else {
                for (int i = lane; i < 32 * L; i += 32) {
                    int g = i / L, k = i - g * L;
                    if ((vis_mask >> g) & 1u) ws[g * SH_ROW_PAD + k] = __ldg(src + (size_t)g * row + k);
                }
            }
            __syncwarp();
            if (visible) {
                // dir = (pos - campos) / length, forward.cu:25-27
                float dx = p_orig.x - s_cam[0], dy = p_orig.y - s_cam[1], dz = p_orig.z - s_cam[2];
                float len = sqrtf(dx * dx + dy * dy + dz * dz);
                dx = dx / len; dy = dy / len; dz = dz / len;
                const float* lo = split ? ws + lane * 3 : ws + lane * SH_ROW_PAD;
                const float* hi = split ? ws + SPLIT_OFF + lane * rrow - 3 : lo;
#pragma unroll
                for (int c = 0; c < 3; c++) {
                    float v = sh_channel(lo + c, hi + c, p.D, dx, dy, dz);
                    if (v < 0.f) clamped_bits |= (1 << c);
                    rgb[c] = fmaxf(v, 0.0f);
                }
            }
        }
    }

    const int tiles = visible ? (int)((maxy - miny) * (maxx - minx)) : 0;
    if (in_range) {
        radii[idx] = visible ? (int)my_radius : 0;
        // invisible Gaussians sort behind every visible one and have an empty rectangle
        depth_keys[idx] = ((uint64_t)(visible ? __float_as_uint(depth) : 0xFFFFFFFFu) << 32) | (uint32_t)idx;
        rects[idx] = visible ? make_uint2(minx | (miny << 16), maxx | (maxy << 16)) : make_uint2(0u, 0u);
        if (visible) {
            GsRecord r;
            r.q0 = make_float4(pix.x, pix.y, conic.x, conic.y);
            r.q1 = make_float4(conic.z, opacity_aa, rgb[0], rgb[1]);
            const float2 ext = cull_extent(conic.x, conic.y, conic.z, opacity_aa);     // render_common.cuh
            r.q2 = make_float4(rgb[2], 1.f / depth, ext.x, ext.y);
            r.q3 = make_float4(__int_as_float(clamped_bits), __int_as_float(tiles), depth, __int_as_float((int)my_radius));
            rec[idx] = r;
        }
    }
    // num_rendered (the reference's InclusiveSum total, rasterizer_impl.cu:309-313) and the number of visible
    // Gaussians, one pair of global adds per CTA (a CTA's tiles fit 32 bits: 128 x 65535 x 2048 does not, but the
    // sum is only meaningful below 2^31 and the 64-bit total still flags the overflow)
    const unsigned vis = __ballot_sync(0xffffffffu, visible);
    uint32_t tsum = (uint32_t)tiles;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) tsum += __shfl_xor_sync(0xffffffffu, tsum, o);
    if (lane == 0 && vis != 0) {
        atomicAdd(&s_visible, (uint32_t)__popc(vis));
        atomicAdd(&s_tiles, tsum);
    }
    __syncthreads();
    if (threadIdx.x == 0 && s_visible != 0) {
        atomicAdd(reinterpret_cast<unsigned long long*>(counters + CNT_RENDERED), (unsigned long long)s_tiles);
        atomicAdd(reinterpret_cast<uint32_t*>(counters + CNT_VISIBLE), s_visible);
    }
}

__global__ void mark_visible_kernel(int P, const float* __restrict__ means,
                                    const float* __restrict__ view, uint8_t* __restrict__ present) {
    int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    float3 p_orig = make_float3(means[3 * idx], means[3 * idx + 1], means[3 * idx + 2]);
    float3 p_view = transform_point_4x3(p_orig, view);
    present[idx] = p_view.z > 0.2f ? 1 : 0;
}

int launch_preprocess(const GsrParams& p, GeomState g, Scratch s, int32_t* radii, cudaStream_t st) {
    const float focal_y = p.H / (2.0f * p.tanfovy);   // rasterizer_impl.cu:253-254
    const float focal_x = p.W / (2.0f * p.tanfovx);
    const int gx = (p.W + TILE_X - 1) / TILE_X, gy = (p.H + TILE_Y - 1) / TILE_Y;
    const int blocks = (p.P + PRE_THREADS - 1) / PRE_THREADS;
    StageTimer tm(GSR_STAGE_PREPROCESS, st);
    GSR_CUDA_TRY(cudaMemsetAsync(s.counters, 0, 8 * sizeof(int32_t), st));
    GSR_LAUNCH(preprocess_fwd_kernel, blocks, PRE_THREADS, 0, st, p, focal_x, focal_y, gx, gy, g.rec, s.rects,
                                                          s.depth_a, radii, s.counters);
    return check_stage("preprocess", p.debug, st);
}

int launch_mark_visible(int P, const float* means, const float* view, uint8_t* present, cudaStream_t st) {
    GSR_LAUNCH(mark_visible_kernel, (P + 255) / 256, 256, 0, st, P, means, view, present);
    return check_stage("markVisible", false, st);
}

}  // namespace gsr
