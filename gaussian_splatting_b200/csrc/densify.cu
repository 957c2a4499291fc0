// densify.cu — adaptive density control on the device and the Gaussian file rows
// (8f rank 4).  Replaces the tensor surgery of scene/gaussian_model.py:400-621 (boolean-mask indexing,
// torch.cat, per-group optimizer-state rebuilds: ~150 torch kernels and a host synchronisation per mask) by
//   plan : classify (1 pass over P) -> 3-component scan -> destination rows
//   apply: ONE launch that moves every parameter tensor and both Adam moments (grid.y = tensor)
// Per-element arithmetic lives in densify_core.cuh (shared with the CPU emulation test).
//
// HBM-bound.  Algorithmic bytes: plan 25*P read + 13*P written; apply reads every source
// element once (708 B per Gaussian at SH degree 3 with both moments) and writes 708 B per result row.
#include <algorithm>
#include "common.cuh"
#include "block_scan.cuh"
#include "densify_core.cuh"

namespace gsr {

#ifndef GSR_INDEX32_LIMIT
#define GSR_INDEX32_LIMIT (1LL << 32)     // element counts below this use 32-bit index math (tests lower it)
#endif
constexpr int DP_THREADS = 256;
constexpr int DP_ITEMS = 4;
constexpr int DP_CHUNK = DP_THREADS * DP_ITEMS;    // Gaussians per block of the plan kernels

// plan layout (u32 words): dstA[P] dstB[P] dstC[P] | block sums [3][nb] | device counts [4] | codes (P bytes)
struct PlanView {
    uint32_t *dstA, *dstB, *dstC, *block_sums, *codes_w;
    int32_t* counts;
    uint32_t nb;
};
__host__ __device__ inline uint32_t plan_blocks(uint32_t P) { return (P + DP_CHUNK - 1) / DP_CHUNK; }
static PlanView plan_view(uint32_t* plan, uint32_t P) {
    PlanView v;
    v.nb = plan_blocks(P);
    v.dstA = plan;
    v.dstB = plan + (size_t)P;
    v.dstC = plan + 2 * (size_t)P;
    v.block_sums = plan + 3 * (size_t)P;
    v.counts = reinterpret_cast<int32_t*>(v.block_sums + 3 * (size_t)v.nb);
    v.codes_w = reinterpret_cast<uint32_t*>(v.counts + 4);
    return v;
}

// Pass 1: decision code per Gaussian + per-block counts of the three segments.
__global__ void __launch_bounds__(DP_THREADS)
densify_classify_kernel(GsrDensifyPlanArgs a, uint8_t* __restrict__ codes, uint32_t* __restrict__ block_sums,
                        uint32_t nb) {
    __shared__ uint32_t s_w[8];
    const uint32_t P = (uint32_t)a.P;
    const uint32_t base = blockIdx.x * DP_CHUNK;
    uint32_t nA = 0, nB = 0, nC = 0;
#pragma unroll
    for (int j = 0; j < DP_ITEMS; j++) {
        const uint32_t i = base + j * DP_THREADS + threadIdx.x;     // coalesced
        if (i < P) {
            uint32_t code;
            if (a.prune_mask) {
                code = a.prune_mask[i] ? 0u : CODE_KEEP;
            } else {
                const float sc[3] = {a.scaling[3 * (size_t)i], a.scaling[3 * (size_t)i + 1], a.scaling[3 * (size_t)i + 2]};
                code = densify_classify_one(a.xyz_gradient_accum[i], a.denom[i], sc, a.opacity[i], a);
            }
            codes[i] = (uint8_t)code;
            nA += code & 1u;
            nB += (code >> 1) & 1u;
            nC += (code >> 2) & 1u;
        }
    }
    uint32_t tA, tB, tC;
    block_excl_scan_256(nA, s_w, &tA);
    block_excl_scan_256(nB, s_w, &tB);
    block_excl_scan_256(nC, s_w, &tC);
    if (threadIdx.x == 0) {
        block_sums[blockIdx.x] = tA;
        block_sums[nb + blockIdx.x] = tB;
        block_sums[2 * nb + blockIdx.x] = tC;
    }
}

// Pass 2 (one block): exclusive scan of the block sums of each segment in place; totals -> counts.
__global__ void __launch_bounds__(256)
densify_blocksums_kernel(uint32_t* __restrict__ block_sums, uint32_t nb, int32_t* __restrict__ counts) {
    __shared__ uint32_t s_w[8];
    uint32_t totals[3];
    for (int c = 0; c < 3; c++) {
        uint32_t* sums = block_sums + (size_t)c * nb;
        uint32_t carry = 0;
        for (uint32_t b0 = 0; b0 < nb; b0 += 256) {
            const uint32_t b = b0 + threadIdx.x;
            const uint32_t v = b < nb ? sums[b] : 0u;
            uint32_t total;
            const uint32_t ex = block_excl_scan_256(v, s_w, &total);
            if (b < nb) sums[b] = carry + ex;
            carry += total;
        }
        totals[c] = carry;
    }
    if (threadIdx.x == 0) {
        counts[0] = (int32_t)totals[0];
        counts[1] = (int32_t)totals[1];
        counts[2] = (int32_t)totals[2];
        const unsigned long long pn = (unsigned long long)totals[0] + totals[1] + 2ull * totals[2];
        counts[3] = pn > 0x7fffffffull ? -1 : (int32_t)pn;
    }
}

// Pass 3: rank of every produced row inside its segment (source order), DENSIFY_NONE where nothing is produced.
__global__ void __launch_bounds__(DP_THREADS)
densify_dst_kernel(const uint8_t* __restrict__ codes, uint32_t P, const uint32_t* __restrict__ block_sums, uint32_t nb,
                   uint32_t* __restrict__ dstA, uint32_t* __restrict__ dstB, uint32_t* __restrict__ dstC) {
    __shared__ uint32_t s_w[8];
    const uint32_t first = blockIdx.x * DP_CHUNK + threadIdx.x * DP_ITEMS;   // DP_ITEMS consecutive Gaussians
    uint32_t code[DP_ITEMS];
    uint32_t nA = 0, nB = 0, nC = 0;
#pragma unroll
    for (int j = 0; j < DP_ITEMS; j++) {
        code[j] = (first + j < P) ? codes[first + j] : 0u;
        nA += code[j] & 1u;
        nB += (code[j] >> 1) & 1u;
        nC += (code[j] >> 2) & 1u;
    }
    uint32_t rA = block_excl_scan_256(nA, s_w, nullptr) + block_sums[blockIdx.x];
    uint32_t rB = block_excl_scan_256(nB, s_w, nullptr) + block_sums[nb + blockIdx.x];
    uint32_t rC = block_excl_scan_256(nC, s_w, nullptr) + block_sums[2 * nb + blockIdx.x];
#pragma unroll
    for (int j = 0; j < DP_ITEMS; j++) {
        if (first + j < P) {
            dstA[first + j] = (code[j] & CODE_KEEP) ? rA++ : DENSIFY_NONE;
            dstB[first + j] = (code[j] & CODE_CLONE) ? rB++ : DENSIFY_NONE;
            dstC[first + j] = (code[j] & CODE_CHILDREN) ? rC++ : DENSIFY_NONE;
        }
    }
}

// Apply: grid.y = tensor; grid-stride over the tensor's source elements (coalesced reads; writes are
// coalesced within each run of surviving rows because every segment keeps source order).  The element -> (row,
// column) split is a division by the row width: compile-time for the widths of the Gaussian model (1, 3, 4, 45 =
// 15 x 3 higher-order SH floats) so that it is a multiply-shift, 32-bit whenever P * width < 2^32.
template <int WIDTH, typename index_t>
__device__ __forceinline__ void densify_move_rows(const GsrDensifyTensor& t, const GsrDensifyApplyArgs& a,
                                                  const uint32_t* __restrict__ dstA, const uint32_t* __restrict__ dstB,
                                                  const uint32_t* __restrict__ dstC) {
    const index_t w = WIDTH ? (index_t)WIDTH : (index_t)t.width;
    const index_t total = (index_t)a.P * w;
    const index_t stride = (index_t)gridDim.x * blockDim.x;
    for (index_t e = (index_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride) {
        const index_t i = e / w;
        densify_move_element(t, (long long)i, (int)(e - i * w), dstA, dstB, dstC, a);
    }
}

template <typename index_t>
__global__ void __launch_bounds__(256)
densify_move_kernel(const __grid_constant__ GsrDensifyApplyArgs a, const uint32_t* __restrict__ dstA,
                    const uint32_t* __restrict__ dstB, const uint32_t* __restrict__ dstC) {
    const GsrDensifyTensor& t = a.tensors[blockIdx.y];
    switch (t.width) {                       // uniform per block
        case 1:  densify_move_rows<1, index_t>(t, a, dstA, dstB, dstC); break;
        case 3:  densify_move_rows<3, index_t>(t, a, dstA, dstB, dstC); break;
        case 4:  densify_move_rows<4, index_t>(t, a, dstA, dstB, dstC); break;
        case 45: densify_move_rows<45, index_t>(t, a, dstA, dstB, dstC); break;
        default: densify_move_rows<0, index_t>(t, a, dstA, dstB, dstC); break;
    }
}

// n = P with a byte mask (visible) or with neither (then visible <=> radii[i] > 0, render()'s own definition,
// gaussian_renderer/__init__.py:152); n = number of indices with an index list (what render() actually returns:
// `(radii > 0).nonzero()`).
__global__ void __launch_bounds__(256)
densify_stats_kernel(uint32_t n, const float* __restrict__ grad, const uint8_t* __restrict__ visible,
                     const int64_t* __restrict__ indices, const int32_t* __restrict__ radii,
                     float* __restrict__ accum, float* __restrict__ denom, float* __restrict__ max_radii) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    size_t i = t;
    if (indices) i = (size_t)indices[t];
    else if (visible ? !visible[t] : radii[t] <= 0) return;
    const float g3[3] = {grad[3 * i], grad[3 * i + 1], 0.0f};
    densify_stats_one(g3, radii ? radii + i : nullptr, accum + i, denom + i, max_radii ? max_radii + i : nullptr);
}

__global__ void __launch_bounds__(256)
reset_opacity_kernel(uint32_t P, float* __restrict__ opacity, float* __restrict__ m, float* __restrict__ v) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    opacity[i] = reset_opacity_one(opacity[i]);
    if (m) m[i] = 0.0f;
    if (v) v[i] = 0.0f;
}

// ---------------------------------------------------------------- PLY rows
// One thread per row element: the row side is fully coalesced, the model side touches every float exactly once.
template <bool PACK, typename index_t>
__global__ void __launch_bounds__(256)
ply_rows_kernel(index_t total, int K, float* __restrict__ xyz, float* __restrict__ f_dc, float* __restrict__ f_rest,
                float* __restrict__ opacity, float* __restrict__ scaling, float* __restrict__ rotation,
                float* __restrict__ rows) {
    const index_t Wd = (index_t)ply_row_floats(K);
    const index_t stride = (index_t)gridDim.x * blockDim.x;
    for (index_t e = (index_t)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride) {
        const index_t i = e / Wd;
        float* cell = ply_cell((long long)i, (int)(e - i * Wd), K, xyz, f_dc, f_rest, opacity, scaling, rotation);
        if (PACK) rows[e] = cell ? *cell : 0.0f;
        else if (cell) *cell = rows[e];
    }
}

}  // namespace gsr

using namespace gsr;

extern "C" int64_t gsr_densify_plan_words(int32_t P) {
    if (P < 0) return 0;
    const uint32_t nb = plan_blocks((uint32_t)P);
    return 3 * (int64_t)P + 3 * (int64_t)nb + 4 + ((int64_t)P + 3) / 4 + 4;
}

extern "C" int gsr_densify_plan(const GsrDensifyPlanArgs* a, uint32_t* plan, int32_t* counts_host, void* stream) {
    if (!a || !plan || !counts_host || a->P <= 0) {
        set_error("gsr_densify_plan: bad arguments");
        return GSR_ERR_INVALID_ARG;
    }
    if (!a->prune_mask && (!a->xyz_gradient_accum || !a->denom || !a->scaling || !a->opacity)) {
        set_error("gsr_densify_plan: xyz_gradient_accum, denom, scaling and opacity are required");
        return GSR_ERR_INVALID_ARG;
    }
    cudaStream_t st = (cudaStream_t)stream;
    const uint32_t P = (uint32_t)a->P;
    PlanView v = plan_view(plan, P);
    uint8_t* codes = reinterpret_cast<uint8_t*>(v.codes_w);
    GSR_LAUNCH(densify_classify_kernel, v.nb, DP_THREADS, 0, st, *a, codes, v.block_sums, v.nb);
    GSR_LAUNCH(densify_blocksums_kernel, 1, 256, 0, st, v.block_sums, v.nb, v.counts);
    GSR_LAUNCH(densify_dst_kernel, v.nb, DP_THREADS, 0, st, codes, P, v.block_sums, v.nb, v.dstA, v.dstB, v.dstC);
    GSR_CUDA_TRY(cudaMemcpyAsync(counts_host, v.counts, 4 * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    return check_stage("densify plan", false, st);
}

extern "C" int gsr_densify_apply(const GsrDensifyApplyArgs* a, const uint32_t* plan, void* stream) {
    if (!a || !plan || a->P <= 0 || a->A < 0 || a->B < 0 || a->C < 0 || a->n_tensors < 0 ||
        a->n_tensors > GSR_DENSIFY_MAX_TENSORS) {
        set_error("gsr_densify_apply: bad arguments");
        return GSR_ERR_INVALID_ARG;
    }
    int max_width = 0;
    for (int t = 0; t < a->n_tensors; t++) {
        const GsrDensifyTensor& e = a->tensors[t];
        if (!e.src || e.width <= 0 || e.kind < GSR_DENSIFY_COPY || e.kind > GSR_DENSIFY_SCALING ||
            (!e.dst && (long long)a->A + a->B + 2LL * a->C > 0)) {
            set_error("gsr_densify_apply: tensor %d is malformed", t);
            return GSR_ERR_INVALID_ARG;
        }
        if (e.kind == GSR_DENSIFY_XYZ && a->C > 0 &&
            (e.width != 3 || !a->scaling || !a->rotation || !a->noise || a->noise_stride < a->P)) {
            set_error("gsr_densify_apply: split children need scaling, rotation and noise (stride >= P)");
            return GSR_ERR_INVALID_ARG;
        }
        max_width = std::max(max_width, e.width);
    }
    if (a->n_tensors == 0 || (long long)a->A + a->B + 2LL * a->C == 0) return GSR_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const uint32_t P = (uint32_t)a->P;
    const uint32_t* dstA = plan;
    const uint32_t* dstB = plan + (size_t)P;
    const uint32_t* dstC = plan + 2 * (size_t)P;
    const long long widest = (long long)P * max_width;
    const unsigned gx = (unsigned)std::min<long long>((widest + 255) / 256, 148LL * 8);
    if (widest + 256LL * gx < GSR_INDEX32_LIMIT)  // element index + one grid stride fits in 32 bits
        GSR_LAUNCH(densify_move_kernel<uint32_t>, dim3(gx, (unsigned)a->n_tensors), 256, 0, st, *a, dstA, dstB, dstC);
    else
        GSR_LAUNCH(densify_move_kernel<long long>, dim3(gx, (unsigned)a->n_tensors), 256, 0, st, *a, dstA, dstB, dstC);
    return check_stage("densify apply", false, st);
}

extern "C" int gsr_densify_stats(int32_t P, const float* viewspace_grad, const uint8_t* visible, const int64_t* indices,
                                 int64_t n_indices, const int32_t* radii, float* xyz_gradient_accum, float* denom,
                                 float* max_radii2D, void* stream) {
    if (P < 0 || n_indices < 0 || n_indices > 0x7fffffffLL ||
        (P > 0 && (!viewspace_grad || !xyz_gradient_accum || !denom)) || (visible && indices) ||
        (P > 0 && !visible && !indices && !radii)) {
        set_error("gsr_densify_stats: bad arguments (one of visible / indices / radii selects the Gaussians)");
        return GSR_ERR_INVALID_ARG;
    }
    const uint32_t n = indices ? (uint32_t)n_indices : (uint32_t)P;
    if (P == 0 || n == 0) return GSR_OK;
    cudaStream_t st = (cudaStream_t)stream;
    GSR_LAUNCH(densify_stats_kernel, div_up(n, 256u), 256, 0, st, n, viewspace_grad, visible, indices, radii,
               xyz_gradient_accum, denom, max_radii2D);
    return check_stage("densification stats", false, st);
}

extern "C" int gsr_reset_opacity(int32_t P, float* opacity, float* exp_avg, float* exp_avg_sq, void* stream) {
    if (P < 0 || (P > 0 && !opacity)) {
        set_error("gsr_reset_opacity: bad arguments");
        return GSR_ERR_INVALID_ARG;
    }
    if (P == 0) return GSR_OK;
    cudaStream_t st = (cudaStream_t)stream;
    GSR_LAUNCH(reset_opacity_kernel, div_up((uint32_t)P, 256u), 256, 0, st, (uint32_t)P, opacity, exp_avg, exp_avg_sq);
    return check_stage("reset opacity", false, st);
}

static int ply_rows(bool pack, int32_t P, int32_t K, float* xyz, float* f_dc, float* f_rest, float* opacity,
                    float* scaling, float* rotation, float* rows, void* stream) {
    if (P < 0 || K < 0 || (P > 0 && (!xyz || !f_dc || (K > 0 && !f_rest) || !opacity || !scaling || !rotation || !rows))) {
        set_error("gsr_ply_%s: bad arguments", pack ? "pack" : "unpack");
        return GSR_ERR_INVALID_ARG;
    }
    if (P == 0) return GSR_OK;
    cudaStream_t st = (cudaStream_t)stream;
    const long long total = (long long)P * ply_row_floats(K);
    const unsigned gx = (unsigned)std::min<long long>((total + 255) / 256, 148LL * 16);
#define PLY_LAUNCH(PACK, T) GSR_LAUNCH((ply_rows_kernel<PACK, T>), gx, 256, 0, st, (T)total, (int)K, xyz, f_dc, f_rest, \
                                       opacity, scaling, rotation, rows)
    const bool small = total + 256LL * gx < GSR_INDEX32_LIMIT;
    if (pack) { if (small) PLY_LAUNCH(true, uint32_t); else PLY_LAUNCH(true, long long); }
    else      { if (small) PLY_LAUNCH(false, uint32_t); else PLY_LAUNCH(false, long long); }
#undef PLY_LAUNCH
    return check_stage(pack ? "ply pack" : "ply unpack", false, st);
}

extern "C" int gsr_ply_pack(int32_t P, int32_t K_rest, const float* xyz, const float* f_dc, const float* f_rest,
                            const float* opacity, const float* scaling, const float* rotation, float* rows,
                            void* stream) {
    return ply_rows(true, P, K_rest, const_cast<float*>(xyz), const_cast<float*>(f_dc), const_cast<float*>(f_rest),
                    const_cast<float*>(opacity), const_cast<float*>(scaling), const_cast<float*>(rotation), rows, stream);
}

extern "C" int gsr_ply_unpack(int32_t P, int32_t K_rest, const float* rows, float* xyz, float* f_dc, float* f_rest,
                              float* opacity, float* scaling, float* rotation, void* stream) {
    return ply_rows(false, P, K_rest, xyz, f_dc, f_rest, opacity, scaling, rotation, const_cast<float*>(rows), stream);
}
