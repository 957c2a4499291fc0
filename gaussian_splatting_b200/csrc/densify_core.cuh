// densify_core.cuh — per-element arithmetic of the density-control kernels (densify.cu), kept free of
// CUDA-only constructs so that tests/host_emul/densify_emul.cpp can run the SAME functions in plain loops
// on the CPU (kernel-logic check in the GPU-less authoring container; the parity tests proper are the
// -m gpu tests against oracle/densify_ref.py).
//
// Reference: scene/gaussian_model.py:510-621 (densify_and_split / _clone / _prune), :323-327
// (reset_opacity), :630-633 (add_densification_stats), utils/general_utils.py:78-100 (build_rotation).
#pragma once
#include <math.h>
#include <stdint.h>
#include "../../include/gsrast.h"

#ifdef __CUDACC__
#define GSD_HD __host__ __device__ __forceinline__
#else
#define GSD_HD inline
#endif

namespace gsr {

constexpr uint32_t DENSIFY_NONE = 0xFFFFFFFFu;
constexpr uint32_t CODE_KEEP = 1u, CODE_CLONE = 2u, CODE_CHILDREN = 4u;
// x / (0.8 * N) with N = 2: on CUDA tensors torch divides by a Python scalar by multiplying with its fp32
// reciprocal (ATen div_true_kernel_cuda), and 1.0f / 1.6f rounds to exactly 0.625f.
constexpr float SPLIT_SCALE = 0.625f;

GSD_HD float sigmoid_f(float x) { return 1.0f / (1.0f + expf(-x)); }

// Decision for source Gaussian i: which of {original, clone, children} exist in the result.
// Follows densify_and_prune (:589-611) with the intermediate models folded:
//   * clone  (:554-556): |grad| >= threshold and max(exp(scaling)) <= percent_dense * extent
//   * split  (:513-516): grad >= threshold and max(exp(scaling)) > percent_dense * extent; the original is
//     removed (:549-550); clones carry zero padded gradient (:512-513) and are small, so they never split
//   * final prune (:598-611) on every staged row: sigmoid(opacity) < min_opacity, or (if max_screen_size)
//     max_radii2D (zeroed at :508, hence 0) > max_screen_size or max(exp(scaling)) > 0.1 * extent.
//     A clone shares its source's opacity and scale; children share the opacity and have scale * 0.625.
GSD_HD uint32_t densify_classify_one(float accum, float denom, const float* scaling3, float opacity_raw,
                                     const GsrDensifyPlanArgs& a) {
    float g = accum / denom;
    if (g != g) g = 0.0f;                                   // grads[grads.isnan()] = 0.0 (:591)
    const float s0 = expf(scaling3[0]), s1 = expf(scaling3[1]), s2 = expf(scaling3[2]);
    const float maxs = fmaxf(s0, fmaxf(s1, s2));
    const bool sel_clone = fabsf(g) >= a.grad_threshold && maxs <= a.dense_extent;
    const bool sel_split = g >= a.grad_threshold && maxs > a.dense_extent;
    const bool low_opacity = sigmoid_f(opacity_raw) < a.min_opacity;
    const bool screen = a.use_screen_size != 0;
    const bool big_vs = 0.0f > a.max_screen_size;
    const bool pruned = low_opacity || (screen && (big_vs || maxs > a.world_extent_limit));
    uint32_t code = 0;
    if (!sel_split && !pruned) code |= CODE_KEEP;
    if (sel_clone && !pruned) code |= CODE_CLONE;
    if (sel_split) {
        // get_scaling of a child = exp(log(exp(scaling) * 0.625)) (:524, then :602 on the new model)
        const float c0 = expf(logf(s0 * SPLIT_SCALE)), c1 = expf(logf(s1 * SPLIT_SCALE)), c2 = expf(logf(s2 * SPLIT_SCALE));
        const float maxc = fmaxf(c0, fmaxf(c1, c2));
        const bool pruned_child = low_opacity || (screen && (big_vs || maxc > a.world_extent_limit));
        if (!pruned_child) code |= CODE_CHILDREN;
    }
    return code;
}

// Component k of a child's position: (R(q) * (exp(scaling) * z))[k] + xyz[k]   (:517-523).
GSD_HD float densify_child_xyz(int k, float xyz_k, const float* scaling3, const float* rot4, const float* z3) {
    const float qr = rot4[0], qx = rot4[1], qy = rot4[2], qz = rot4[3];
    const float norm = sqrtf(qr * qr + qx * qx + qy * qy + qz * qz);      // build_rotation normalises (:79-81)
    const float r = qr / norm, x = qx / norm, y = qy / norm, z = qz / norm;
    float R0, R1, R2;
    if (k == 0)      { R0 = 1.f - 2.f * (y * y + z * z); R1 = 2.f * (x * y - r * z);       R2 = 2.f * (x * z + r * y); }
    else if (k == 1) { R0 = 2.f * (x * y + r * z);       R1 = 1.f - 2.f * (x * x + z * z); R2 = 2.f * (y * z - r * x); }
    else             { R0 = 2.f * (x * z - r * y);       R1 = 2.f * (y * z + r * x);       R2 = 1.f - 2.f * (x * x + y * y); }
    const float v0 = expf(scaling3[0]) * z3[0], v1 = expf(scaling3[1]) * z3[1], v2 = expf(scaling3[2]) * z3[2];
    return (R0 * v0 + R1 * v1 + R2 * v2) + xyz_k;
}

GSD_HD float densify_child_scaling(float scaling_raw) { return logf(expf(scaling_raw) * SPLIT_SCALE); }

// Moves column k of source row i of one tensor to every destination row derived from source row i.
// dstA/dstB/dstC: rank of the row inside its segment or DENSIFY_NONE.
GSD_HD void densify_move_element(const GsrDensifyTensor& t, long long i, int k, const uint32_t* dstA, const uint32_t* dstB,
                                 const uint32_t* dstC, const GsrDensifyApplyArgs& a) {
    const float v = t.src[i * t.width + k];
    const uint32_t da = dstA[i], db = dstB[i], dc = dstC[i];
    const bool moment = t.kind == GSR_DENSIFY_MOMENT;
    if (da != DENSIFY_NONE) t.dst[(long long)da * t.width + k] = v;
    if (db != DENSIFY_NONE) t.dst[((long long)a.A + db) * t.width + k] = moment ? 0.0f : v;
    if (dc != DENSIFY_NONE) {
        for (int copy = 0; copy < 2; copy++) {
            float w = v;
            if (moment) w = 0.0f;
            else if (t.kind == GSR_DENSIFY_SCALING) w = densify_child_scaling(v);
            else if (t.kind == GSR_DENSIFY_XYZ)
                w = densify_child_xyz(k, v, a.scaling + 3 * i, a.rotation + 4 * i,
                                      a.noise + ((long long)copy * a.noise_stride + i) * 3);
            t.dst[((long long)a.A + a.B + (long long)copy * a.C + dc) * t.width + k] = w;
        }
    }
}

// add_densification_stats (:630-633) + train.py:212-213 for one visible Gaussian.
GSD_HD void densify_stats_one(const float* grad3, const int32_t* radius, float* accum, float* denom, float* max_radius) {
    if (radius && max_radius) *max_radius = fmaxf(*max_radius, (float)*radius);
    *accum += sqrtf(grad3[0] * grad3[0] + grad3[1] * grad3[1]);
    *denom += 1.0f;
}

// reset_opacity (:323-327): inverse_sigmoid(min(sigmoid(o), 0.01)), inverse_sigmoid(x) = log(x / (1 - x)).
GSD_HD float reset_opacity_one(float opacity_raw) {
    const float x = fminf(sigmoid_f(opacity_raw), 0.01f);
    return logf(x / (1.0f - x));
}

// PLY row layout (gaussian_model.py:285-321): x y z nx ny nz f_dc_0..2 f_rest_0..3K-1 opacity scale_0..2 rot_0..3.
GSD_HD int ply_row_floats(int K) { return 17 + 3 * K; }

// The model element behind column c of row i; nullptr for the normals (zeros, :305).
GSD_HD float* ply_cell(long long i, int c, int K, float* xyz, float* f_dc, float* f_rest, float* opacity,
                       float* scaling, float* rotation) {
    if (c < 3) return xyz + 3 * i + c;
    if (c < 6) return nullptr;
    if (c < 9) return f_dc + 3 * i + (c - 6);
    if (c < 9 + 3 * K) {
        const int r = c - 9, ch = r / K, k = r - ch * K;     // f_rest_{ch*K + k} = _features_rest[i, k, ch] (:309)
        return f_rest + (i * K + k) * 3 + ch;
    }
    if (c == 9 + 3 * K) return opacity + i;
    if (c < 13 + 3 * K) return scaling + 3 * i + (c - 10 - 3 * K);
    return rotation + 4 * i + (c - 13 - 3 * K);
}

}  // namespace gsr
