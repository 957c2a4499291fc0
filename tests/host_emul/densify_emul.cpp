// densify_emul.cpp — CPU emulation of the launch structure of gaussian_splatting_b200/csrc/densify.cu: the SAME
// per-element functions (densify_core.cuh), with the block scans replaced by serial prefix sums and the
// grid-stride loops by plain loops.  Test infrastructure for the GPU-less container (tests/test_densify_emul_cpu.py
// compares it with oracle/densify_ref.py); it is not part of the product and is never loaded by it.
#include <stdint.h>
#include "../../gaussian_splatting_b200/csrc/densify_core.cuh"

using namespace gsr;

extern "C" int emul_densify_plan(const GsrDensifyPlanArgs* a, uint32_t* dstA, uint32_t* dstB, uint32_t* dstC,
                                 int32_t* counts) {
    uint32_t rA = 0, rB = 0, rC = 0;
    for (int32_t i = 0; i < a->P; i++) {
        uint32_t code;
        if (a->prune_mask) code = a->prune_mask[i] ? 0u : CODE_KEEP;
        else code = densify_classify_one(a->xyz_gradient_accum[i], a->denom[i], a->scaling + 3 * (size_t)i, a->opacity[i], *a);
        dstA[i] = (code & CODE_KEEP) ? rA++ : DENSIFY_NONE;
        dstB[i] = (code & CODE_CLONE) ? rB++ : DENSIFY_NONE;
        dstC[i] = (code & CODE_CHILDREN) ? rC++ : DENSIFY_NONE;
    }
    counts[0] = (int32_t)rA; counts[1] = (int32_t)rB; counts[2] = (int32_t)rC;
    counts[3] = (int32_t)(rA + rB + 2 * rC);
    return 0;
}

extern "C" int emul_densify_apply(const GsrDensifyApplyArgs* a, const uint32_t* dstA, const uint32_t* dstB,
                                  const uint32_t* dstC) {
    for (int t = 0; t < a->n_tensors; t++) {
        const GsrDensifyTensor& e = a->tensors[t];
        for (long long i = 0; i < a->P; i++)
            for (int k = 0; k < e.width; k++) densify_move_element(e, i, k, dstA, dstB, dstC, *a);
    }
    return 0;
}

extern "C" int emul_densify_stats(int32_t P, const float* grad, const uint8_t* visible, const int64_t* indices,
                                  int64_t n_indices, const int32_t* radii, float* accum, float* denom, float* max_radii) {
    const int64_t n = indices ? n_indices : P;
    for (int64_t t = 0; t < n; t++) {
        size_t i = (size_t)t;
        if (indices) i = (size_t)indices[t];
        else if (visible ? !visible[t] : radii[t] <= 0) continue;
        const float g3[3] = {grad[3 * i], grad[3 * i + 1], 0.0f};
        densify_stats_one(g3, radii ? radii + i : nullptr, accum + i, denom + i, max_radii ? max_radii + i : nullptr);
    }
    return 0;
}

extern "C" int emul_reset_opacity(int32_t P, float* opacity, float* m, float* v) {
    for (int32_t i = 0; i < P; i++) {
        opacity[i] = reset_opacity_one(opacity[i]);
        if (m) m[i] = 0.0f;
        if (v) v[i] = 0.0f;
    }
    return 0;
}

extern "C" int emul_ply_rows(int pack, int32_t P, int32_t K, float* xyz, float* f_dc, float* f_rest, float* opacity,
                             float* scaling, float* rotation, float* rows) {
    const int Wd = ply_row_floats(K);
    for (long long e = 0; e < (long long)P * Wd; e++) {
        const long long i = e / Wd;
        float* cell = ply_cell(i, (int)(e - i * Wd), K, xyz, f_dc, f_rest, opacity, scaling, rotation);
        if (pack) rows[e] = cell ? *cell : 0.0f;
        else if (cell) *cell = rows[e];
    }
    return 0;
}
