/*
 * gs_oracle.c — CPU restatement of the reference's differentiable Gaussian rasterizer.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (gaussian_splatting_b200/,
 * libgsrast.so) may link, import or call this file; it is the checker used by tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
 *
 * The reference (mango1118/gaussian_splatting, submodules/diff-gaussian-rasterization,
 * "DGR") has no CPU implementation: every stage is a __global__/__device__ function.
 * Each function below restates one of them in plain C (fp32 arithmetic, no FMA
 * contraction: build with -ffp-contract=off) and cites the file:line it follows.
 * Paths are relative to DGR/cuda_rasterizer/ unless stated.
 *
 * Parity status: pinned.  tests/golden/raster_golden.npz holds outputs of the reference's own CUDA
 * rasterizer (oracle/_ref, built from /root/reference by oracle/build_ref.py) run on a B200 by
 * tests/golden/make_raster_golden_gpu.py; tests/test_raster_golden_cpu.py checks this oracle against it on
 * every CPU run, and tests/test_parity_gpu.py checks the CUDA path against both.  The reference ships no
 * golden vectors of its own for this path.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include "gso_parallel.h"

int gso_num_threads = 1;
void gso_set_threads(int n) { gso_num_threads = n < 1 ? 1 : (n > 256 ? 256 : n); }
int gso_get_threads(void) { return gso_num_threads; }

#define BLOCK_X 16 /* config.h:16-17 */
#define BLOCK_Y 16
#define BLOCK_SIZE (BLOCK_X * BLOCK_Y)

/* ---- SH constants, auxiliary.h:23-40 ---- */
static const float SH_C0 = 0.28209479177387814f;
static const float SH_C1 = 0.4886025119029199f;
static const float SH_C2[] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                              -1.0925484305920792f, 0.5462742152960396f};
static const float SH_C3[] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                              0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                              -0.5900435899266435f};

/* glm::mat3 semantics: column-major, m[c][r]; product sums k = 0,1,2 left to right
 * (third_party/glm/glm/detail/type_mat3x3.inl:486-519). */
typedef struct { float m[3][3]; } mat3;

static mat3 m3(float a, float b, float c, float d, float e, float f, float g, float h, float i) {
    mat3 r = {{{a, b, c}, {d, e, f}, {g, h, i}}};
    return r;
}
static mat3 m3mul(mat3 A, mat3 B) {
    mat3 R;
    for (int c = 0; c < 3; c++)
        for (int r = 0; r < 3; r++)
            R.m[c][r] = A.m[0][r] * B.m[c][0] + A.m[1][r] * B.m[c][1] + A.m[2][r] * B.m[c][2];
    return R;
}
static mat3 m3t(mat3 A) {
    mat3 R;
    for (int c = 0; c < 3; c++)
        for (int r = 0; r < 3; r++) R.m[c][r] = A.m[r][c];
    return R;
}

/* auxiliary.h:43-46 (double arithmetic: 1.0 and 0.5 are double literals) */
static float ndc2Pix(float v, int S) { return (float)(((v + 1.0) * S - 1.0) * 0.5); }

static int imin(int a, int b) { return a < b ? a : b; }
static int imax(int a, int b) { return a > b ? a : b; }

/* auxiliary.h:49-59 */
static void getRect(float px, float py, int max_radius, int gx, int gy, uint32_t* rmin, uint32_t* rmax) {
    rmin[0] = (uint32_t)imin(gx, imax(0, (int)((px - max_radius) / BLOCK_X)));
    rmin[1] = (uint32_t)imin(gy, imax(0, (int)((py - max_radius) / BLOCK_Y)));
    rmax[0] = (uint32_t)imin(gx, imax(0, (int)((px + max_radius + BLOCK_X - 1) / BLOCK_X)));
    rmax[1] = (uint32_t)imin(gy, imax(0, (int)((py + max_radius + BLOCK_Y - 1) / BLOCK_Y)));
}

/* auxiliary.h:75-95 */
static void transformPoint4x3(const float* p, const float* mt, float* o) {
    o[0] = mt[0] * p[0] + mt[4] * p[1] + mt[8] * p[2] + mt[12];
    o[1] = mt[1] * p[0] + mt[5] * p[1] + mt[9] * p[2] + mt[13];
    o[2] = mt[2] * p[0] + mt[6] * p[1] + mt[10] * p[2] + mt[14];
}
static void transformPoint4x4(const float* p, const float* mt, float* o) {
    o[0] = mt[0] * p[0] + mt[4] * p[1] + mt[8] * p[2] + mt[12];
    o[1] = mt[1] * p[0] + mt[5] * p[1] + mt[9] * p[2] + mt[13];
    o[2] = mt[2] * p[0] + mt[6] * p[1] + mt[10] * p[2] + mt[14];
    o[3] = mt[3] * p[0] + mt[7] * p[1] + mt[11] * p[2] + mt[15];
}
/* auxiliary.h:109-117 */
static void transformVec4x3Transpose(const float* p, const float* mt, float* o) {
    o[0] = mt[0] * p[0] + mt[1] * p[1] + mt[2] * p[2];
    o[1] = mt[4] * p[0] + mt[5] * p[1] + mt[6] * p[2];
    o[2] = mt[8] * p[0] + mt[9] * p[1] + mt[10] * p[2];
}

/* forward.cu:149-190 */
static void computeCov3D(const float* scale, float mod, const float* rot, float* cov3D) {
    mat3 S = m3(1, 0, 0, 0, 1, 0, 0, 0, 1);
    S.m[0][0] = mod * scale[0];
    S.m[1][1] = mod * scale[1];
    S.m[2][2] = mod * scale[2];
    float r = rot[0], x = rot[1], y = rot[2], z = rot[3];
    mat3 R = m3(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
    mat3 M = m3mul(S, R);
    mat3 Sigma = m3mul(m3t(M), M);
    cov3D[0] = Sigma.m[0][0];
    cov3D[1] = Sigma.m[0][1];
    cov3D[2] = Sigma.m[0][2];
    cov3D[3] = Sigma.m[1][1];
    cov3D[4] = Sigma.m[1][2];
    cov3D[5] = Sigma.m[2][2];
}

typedef struct {
    float t[3], txtz, tytz, limx, limy;
    mat3 T;
} cov2d_ctx;

/* forward.cu:89-141 */
static void computeCov2D(const float* mean, float focal_x, float focal_y, float tan_fovx, float tan_fovy,
                         const float* cov3D, const float* viewmatrix, float* cov, cov2d_ctx* ctx) {
    float t[3];
    transformPoint4x3(mean, viewmatrix, t);
    const float limx = 1.3f * tan_fovx;
    const float limy = 1.3f * tan_fovy;
    const float txtz = t[0] / t[2];
    const float tytz = t[1] / t[2];
    t[0] = fminf(limx, fmaxf(-limx, txtz)) * t[2];
    t[1] = fminf(limy, fmaxf(-limy, tytz)) * t[2];
    mat3 J = m3(focal_x / t[2], 0.0f, -(focal_x * t[0]) / (t[2] * t[2]),
                0.0f, focal_y / t[2], -(focal_y * t[1]) / (t[2] * t[2]),
                0, 0, 0);
    mat3 W = m3(viewmatrix[0], viewmatrix[4], viewmatrix[8],
                viewmatrix[1], viewmatrix[5], viewmatrix[9],
                viewmatrix[2], viewmatrix[6], viewmatrix[10]);
    mat3 T = m3mul(W, J);
    mat3 Vrk = m3(cov3D[0], cov3D[1], cov3D[2], cov3D[1], cov3D[3], cov3D[4], cov3D[2], cov3D[4], cov3D[5]);
    mat3 c = m3mul(m3mul(m3t(T), m3t(Vrk)), T);
    cov[0] = c.m[0][0];
    cov[1] = c.m[0][1];
    cov[2] = c.m[1][1];
    if (ctx) {
        ctx->t[0] = t[0]; ctx->t[1] = t[1]; ctx->t[2] = t[2];
        ctx->txtz = txtz; ctx->tytz = tytz; ctx->limx = limx; ctx->limy = limy; ctx->T = T;
    }
}

/* forward.cu:22-80. sh = (M,3) coefficients of one Gaussian. */
static void computeColorFromSH(int deg, const float* pos, const float* campos, const float* sh,
                               float* rgb, uint8_t* clamped) {
    float dir[3] = {pos[0] - campos[0], pos[1] - campos[1], pos[2] - campos[2]};
    float len = sqrtf(dir[0] * dir[0] + dir[1] * dir[1] + dir[2] * dir[2]);
    float x = dir[0] / len, y = dir[1] / len, z = dir[2] / len;
    for (int c = 0; c < 3; c++) {
#define SH(k) sh[(k) * 3 + c]
        float result = SH_C0 * SH(0);
        if (deg > 0) {
            result = result - SH_C1 * y * SH(1) + SH_C1 * z * SH(2) - SH_C1 * x * SH(3);
            if (deg > 1) {
                float xx = x * x, yy = y * y, zz = z * z;
                float xy = x * y, yz = y * z, xz = x * z;
                result = result + SH_C2[0] * xy * SH(4) + SH_C2[1] * yz * SH(5) +
                         SH_C2[2] * (2.0f * zz - xx - yy) * SH(6) + SH_C2[3] * xz * SH(7) +
                         SH_C2[4] * (xx - yy) * SH(8);
                if (deg > 2) {
                    result = result + SH_C3[0] * y * (3.0f * xx - yy) * SH(9) + SH_C3[1] * xy * z * SH(10) +
                             SH_C3[2] * y * (4.0f * zz - xx - yy) * SH(11) +
                             SH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * SH(12) +
                             SH_C3[4] * x * (4.0f * zz - xx - yy) * SH(13) + SH_C3[5] * z * (xx - yy) * SH(14) +
                             SH_C3[6] * x * (xx - 3.0f * yy) * SH(15);
                }
            }
        }
#undef SH
        result += 0.5f;
        clamped[c] = (result < 0);
        rgb[c] = fmaxf(result, 0.0f);
    }
}

/* rasterizer_impl.cu:36-51 */
uint32_t gso_higher_msb(uint32_t n) {
    uint32_t msb = sizeof(n) * 4;
    uint32_t step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step; else msb -= step;
    }
    if (n >> msb) msb++;
    return msb;
}

/* checkFrustum / in_frustum, rasterizer_impl.cu:56-73, auxiliary.h:164-190 */
void gso_mark_visible(int P, const float* means3D, const float* viewmatrix, uint8_t* present) {
    for (int i = 0; i < P; i++) {
        float pv[3];
        transformPoint4x3(means3D + 3 * i, viewmatrix, pv);
        present[i] = pv[2] > 0.2f;
    }
}

/*
 * preprocessCUDA forward, forward.cu:222-351 (order of operations and early-outs).
 * NULL shs => colors_precomp mode (rgb/clamped untouched, as the reference); NULL scales
 * => cov3D_precomp mode.  All outputs are P-sized arrays; entries of culled Gaussians are
 * left as passed in except radii and tiles_touched, which are zeroed first (:254-255).
 */
typedef struct {
    int P, D, M;
    const float *means3D, *scales; float scale_modifier;
    const float *rotations, *opacities, *shs, *cov3D_precomp, *colors_precomp, *viewmatrix, *projmatrix, *campos;
    int W, H; float tan_fovx, tan_fovy; int antialiasing;
    int32_t* radii; float *means2D, *depths, *cov3Ds, *rgb; uint8_t* clamped; float* conic_opacity;
    uint32_t* tiles_touched;
} pre_ctx;

static void pre_range(int lo, int hi, int tid, void* vp) {
    (void)tid;
    const pre_ctx* a = (const pre_ctx*)vp;
    const int D = a->D, M = a->M, W = a->W, H = a->H, antialiasing = a->antialiasing;
    const float *means3D = a->means3D, *scales = a->scales, *rotations = a->rotations, *opacities = a->opacities,
                *shs = a->shs, *cov3D_precomp = a->cov3D_precomp, *colors_precomp = a->colors_precomp,
                *viewmatrix = a->viewmatrix, *projmatrix = a->projmatrix, *campos = a->campos;
    const float scale_modifier = a->scale_modifier, tan_fovx = a->tan_fovx, tan_fovy = a->tan_fovy;
    int32_t* radii = a->radii; float *means2D = a->means2D, *depths = a->depths, *cov3Ds = a->cov3Ds, *rgb = a->rgb;
    uint8_t* clamped = a->clamped; float* conic_opacity = a->conic_opacity; uint32_t* tiles_touched = a->tiles_touched;
    const float focal_y = H / (2.0f * tan_fovy); /* rasterizer_impl.cu:253-254 */
    const float focal_x = W / (2.0f * tan_fovx);
    const int gx = (W + BLOCK_X - 1) / BLOCK_X, gy = (H + BLOCK_Y - 1) / BLOCK_Y;
    for (int idx = lo; idx < hi; idx++) {
        radii[idx] = 0;
        tiles_touched[idx] = 0;
        const float* p_orig = means3D + 3 * idx;
        float p_view[3];
        transformPoint4x3(p_orig, viewmatrix, p_view);
        if (p_view[2] <= 0.2f) continue;
        float p_hom[4];
        transformPoint4x4(p_orig, projmatrix, p_hom);
        float p_w = 1.0f / (p_hom[3] + 0.0000001f);
        float p_proj[3] = {p_hom[0] * p_w, p_hom[1] * p_w, p_hom[2] * p_w};
        const float* cov3D;
        if (cov3D_precomp) {
            cov3D = cov3D_precomp + idx * 6;
        } else {
            computeCov3D(scales + 3 * idx, scale_modifier, rotations + 4 * idx, cov3Ds + idx * 6);
            cov3D = cov3Ds + idx * 6;
        }
        float cov[3];
        computeCov2D(p_orig, focal_x, focal_y, tan_fovx, tan_fovy, cov3D, viewmatrix, cov, NULL);
        const float h_var = 0.3f;
        const float det_cov = cov[0] * cov[2] - cov[1] * cov[1];
        cov[0] += h_var;
        cov[2] += h_var;
        const float det_cov_plus_h_cov = cov[0] * cov[2] - cov[1] * cov[1];
        float h_convolution_scaling = 1.0f;
        if (antialiasing) h_convolution_scaling = sqrtf(fmaxf(0.000025f, det_cov / det_cov_plus_h_cov));
        const float det = det_cov_plus_h_cov;
        if (det == 0.0f) continue;
        float det_inv = 1.f / det;
        float conic[3] = {cov[2] * det_inv, -cov[1] * det_inv, cov[0] * det_inv};
        float mid = 0.5f * (cov[0] + cov[2]);
        float lambda1 = mid + sqrtf(fmaxf(0.1f, mid * mid - det));
        float lambda2 = mid - sqrtf(fmaxf(0.1f, mid * mid - det));
        float my_radius = ceilf(3.f * sqrtf(fmaxf(lambda1, lambda2)));
        float px = ndc2Pix(p_proj[0], W), py = ndc2Pix(p_proj[1], H);
        uint32_t rmin[2], rmax[2];
        getRect(px, py, (int)my_radius, gx, gy, rmin, rmax);
        if ((rmax[0] - rmin[0]) * (rmax[1] - rmin[1]) == 0) continue;
        if (!colors_precomp) computeColorFromSH(D, p_orig, campos, shs + (size_t)idx * M * 3, rgb + 3 * idx, clamped + 3 * idx);
        depths[idx] = p_view[2];
        radii[idx] = (int32_t)my_radius;
        means2D[2 * idx] = px;
        means2D[2 * idx + 1] = py;
        conic_opacity[4 * idx + 0] = conic[0];
        conic_opacity[4 * idx + 1] = conic[1];
        conic_opacity[4 * idx + 2] = conic[2];
        conic_opacity[4 * idx + 3] = opacities[idx] * h_convolution_scaling;
        tiles_touched[idx] = (rmax[1] - rmin[1]) * (rmax[0] - rmin[0]);
    }
}

void gso_preprocess(int P, int D, int M, const float* means3D, const float* scales, float scale_modifier,
                    const float* rotations, const float* opacities, const float* shs,
                    const float* cov3D_precomp, const float* colors_precomp, const float* viewmatrix,
                    const float* projmatrix, const float* campos, int W, int H, float tan_fovx,
                    float tan_fovy, int antialiasing,
                    int32_t* radii, float* means2D, float* depths, float* cov3Ds, float* rgb,
                    uint8_t* clamped, float* conic_opacity, uint32_t* tiles_touched) {
    pre_ctx a = {P, D, M, means3D, scales, scale_modifier, rotations, opacities, shs, cov3D_precomp, colors_precomp,
                 viewmatrix, projmatrix, campos, W, H, tan_fovx, tan_fovy, antialiasing, radii, means2D, depths,
                 cov3Ds, rgb, clamped, conic_opacity, tiles_touched};
    gso_parallel_for(0, P, 4096, pre_range, &a);
}

/*
 * Binning: InclusiveSum + duplicateWithKeys + SortPairs + identifyTileRanges
 * (rasterizer_impl.cu:78-164, 309-350).  keys/values must hold R = sum(tiles_touched)
 * entries; tmp_keys/tmp_values are scratch of the same size; ranges is T*2 uint32.
 * The sort is a stable LSD radix sort restricted to bits [0, 32+getHigherMsb(T)) exactly
 * like the reference's cub call (:335-340) — bits above that are ignored, as in CUB.
 * Returns R.
 */
int64_t gso_binning(int P, int W, int H, const int32_t* radii, const float* means2D, const float* depths,
                    const uint32_t* tiles_touched, uint64_t* keys, uint32_t* values, uint64_t* tmp_keys,
                    uint32_t* tmp_values, uint32_t* ranges, int64_t capacity) {
    const int gx = (W + BLOCK_X - 1) / BLOCK_X, gy = (H + BLOCK_Y - 1) / BLOCK_Y;
    int64_t R = 0;
    for (int i = 0; i < P; i++) R += tiles_touched[i];
    memset(ranges, 0, (size_t)gx * gy * 2 * sizeof(uint32_t)); /* :342 */
    if (R > capacity || R == 0) return R;
    int64_t off = 0;
    for (int idx = 0; idx < P; idx++) { /* duplicateWithKeys :78-126 */
        if (radii[idx] > 0) {
            uint32_t rmin[2], rmax[2];
            getRect(means2D[2 * idx], means2D[2 * idx + 1], radii[idx], gx, gy, rmin, rmax);
            uint32_t dbits;
            memcpy(&dbits, depths + idx, 4);
            for (uint32_t y = rmin[1]; y < rmax[1]; y++)
                for (uint32_t x = rmin[0]; x < rmax[0]; x++) {
                    uint64_t key = (uint64_t)(y * (uint32_t)gx + x);
                    key <<= 32;
                    key |= dbits;
                    keys[off] = key;
                    values[off] = (uint32_t)idx;
                    off++;
                }
        }
    }
    const int end_bit = 32 + (int)gso_higher_msb((uint32_t)(gx * gy));
    uint64_t *ka = keys, *kb = tmp_keys;
    uint32_t *va = values, *vb = tmp_values;
    for (int shift = 0; shift < end_bit; shift += 8) {
        int nbits = end_bit - shift < 8 ? end_bit - shift : 8;
        uint32_t mask = (1u << nbits) - 1u;
        size_t cnt[257];
        memset(cnt, 0, sizeof(cnt));
        for (int64_t i = 0; i < R; i++) cnt[((ka[i] >> shift) & mask) + 1]++;
        for (int d = 0; d < 256; d++) cnt[d + 1] += cnt[d];
        for (int64_t i = 0; i < R; i++) {
            size_t pos = cnt[(ka[i] >> shift) & mask]++;
            kb[pos] = ka[i];
            vb[pos] = va[i];
        }
        uint64_t* tk = ka; ka = kb; kb = tk;
        uint32_t* tv = va; va = vb; vb = tv;
    }
    if (ka != keys) {
        memcpy(keys, ka, (size_t)R * 8);
        memcpy(values, va, (size_t)R * 4);
    }
    for (int64_t idx = 0; idx < R; idx++) { /* identifyTileRanges :132-164 */
        uint32_t currtile = (uint32_t)(keys[idx] >> 32);
        if (idx == 0)
            ranges[2 * currtile] = 0;
        else {
            uint32_t prevtile = (uint32_t)(keys[idx - 1] >> 32);
            if (currtile != prevtile) {
                ranges[2 * prevtile + 1] = (uint32_t)idx;
                ranges[2 * currtile] = (uint32_t)idx;
            }
        }
        if (idx == R - 1) ranges[2 * currtile + 1] = (uint32_t)R;
    }
    return R;
}
