// gs_math.cuh — per-Gaussian projection math shared by preprocess forward and backward.
//
// The scalar expression trees of the chain  cov3D -> cov2D -> eigen radius  and
// p_proj -> pixel -> tile rect  follow the reference operation by operation
// (DGR/cuda_rasterizer/forward.cu:89-190,251-350; auxiliary.h:43-59) so that nvcc's
// default FMA contraction produces the same bits: radii / tile rects / num_rendered
// are compared bit-exactly.  GLM is not used; Mat3 below reproduces
// GLM's column-major storage and the summation order of its operator*
// (third_party/glm/glm/detail/type_mat3x3.inl:486-519).
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#define GSR_HD __host__ __device__ __forceinline__

namespace gsr {

// Column-major 3x3, m[c][r] like glm::mat3.
struct Mat3 {
    float m[3][3];
};

GSR_HD Mat3 mat3_cols(float a, float b, float c, float d, float e, float f, float g, float h, float i) {
    Mat3 r;
    r.m[0][0] = a; r.m[0][1] = b; r.m[0][2] = c;
    r.m[1][0] = d; r.m[1][1] = e; r.m[1][2] = f;
    r.m[2][0] = g; r.m[2][1] = h; r.m[2][2] = i;
    return r;
}

GSR_HD Mat3 mat3_mul(const Mat3& A, const Mat3& B) {
    Mat3 R;
#pragma unroll
    for (int c = 0; c < 3; c++) {
#pragma unroll
        for (int r = 0; r < 3; r++) {
            R.m[c][r] = A.m[0][r] * B.m[c][0] + A.m[1][r] * B.m[c][1] + A.m[2][r] * B.m[c][2];
        }
    }
    return R;
}

GSR_HD Mat3 mat3_transpose(const Mat3& A) {
    Mat3 R;
#pragma unroll
    for (int c = 0; c < 3; c++)
#pragma unroll
        for (int r = 0; r < 3; r++) R.m[c][r] = A.m[r][c];
    return R;
}

// auxiliary.h:75-95
GSR_HD float3 transform_point_4x3(const float3& p, const float* mt) {
    float3 t = {
        mt[0] * p.x + mt[4] * p.y + mt[8] * p.z + mt[12],
        mt[1] * p.x + mt[5] * p.y + mt[9] * p.z + mt[13],
        mt[2] * p.x + mt[6] * p.y + mt[10] * p.z + mt[14],
    };
    return t;
}

GSR_HD float4 transform_point_4x4(const float3& p, const float* mt) {
    float4 t = {
        mt[0] * p.x + mt[4] * p.y + mt[8] * p.z + mt[12],
        mt[1] * p.x + mt[5] * p.y + mt[9] * p.z + mt[13],
        mt[2] * p.x + mt[6] * p.y + mt[10] * p.z + mt[14],
        mt[3] * p.x + mt[7] * p.y + mt[11] * p.z + mt[15]};
    return t;
}

// auxiliary.h:109-117
GSR_HD float3 transform_vec_4x3_transpose(const float3& p, const float* mt) {
    float3 t = {
        mt[0] * p.x + mt[1] * p.y + mt[2] * p.z,
        mt[4] * p.x + mt[5] * p.y + mt[6] * p.z,
        mt[8] * p.x + mt[9] * p.y + mt[10] * p.z,
    };
    return t;
}

// auxiliary.h:43-46 — the literals are double, so the whole expression is double.
GSR_HD float ndc2pix(float v, int S) { return ((v + 1.0) * S - 1.0) * 0.5; }

// Rotation of the (unnormalised) quaternion (r,x,y,z), GLM column-major as forward.cu:166-170.
GSR_HD Mat3 quat_to_R(float r, float x, float y, float z) {
    return mat3_cols(
        1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
        2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
        2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
}

// forward.cu:149-190. cov3D = upper triangle of Sigma = (S R)^T (S R).
GSR_HD void compute_cov3D(float3 scale, float mod, float4 rot, float* cov3D) {
    Mat3 S = mat3_cols(1.f, 0.f, 0.f, 0.f, 1.f, 0.f, 0.f, 0.f, 1.f);
    S.m[0][0] = mod * scale.x;
    S.m[1][1] = mod * scale.y;
    S.m[2][2] = mod * scale.z;
    Mat3 R = quat_to_R(rot.x, rot.y, rot.z, rot.w);
    Mat3 M = mat3_mul(S, R);
    Mat3 Sigma = mat3_mul(mat3_transpose(M), M);
    cov3D[0] = Sigma.m[0][0];
    cov3D[1] = Sigma.m[0][1];
    cov3D[2] = Sigma.m[0][2];
    cov3D[3] = Sigma.m[1][1];
    cov3D[4] = Sigma.m[1][2];
    cov3D[5] = Sigma.m[2][2];
}

struct Cov2DCtx {   // intermediates the backward pass needs again
    float3 t;       // view-space mean after the 1.3*tanfov clamp
    float txtz, tytz, limx, limy;
    Mat3 T;         // W * J (GLM sense)
};

// forward.cu:89-141.
GSR_HD float3 compute_cov2D(const float3& mean, float focal_x, float focal_y, float tan_fovx,
                            float tan_fovy, const float* cov3D, const float* view, Cov2DCtx* ctx) {
    float3 t = transform_point_4x3(mean, view);
    const float limx = 1.3f * tan_fovx;
    const float limy = 1.3f * tan_fovy;
    const float txtz = t.x / t.z;
    const float tytz = t.y / t.z;
    t.x = fminf(limx, fmaxf(-limx, txtz)) * t.z;
    t.y = fminf(limy, fmaxf(-limy, tytz)) * t.z;

    Mat3 J = mat3_cols(
        focal_x / t.z, 0.0f, -(focal_x * t.x) / (t.z * t.z),
        0.0f, focal_y / t.z, -(focal_y * t.y) / (t.z * t.z),
        0, 0, 0);
    Mat3 W = mat3_cols(
        view[0], view[4], view[8],
        view[1], view[5], view[9],
        view[2], view[6], view[10]);
    Mat3 T = mat3_mul(W, J);
    Mat3 Vrk = mat3_cols(
        cov3D[0], cov3D[1], cov3D[2],
        cov3D[1], cov3D[3], cov3D[4],
        cov3D[2], cov3D[4], cov3D[5]);
    Mat3 cov = mat3_mul(mat3_mul(mat3_transpose(T), mat3_transpose(Vrk)), T);
    if (ctx) {
        ctx->t = t; ctx->txtz = txtz; ctx->tytz = tytz; ctx->limx = limx; ctx->limy = limy; ctx->T = T;
    }
    return make_float3(cov.m[0][0], cov.m[0][1], cov.m[1][1]);
}

// auxiliary.h:49-59 with `grid` as plain ints.  max_radius is the float radius
// converted to int at the call (forward.cu:320 passes a float to an int parameter).
GSR_HD void get_rect(float px, float py, int max_radius, int gx, int gy,
                     uint32_t& minx, uint32_t& miny, uint32_t& maxx, uint32_t& maxy) {
    minx = (uint32_t)min(gx, max((int)0, (int)((px - max_radius) / GSR_TILE_X)));
    miny = (uint32_t)min(gy, max((int)0, (int)((py - max_radius) / GSR_TILE_Y)));
    maxx = (uint32_t)min(gx, max((int)0, (int)((px + max_radius + GSR_TILE_X - 1) / GSR_TILE_X)));
    maxy = (uint32_t)min(gy, max((int)0, (int)((py + max_radius + GSR_TILE_Y - 1) / GSR_TILE_Y)));
}

// SH constants, auxiliary.h:23-40.
#define GSR_SH_C0 0.28209479177387814f
#define GSR_SH_C1 0.4886025119029199f
#define GSR_SH_C2_0 1.0925484305920792f
#define GSR_SH_C2_1 -1.0925484305920792f
#define GSR_SH_C2_2 0.31539156525252005f
#define GSR_SH_C2_3 -1.0925484305920792f
#define GSR_SH_C2_4 0.5462742152960396f
#define GSR_SH_C3_0 -0.5900435899266435f
#define GSR_SH_C3_1 2.890611442640554f
#define GSR_SH_C3_2 -0.4570457994644658f
#define GSR_SH_C3_3 0.3731763325901154f
#define GSR_SH_C3_4 -0.4570457994644658f
#define GSR_SH_C3_5 1.445305721320277f
#define GSR_SH_C3_6 -0.5900435899266435f

// ---- the model's activations (scene/gaussian_model.py:32-55), applied in registers when GsrParams.raw_inputs asks
// for it.  Same float functions torch's CUDA kernels call: expf for torch.exp, 1 / (1 + expf(-x)) for torch.sigmoid,
// x / max(||x||_2, 1e-12) for torch.nn.functional.normalize.
__device__ __forceinline__ float3 act_scale(float3 raw) { return make_float3(expf(raw.x), expf(raw.y), expf(raw.z)); }
__device__ __forceinline__ float act_opacity(float raw) { return 1.0f / (1.0f + expf(-raw)); }
__device__ __forceinline__ float4 act_rotation(float4 raw, float* inv_norm) {
    const float n = sqrtf(raw.x * raw.x + raw.y * raw.y + raw.z * raw.z + raw.w * raw.w);
    const float inv = 1.0f / fmaxf(n, 1e-12f);
    if (inv_norm) *inv_norm = inv;
    return make_float4(raw.x * inv, raw.y * inv, raw.z * inv, raw.w * inv);
}
// their backward: gradient w.r.t. the raw parameter from the gradient w.r.t. the activated one
__device__ __forceinline__ float3 act_scale_bwd(float3 s, float3 g) { return make_float3(g.x * s.x, g.y * s.y, g.z * s.z); }
__device__ __forceinline__ float act_opacity_bwd(float o, float g) { return g * o * (1.0f - o); }
__device__ __forceinline__ float4 act_rotation_bwd(float4 q, float inv_norm, float4 g) {
    const float d = q.x * g.x + q.y * g.y + q.z * g.z + q.w * g.w;      // the component along q does not move q / |q|
    return make_float4((g.x - q.x * d) * inv_norm, (g.y - q.y * d) * inv_norm, (g.z - q.z * d) * inv_norm,
                       (g.w - q.w * d) * inv_norm);
}

}  // namespace gsr
