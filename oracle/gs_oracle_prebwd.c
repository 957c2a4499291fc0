/*
 * gs_oracle_prebwd.c — CPU restatement of BACKWARD::preprocess
 * (DGR/cuda_rasterizer/backward.cu:153-290 computeCov2DCUDA, :296-365 computeCov3D,
 *  :372-429 preprocessCUDA, :12-146 computeColorFromSH backward).
 * TEST INFRASTRUCTURE ONLY (see the header of gs_oracle.c, which this unit includes for
 * the shared static helpers).
 *
 * `restored` selects the semantics:
 *   0  as shipped: computeCov2DCUDA stops after dL_dc_xx/xy/yy (backward.cu:285-290), so
 *      dL_dcov3D stays zero (=> zero scale/rotation gradients) and dL_dmeans3D has only the
 *      projection and SH terms;
 *   1  with the elided tail restored from the chain rule of cov2D = T^T Vrk^T T
 *      (SURVEY.md App. A.6).
 */
#include "gs_oracle.c"

static float sq(float x) { return x * x; }

/* auxiliary.h:128-139 */
static void dnormvdv3(const float* v, const float* dv, float* o) {
    float sum2 = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
    float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
    o[0] = ((+sum2 - v[0] * v[0]) * dv[0] - v[1] * v[0] * dv[1] - v[2] * v[0] * dv[2]) * invsum32;
    o[1] = (-v[0] * v[1] * dv[0] + (sum2 - v[1] * v[1]) * dv[1] - v[2] * v[1] * dv[2]) * invsum32;
    o[2] = (-v[0] * v[2] * dv[0] - v[1] * v[2] * dv[1] + (sum2 - v[2] * v[2]) * dv[2]) * invsum32;
}

/* backward.cu:12-146 */
static void shBackward(int deg, int max_coeffs, const float* pos, const float* campos, const float* sh,
                       const uint8_t* clamped, const float* dL_dcolor, float* dL_dmean, float* dL_dsh) {
    float dir_orig[3] = {pos[0] - campos[0], pos[1] - campos[1], pos[2] - campos[2]};
    float len = sqrtf(dir_orig[0] * dir_orig[0] + dir_orig[1] * dir_orig[1] + dir_orig[2] * dir_orig[2]);
    float x = dir_orig[0] / len, y = dir_orig[1] / len, z = dir_orig[2] / len;
    float dL_dRGB[3];
    for (int c = 0; c < 3; c++) dL_dRGB[c] = dL_dcolor[c] * (clamped[c] ? 0 : 1);
    float w[16];
    memset(w, 0, sizeof(w));
    float dRGBdx[3] = {0, 0, 0}, dRGBdy[3] = {0, 0, 0}, dRGBdz[3] = {0, 0, 0};
#define SH(k) sh[(k) * 3 + c]
    w[0] = SH_C0;
    if (deg > 0) {
        w[1] = -SH_C1 * y;
        w[2] = SH_C1 * z;
        w[3] = -SH_C1 * x;
        for (int c = 0; c < 3; c++) {
            dRGBdx[c] = -SH_C1 * SH(3);
            dRGBdy[c] = -SH_C1 * SH(1);
            dRGBdz[c] = SH_C1 * SH(2);
        }
        if (deg > 1) {
            float xx = x * x, yy = y * y, zz = z * z;
            float xy = x * y, yz = y * z, xz = x * z;
            w[4] = SH_C2[0] * xy;
            w[5] = SH_C2[1] * yz;
            w[6] = SH_C2[2] * (2.f * zz - xx - yy);
            w[7] = SH_C2[3] * xz;
            w[8] = SH_C2[4] * (xx - yy);
            for (int c = 0; c < 3; c++) {
                dRGBdx[c] += SH_C2[0] * y * SH(4) + SH_C2[2] * 2.f * -x * SH(6) + SH_C2[3] * z * SH(7) + SH_C2[4] * 2.f * x * SH(8);
                dRGBdy[c] += SH_C2[0] * x * SH(4) + SH_C2[1] * z * SH(5) + SH_C2[2] * 2.f * -y * SH(6) + SH_C2[4] * 2.f * -y * SH(8);
                dRGBdz[c] += SH_C2[1] * y * SH(5) + SH_C2[2] * 2.f * 2.f * z * SH(6) + SH_C2[3] * x * SH(7);
            }
            if (deg > 2) {
                w[9] = SH_C3[0] * y * (3.f * xx - yy);
                w[10] = SH_C3[1] * xy * z;
                w[11] = SH_C3[2] * y * (4.f * zz - xx - yy);
                w[12] = SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy);
                w[13] = SH_C3[4] * x * (4.f * zz - xx - yy);
                w[14] = SH_C3[5] * z * (xx - yy);
                w[15] = SH_C3[6] * x * (xx - 3.f * yy);
                for (int c = 0; c < 3; c++) {
                    dRGBdx[c] += (SH_C3[0] * SH(9) * 3.f * 2.f * xy + SH_C3[1] * SH(10) * yz +
                                  SH_C3[2] * SH(11) * -2.f * xy + SH_C3[3] * SH(12) * -3.f * 2.f * xz +
                                  SH_C3[4] * SH(13) * (-3.f * xx + 4.f * zz - yy) + SH_C3[5] * SH(14) * 2.f * xz +
                                  SH_C3[6] * SH(15) * 3.f * (xx - yy));
                    dRGBdy[c] += (SH_C3[0] * SH(9) * 3.f * (xx - yy) + SH_C3[1] * SH(10) * xz +
                                  SH_C3[2] * SH(11) * (-3.f * yy + 4.f * zz - xx) + SH_C3[3] * SH(12) * -3.f * 2.f * yz +
                                  SH_C3[4] * SH(13) * -2.f * xy + SH_C3[5] * SH(14) * -2.f * yz +
                                  SH_C3[6] * SH(15) * -3.f * 2.f * xy);
                    dRGBdz[c] += (SH_C3[1] * SH(10) * xy + SH_C3[2] * SH(11) * 4.f * 2.f * yz +
                                  SH_C3[3] * SH(12) * 3.f * (2.f * zz - xx - yy) + SH_C3[4] * SH(13) * 4.f * 2.f * xz +
                                  SH_C3[5] * SH(14) * (xx - yy));
                }
            }
        }
    }
#undef SH
    int ncoef = (deg + 1) * (deg + 1);
    (void)max_coeffs;
    for (int k = 0; k < ncoef; k++)
        for (int c = 0; c < 3; c++) dL_dsh[k * 3 + c] = w[k] * dL_dRGB[c];
    float dL_ddir[3] = {dRGBdx[0] * dL_dRGB[0] + dRGBdx[1] * dL_dRGB[1] + dRGBdx[2] * dL_dRGB[2],
                        dRGBdy[0] * dL_dRGB[0] + dRGBdy[1] * dL_dRGB[1] + dRGBdy[2] * dL_dRGB[2],
                        dRGBdz[0] * dL_dRGB[0] + dRGBdz[1] * dL_dRGB[1] + dRGBdz[2] * dL_dRGB[2]};
    float dm[3];
    dnormvdv3(dir_orig, dL_ddir, dm);
    dL_dmean[0] += dm[0];
    dL_dmean[1] += dm[1];
    dL_dmean[2] += dm[2];
}

/* backward.cu:296-365 */
static void cov3DBackward(const float* scale, float mod, const float* rot, const float* dL_dcov3D,
                          float* dL_dscale, float* dL_drot) {
    float r = rot[0], x = rot[1], y = rot[2], z = rot[3];
    mat3 R = m3(1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y),
                2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x),
                2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y));
    mat3 S = m3(1, 0, 0, 0, 1, 0, 0, 0, 1);
    float s[3] = {mod * scale[0], mod * scale[1], mod * scale[2]};
    S.m[0][0] = s[0]; S.m[1][1] = s[1]; S.m[2][2] = s[2];
    mat3 M = m3mul(S, R);
    mat3 dL_dSigma = m3(dL_dcov3D[0], 0.5f * dL_dcov3D[1], 0.5f * dL_dcov3D[2],
                        0.5f * dL_dcov3D[1], dL_dcov3D[3], 0.5f * dL_dcov3D[4],
                        0.5f * dL_dcov3D[2], 0.5f * dL_dcov3D[4], dL_dcov3D[5]);
    mat3 M2 = M;
    for (int c = 0; c < 3; c++) for (int rr = 0; rr < 3; rr++) M2.m[c][rr] = 2.0f * M.m[c][rr];
    mat3 dL_dM = m3mul(M2, dL_dSigma);
    mat3 Rt = m3t(R);
    mat3 D = m3t(dL_dM); /* dL_dMt */
    for (int k = 0; k < 3; k++)
        dL_dscale[k] = Rt.m[k][0] * D.m[k][0] + Rt.m[k][1] * D.m[k][1] + Rt.m[k][2] * D.m[k][2];
    for (int k = 0; k < 3; k++) for (int rr = 0; rr < 3; rr++) D.m[k][rr] *= s[k];
    dL_drot[0] = 2 * z * (D.m[0][1] - D.m[1][0]) + 2 * y * (D.m[2][0] - D.m[0][2]) + 2 * x * (D.m[1][2] - D.m[2][1]);
    dL_drot[1] = 2 * y * (D.m[1][0] + D.m[0][1]) + 2 * z * (D.m[2][0] + D.m[0][2]) + 2 * r * (D.m[1][2] - D.m[2][1]) - 4 * x * (D.m[2][2] + D.m[1][1]);
    dL_drot[2] = 2 * x * (D.m[1][0] + D.m[0][1]) + 2 * r * (D.m[2][0] - D.m[0][2]) + 2 * z * (D.m[1][2] + D.m[2][1]) - 4 * y * (D.m[2][2] + D.m[0][0]);
    dL_drot[3] = 2 * r * (D.m[0][1] - D.m[1][0]) + 2 * x * (D.m[2][0] + D.m[0][2]) + 2 * y * (D.m[1][2] + D.m[2][1]) - 4 * z * (D.m[1][1] + D.m[0][0]);
}

/*
 * BACKWARD::preprocess (backward.cu:614-657).  Inputs dL_dmean2D (P,3), dL_dconic (P,4),
 * dL_dinvdepth (P or NULL), dL_dcolor (P,3) come from gso_render_backward.  dL_dopacity is
 * read and rewritten in place (:239-241).  Outputs dL_dmeans (P,3), dL_dcov3D (P,6),
 * dL_dsh (P,M,3), dL_dscale (P,3), dL_drot (P,4) must be zero-initialised by the caller
 * (rasterize_points.cu:186-204) and are only touched for radii > 0.
 */
void gso_preprocess_backward(int P, int D, int M, const float* means3D, const int32_t* radii,
                             const float* shs, const uint8_t* clamped, const float* opacities,
                             const float* scales, const float* rotations, float scale_modifier,
                             const float* cov3Ds, const float* viewmatrix, const float* projmatrix,
                             int W, int H, float tan_fovx, float tan_fovy, const float* campos,
                             const float* dL_dmean2D, const float* dL_dconics, const float* dL_dinvdepth,
                             float* dL_dopacity, float* dL_dmeans, const float* dL_dcolor, float* dL_dcov3D,
                             float* dL_dsh, float* dL_dscale, float* dL_drot, int antialiasing,
                             int restored) {
    const float h_y = H / (2.0f * tan_fovy); /* rasterizer_impl.cu:421-422 */
    const float h_x = W / (2.0f * tan_fovx);
    for (int idx = 0; idx < P; idx++) {
        if (!(radii[idx] > 0)) continue;
        /* ---------------- computeCov2DCUDA, backward.cu:153-290 ---------------- */
        const float* cov3D = cov3Ds + 6 * idx;
        const float* mean = means3D + 3 * idx;
        float dL_dconic[3] = {dL_dconics[4 * idx], dL_dconics[4 * idx + 1], dL_dconics[4 * idx + 3]};
        cov2d_ctx ctx;
        float cov[3];
        computeCov2D(mean, h_x, h_y, tan_fovx, tan_fovy, cov3D, viewmatrix, cov, &ctx);
        const float* t = ctx.t;
        const float x_grad_mul = ctx.txtz < -ctx.limx || ctx.txtz > ctx.limx ? 0 : 1;
        const float y_grad_mul = ctx.tytz < -ctx.limy || ctx.tytz > ctx.limy ? 0 : 1;
        float c_xx = cov[0], c_xy = cov[1], c_yy = cov[2];
        const float h_var = 0.3f;
        float d_inside_root = 0.f;
        if (antialiasing) {
            const float det_cov = c_xx * c_yy - c_xy * c_xy;
            c_xx += h_var;
            c_yy += h_var;
            const float det_cov_plus_h_cov = c_xx * c_yy - c_xy * c_xy;
            const float h_convolution_scaling = sqrtf(fmaxf(0.000025f, det_cov / det_cov_plus_h_cov));
            const float dL_dopacity_v = dL_dopacity[idx];
            const float d_h_convolution_scaling = dL_dopacity_v * opacities[idx];
            dL_dopacity[idx] = dL_dopacity_v * h_convolution_scaling;
            d_inside_root = (det_cov / det_cov_plus_h_cov) <= 0.000025f ? 0.f : d_h_convolution_scaling / (2 * h_convolution_scaling);
        } else {
            c_xx += h_var;
            c_yy += h_var;
        }
        float dL_dc_xx = 0, dL_dc_xy = 0, dL_dc_yy = 0;
        if (antialiasing) {
            const float x = c_xx, y = c_yy, z = c_xy, w = h_var;
            const float denom_f = d_inside_root / sq(w * w + w * (x + y) + x * y - z * z);
            dL_dc_xx = w * (w * y + y * y + z * z) * denom_f;
            dL_dc_yy = w * (w * x + x * x + z * z) * denom_f;
            dL_dc_xy = -2.f * w * z * (w + x + y) * denom_f;
        }
        float denom = c_xx * c_yy - c_xy * c_xy;
        float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
        if (denom2inv != 0) {
            dL_dc_xx += denom2inv * (-c_yy * c_yy * dL_dconic[0] + 2 * c_xy * c_yy * dL_dconic[1] + (denom - c_xx * c_yy) * dL_dconic[2]);
            dL_dc_yy += denom2inv * (-c_xx * c_xx * dL_dconic[2] + 2 * c_xx * c_xy * dL_dconic[1] + (denom - c_xx * c_yy) * dL_dconic[0]);
            dL_dc_xy += denom2inv * 2 * (c_xy * c_yy * dL_dconic[0] - (denom + 2 * c_xy * c_xy) * dL_dconic[1] + c_xx * c_xy * dL_dconic[2]);
        }
        if (restored) {
            /* ---- elided tail (backward.cu:285-290), SURVEY.md App. A.6 ---- */
            const float dL_da = dL_dc_xx, dL_db = dL_dc_xy, dL_dc = dL_dc_yy;
            const mat3 T = ctx.T; /* GLM indexing T.m[c][r] */
            float* dc = dL_dcov3D + 6 * idx;
            dc[0] = (T.m[0][0] * T.m[0][0] * dL_da + T.m[0][0] * T.m[1][0] * dL_db + T.m[1][0] * T.m[1][0] * dL_dc);
            dc[3] = (T.m[0][1] * T.m[0][1] * dL_da + T.m[0][1] * T.m[1][1] * dL_db + T.m[1][1] * T.m[1][1] * dL_dc);
            dc[5] = (T.m[0][2] * T.m[0][2] * dL_da + T.m[0][2] * T.m[1][2] * dL_db + T.m[1][2] * T.m[1][2] * dL_dc);
            dc[1] = 2 * T.m[0][0] * T.m[0][1] * dL_da + (T.m[0][0] * T.m[1][1] + T.m[0][1] * T.m[1][0]) * dL_db + 2 * T.m[1][0] * T.m[1][1] * dL_dc;
            dc[2] = 2 * T.m[0][0] * T.m[0][2] * dL_da + (T.m[0][0] * T.m[1][2] + T.m[0][2] * T.m[1][0]) * dL_db + 2 * T.m[1][0] * T.m[1][2] * dL_dc;
            dc[4] = 2 * T.m[0][2] * T.m[0][1] * dL_da + (T.m[0][1] * T.m[1][2] + T.m[0][2] * T.m[1][1]) * dL_db + 2 * T.m[1][1] * T.m[1][2] * dL_dc;
            /* Vrk (symmetric) */
            const float V[3][3] = {{cov3D[0], cov3D[1], cov3D[2]}, {cov3D[1], cov3D[3], cov3D[4]}, {cov3D[2], cov3D[4], cov3D[5]}};
            float dL_dT0[3], dL_dT1[3];
            for (int k = 0; k < 3; k++) {
                float v0 = T.m[0][0] * V[k][0] + T.m[0][1] * V[k][1] + T.m[0][2] * V[k][2];
                float v1 = T.m[1][0] * V[k][0] + T.m[1][1] * V[k][1] + T.m[1][2] * V[k][2];
                dL_dT0[k] = 2 * v0 * dL_da + v1 * dL_db;
                dL_dT1[k] = 2 * v1 * dL_dc + v0 * dL_db;
            }
            /* T = W * J with W[c][r] = viewmatrix[4r + c] */
            const float* vm = viewmatrix;
            float dL_dJ00 = vm[0] * dL_dT0[0] + vm[4] * dL_dT0[1] + vm[8] * dL_dT0[2];
            float dL_dJ02 = vm[2] * dL_dT0[0] + vm[6] * dL_dT0[1] + vm[10] * dL_dT0[2];
            float dL_dJ11 = vm[1] * dL_dT1[0] + vm[5] * dL_dT1[1] + vm[9] * dL_dT1[2];
            float dL_dJ12 = vm[2] * dL_dT1[0] + vm[6] * dL_dT1[1] + vm[10] * dL_dT1[2];
            float tz = 1.f / t[2];
            float tz2 = tz * tz;
            float tz3 = tz2 * tz;
            float dL_dt[3];
            dL_dt[0] = x_grad_mul * -h_x * tz2 * dL_dJ02;
            dL_dt[1] = y_grad_mul * -h_y * tz2 * dL_dJ12;
            dL_dt[2] = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * t[0]) * tz3 * dL_dJ02 + (2 * h_y * t[1]) * tz3 * dL_dJ12;
            if (dL_dinvdepth) dL_dt[2] -= dL_dinvdepth[idx] / (t[2] * t[2]);
            float dm[3];
            transformVec4x3Transpose(dL_dt, viewmatrix, dm);
            dL_dmeans[3 * idx + 0] = dm[0];
            dL_dmeans[3 * idx + 1] = dm[1];
            dL_dmeans[3 * idx + 2] = dm[2];
        }
        /* ---------------- preprocessCUDA backward, backward.cu:372-429 ---------------- */
        const float* m = mean;
        const float* proj = projmatrix;
        float m_hom[4];
        transformPoint4x4(m, proj, m_hom);
        float m_w = 1.0f / (m_hom[3] + 0.0000001f);
        float mul1 = (proj[0] * m[0] + proj[4] * m[1] + proj[8] * m[2] + proj[12]) * m_w * m_w;
        float mul2 = (proj[1] * m[0] + proj[5] * m[1] + proj[9] * m[2] + proj[13]) * m_w * m_w;
        const float* g2 = dL_dmean2D + 3 * idx;
        dL_dmeans[3 * idx + 0] += (proj[0] * m_w - proj[3] * mul1) * g2[0] + (proj[1] * m_w - proj[3] * mul2) * g2[1];
        dL_dmeans[3 * idx + 1] += (proj[4] * m_w - proj[7] * mul1) * g2[0] + (proj[5] * m_w - proj[7] * mul2) * g2[1];
        dL_dmeans[3 * idx + 2] += (proj[8] * m_w - proj[11] * mul1) * g2[0] + (proj[9] * m_w - proj[11] * mul2) * g2[1];
        if (shs)
            shBackward(D, M, mean, campos, shs + (size_t)idx * M * 3, clamped + 3 * idx, dL_dcolor + 3 * idx,
                       dL_dmeans + 3 * idx, dL_dsh + (size_t)idx * M * 3);
        if (scales)
            cov3DBackward(scales + 3 * idx, scale_modifier, rotations + 4 * idx, dL_dcov3D + 6 * idx,
                          dL_dscale + 3 * idx, dL_drot + 4 * idx);
    }
}
