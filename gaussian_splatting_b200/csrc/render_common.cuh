// render_common.cuh — pieces shared by the forward and backward blend kernels:
// the pixel mapping of a CTA and the conservative sub-tile culling of batch entries.
//
// A CTA of 256 threads owns one 16x16 tile (as the reference, forward.cu:367-381), but the
// 8 warps are laid out as 2 x 4 sub-rectangles of 8 x 4 pixels instead of 16 x 2 strips:
// a Gaussian's footprint is compact, so compact sub-rectangles see fewer partially covered
// warps (less divergence) and can be culled per warp.
//
// Culling.  The reference evaluates every entry of the tile's list at every pixel and drops
// it when  power > 0  or  alpha = min(0.99, o * exp(power)) < 1/255  (forward.cu:468-480).
// alpha >= 1/255 needs  power >= -ln(255 o), i.e. the pixel lies inside the ellipse
//   q(d) = a dx^2 + 2 b dx dy + c dy^2  <=  t2 = 2 ln(255 o).
// Each staged entry gets the axis-aligned bounding box of that ellipse, inflated by a margin that
// covers the fp32 rounding of `power` and of expf (derivation in DESIGN.md §4.3), and an 8-bit
// mask of the sub-rectangles the box overlaps.  A warp only visits entries whose bit is set.
// Entries skipped this way would have been rejected at all 32 pixels by the reference's own
// tests, so images, n_contrib and gradients are unchanged.
#pragma once
#include "common.cuh"

namespace gsr {

constexpr int SUB_W = 8;     // sub-rectangle of one warp
constexpr int SUB_H = 4;

// pixel of thread `tid` (0..255) inside the tile
__device__ __forceinline__ uint32_t sub_px(uint32_t tid) { return ((tid >> 5) & 1u) * SUB_W + (tid & 7u); }
__device__ __forceinline__ uint32_t sub_py(uint32_t tid) { return (tid >> 6) * SUB_H + ((tid >> 3) & 3u); }

// Mask of the 8 sub-rectangles (bit w = warp w) of the tile at pixel origin (tx0, ty0) that the
// entry {x, y, a, b | c, o, ..} can contribute to.  Conservative; NaNs and degenerate conics
// return 0xFF (no culling).
__device__ __forceinline__ uint32_t cull_mask(const float4 q0, const float4 q1, float tx0, float ty0) {
    const float a = q0.z, b = q0.w, c = q1.x, o = q1.y;
    const float det = a * c - b * b;
    // |power_fp32 - (-q/2)| <= ~2.4e-7 * (a dx^2 + c dy^2) and a dx^2 + c dy^2 <= q * 2ac/det
    const float k = 1e-6f * (a * c) / det;
    if (!(det > 0.f) || !(k < 0.25f)) return 0xFFu;
    const float t2 = 2.0f * logf(255.0f * o) + 1e-4f;
    if (t2 < 0.f) return 0u;                       // o < 1/255: alpha < 1/255 everywhere
    if (!(t2 < 1e30f)) return 0xFFu;
    const float t2p = t2 / (1.0f - 2.0f * k) * 1.001f;   // q (1/2 - k) <= t2/2 + eps
    const float inv_det = 1.0f / det;
    const float ex = sqrtf(t2p * c * inv_det) * 1.0001f + 0.01f;
    const float ey = sqrtf(t2p * a * inv_det) * 1.0001f + 0.01f;
    const float x_lo = q0.x - ex - tx0, x_hi = q0.x + ex - tx0;   // relative to the tile origin
    const float y_lo = q0.y - ey - ty0, y_hi = q0.y + ey - ty0;
    // column c covers x in [8c, 8c+7], row r covers y in [4r, 4r+3]; "not outside" keeps NaNs
    uint32_t cols = 0, rows = 0;
#pragma unroll
    for (int i = 0; i < 2; i++)
        if (!(x_hi < (float)(i * SUB_W) || x_lo > (float)(i * SUB_W + SUB_W - 1))) cols |= 1u << i;
#pragma unroll
    for (int i = 0; i < 4; i++)
        if (!(y_hi < (float)(i * SUB_H) || y_lo > (float)(i * SUB_H + SUB_H - 1))) rows |= 1u << i;
    // bit w = (row w>>1, column w&1)
    uint32_t m = 0;
#pragma unroll
    for (int r = 0; r < 4; r++)
        if ((rows >> r) & 1u) m |= cols << (2 * r);
    return m;
}

}  // namespace gsr
