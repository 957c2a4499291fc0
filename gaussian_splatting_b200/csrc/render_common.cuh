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
// Each Gaussian gets the axis-aligned bounding box of that ellipse, inflated by a margin that
// covers the fp32 rounding of `power` and of expf (derivation in DESIGN.md §4.3); each staged entry
// an 8-bit mask of the sub-rectangles the box overlaps.  A warp only visits entries whose bit is set.
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

// Half-extents (ex, ey) of the axis-aligned box around the ellipse  alpha >= 1/255  of one Gaussian with conic
// (a, b, c) and blend opacity o, margins included.  Depends on the Gaussian only, so preprocess computes it once
// and stores it in the record (q2.z, q2.w); the blend kernels test boxes, they never evaluate log / sqrt.
//   +inf : no culling (degenerate conic, NaNs)        -inf : culled everywhere (o < 1/255)
__device__ __forceinline__ float2 cull_extent(float a, float b, float c, float o) {
    const float INF = __int_as_float(0x7f800000);
    const float det = a * c - b * b;
    // |power_fp32 - (-q/2)| <= ~2.4e-7 * (a dx^2 + c dy^2) and a dx^2 + c dy^2 <= q * 2ac/det
    const float k = 1e-6f * (a * c) / det;
    if (!(det > 0.f) || !(k < 0.25f)) return make_float2(INF, INF);
    const float t2 = 2.0f * logf(255.0f * o) + 1e-4f;
    if (t2 < 0.f) return make_float2(-INF, -INF);          // o < 1/255: alpha < 1/255 everywhere
    if (!(t2 < 1e30f)) return make_float2(INF, INF);
    const float t2p = t2 / (1.0f - 2.0f * k) * 1.001f;     // q (1/2 - k) <= t2/2 + eps
    const float inv_det = 1.0f / det;
    return make_float2(sqrtf(t2p * c * inv_det) * 1.0001f + 0.01f, sqrtf(t2p * a * inv_det) * 1.0001f + 0.01f);
}

// Mask of the 8 sub-rectangles (bit w = warp w) of the tile at pixel origin (tx0, ty0) that the box
// (x +- ex, y +- ey) overlaps.  Conservative: "not outside" keeps NaNs.
__device__ __forceinline__ uint32_t cull_mask(float x, float y, float ex, float ey, float tx0, float ty0) {
    const float xr = x - tx0, yr = y - ty0;                // relative to the tile origin
    const float x_lo = xr - ex, x_hi = xr + ex;
    const float y_lo = yr - ey, y_hi = yr + ey;
    // column c covers x in [8c, 8c+7], row r covers y in [4r, 4r+3]
    uint32_t cols = 0, rows = 0;
#pragma unroll
    for (int i = 0; i < 2; i++)
        if (!(x_hi < (float)(i * SUB_W) || x_lo > (float)(i * SUB_W + SUB_W - 1))) cols |= 1u << i;
#pragma unroll
    for (int i = 0; i < 4; i++)
        if (!(y_hi < (float)(i * SUB_H) || y_lo > (float)(i * SUB_H + SUB_H - 1))) rows |= 3u << (2 * i);
    // bit w = (row w>>1, column w&1)
    return rows & (cols * 0x55u);
}

// Per-warp entry lists of a staged batch as bit words.  Thread r has the 8-bit warp mask `m` of entry r; the
// warp that computed the masks of entries 32w..32w+31 turns them into one 32-bit word per target warp
// (8 ballots) and its lane 0 stores the eight words bits[w][0..7] (two 16 B stores).  After a block barrier
// warp t walks the set bits of words bits[0..7][t] in order: no per-warp compaction pass over the batch.
constexpr int CULL_WARPS = TILE_PIX / 32;
static_assert(CULL_WARPS == 8, "two uint4 per source warp");
__device__ __forceinline__ void publish_cull_bits(uint4* bits /* [CULL_WARPS][2] */, uint32_t m, uint32_t lane, uint32_t warp) {
    uint32_t b[CULL_WARPS];
#pragma unroll
    for (int t = 0; t < CULL_WARPS; t++) b[t] = __ballot_sync(0xffffffffu, (m & (1u << t)) != 0u);
    if (lane == 0) {
        bits[warp * 2 + 0] = make_uint4(b[0], b[1], b[2], b[3]);
        bits[warp * 2 + 1] = make_uint4(b[4], b[5], b[6], b[7]);
    }
}

}  // namespace gsr
