// render_fwd.cu — kernel 3 of the path: per-tile front-to-back alpha blending.
//
// Replaces renderCUDA forward (DGR/cuda_rasterizer/forward.cu:367-513).  One CTA per
// 16x16 tile, one thread per pixel, warps laid out as 8x4-pixel sub-rectangles.  The tile's
// depth-ordered Gaussian records are gathered into shared memory by bulk async copies (one
// 48 B `cp.async.bulk` per entry, completing on an mbarrier), double-buffered so batch k+1
// lands while batch k is blended.  Each landed batch is culled per sub-rectangle
// (render_common.cuh) and every warp walks only its own compacted entry list.  A warp leaves
// the inner loop as soon as all of its 32 pixels are saturated; the CTA stops fetching when
// all 256 are (block vote at batch granularity, as the reference).
//
// The per-pixel arithmetic keeps the reference's expression order: the skip tests
// (power > 0, alpha < 1/255, T*(1-alpha) < 1e-4) are discontinuities, so a 1-ulp
// difference in alpha would change n_contrib and move the image by up to 4e-3.
#include "common.cuh"
#include "async_copy.cuh"
#include "render_common.cuh"

namespace gsr {

constexpr int RF_BATCH = TILE_PIX;         // entries per stage (one gather per thread)
constexpr int RF_REC_BYTES = 48;           // q0,q1,q2 of GsRecord
constexpr int RF_STAGES = 2;

__global__ void __launch_bounds__(TILE_PIX)
render_fwd_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                  int W, int H, const GsRecord* __restrict__ rec,
                  const float* __restrict__ bg_color,
                  float* __restrict__ final_T, uint32_t* __restrict__ n_contrib,
                  uint32_t* __restrict__ tile_max_contrib,
                  float* __restrict__ out_color, float* __restrict__ out_invdepth,
                  float4* __restrict__ acc_clear, unsigned long long acc_vec4) {
    __shared__ __align__(128) float4 s_rec[RF_STAGES][RF_BATCH * 3];
    __shared__ __align__(8) uint64_t s_bar[RF_STAGES];
    __shared__ __align__(16) uint32_t s_bits[RF_STAGES][CULL_WARPS][CULL_WARPS];   // [stage][source warp][target warp]
    __shared__ uint32_t s_max;

    const uint32_t rank = threadIdx.x;
    const uint32_t lane = rank & 31u, warp = rank >> 5;
    const uint32_t horizontal_blocks = (W + TILE_X - 1) / TILE_X;
    const uint32_t tile = blockIdx.y * horizontal_blocks + blockIdx.x;
    const uint2 pix = {blockIdx.x * TILE_X + sub_px(rank), blockIdx.y * TILE_Y + sub_py(rank)};
    const uint32_t pix_id = W * pix.y + pix.x;
    const float2 pixf = {(float)pix.x, (float)pix.y};
    const float tx0 = (float)(blockIdx.x * TILE_X), ty0 = (float)(blockIdx.y * TILE_Y);
    const bool inside = pix.x < (uint32_t)W && pix.y < (uint32_t)H;
    bool done = !inside;

    const uint2 range = ranges[tile];
    const int n = (int)(range.y - range.x);
    const int rounds = (n + RF_BATCH - 1) / RF_BATCH;

    // GSR_ACC_CLEAR_IN_FORWARD: this kernel is bound by instruction issue and leaves most of the DRAM bandwidth idle,
    // so the zero fill of the backward's gradient accumulator (48 B per Gaussian, a 25 us memset at 3M Gaussians)
    // rides along: every CTA clears a slice with 16 B stores that nothing below waits for
    if (acc_clear != nullptr) {
        const unsigned long long step = (unsigned long long)gridDim.x * gridDim.y * TILE_PIX;
        for (unsigned long long i = (unsigned long long)tile * TILE_PIX + rank; i < acc_vec4; i += step)
            acc_clear[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }

    if (rank == 0) {
        mbar_init(smem_u32(&s_bar[0]), 1);
        mbar_init(smem_u32(&s_bar[1]), 1);
        mbar_fence_init();
        s_max = 0;
    }
    __syncthreads();

    // gather of batch k into stage k&1: every thread copies one record, thread 0 posts
    // the expected byte count (the only arrival of the phase).
    auto issue = [&](int k) {
        const int stage = k & 1;
        const int cnt = min(RF_BATCH, n - k * RF_BATCH);
        const uint32_t bar = smem_u32(&s_bar[stage]);
        if ((int)rank < cnt) {
            const uint32_t id = point_list[range.x + k * RF_BATCH + rank];
            bulk_g2s(smem_u32(&s_rec[stage][rank * 3]), rec + id, RF_REC_BYTES, bar);
        }
        if (rank == 0) mbar_arrive_expect_tx(bar, (uint32_t)cnt * RF_REC_BYTES);
    };

    float T = 1.0f;
    uint32_t last_contributor = 0;
    float C[3] = {0.f, 0.f, 0.f};
    float expected_invdepth = 0.0f;
    bool warp_done = false;                    // all 32 pixels of this warp are saturated

    if (rounds > 0) issue(0);
    for (int k = 0; k < rounds; k++) {
        const int stage = k & 1;
        const int cnt = min(RF_BATCH, n - k * RF_BATCH);
        const float4* sr = s_rec[stage];
        mbar_wait(smem_u32(&s_bar[stage]), (uint32_t)((k >> 1) & 1));   // batch k has landed
        {
            uint32_t m = 0;
            if ((int)rank < cnt) {
                const float4 e0 = sr[rank * 3], e2 = sr[rank * 3 + 2];
                m = cull_mask(e0.x, e0.y, e2.z, e2.w, tx0, ty0);
            }
            publish_cull_bits(reinterpret_cast<uint4*>(s_bits[stage]), m, lane, warp);
        }
        // all threads are past batch k-1 here, so its stage may be refilled; nothing is
        // in flight when the CTA leaves through the break
        const int num_done = __syncthreads_count(done);
        if (num_done == TILE_PIX) break;
        if (k + 1 < rounds) issue(k + 1);

        const uint32_t base = (uint32_t)k * RF_BATCH;
        // Warp-uniform loop over this warp's entries (set bits, in list order), branch-free body: the reference's
        // rejections (power > 0, alpha < 1/255, pixel already saturated; forward.cu:468-487) zero the blend weight
        // instead of skipping, which leaves C, T and last_contributor bit-identical (fma(T, 0, C) == C).
        for (int w8 = 0; w8 * 32 < cnt && !warp_done; w8++) {
            // redux.sync: the word lands in a uniform register, so the bit walk below is uniform control flow
            uint32_t bits = __reduce_or_sync(0xffffffffu, s_bits[stage][w8][warp]);
            while (bits != 0) {
                const int j = w8 * 32 + __ffs((int)bits) - 1;
                bits &= bits - 1;
                const float4 q0 = sr[j * 3 + 0];      // x, y, conic.x, conic.y
                const float4 q1 = sr[j * 3 + 1];      // conic.z, opacity, r, g
                const float2 d = {q0.x - pixf.x, q0.y - pixf.y};
                const float power = -0.5f * (q0.z * d.x * d.x + q1.x * d.y * d.y) - q0.w * d.x * d.y;
                const float alpha = fminf(0.99f, q1.y * expf(power));
                bool ok = !done && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
                const float test_T = T * (1 - alpha);
                const bool stop = ok && test_T < 0.0001f;
                done = done || stop;
                ok = ok && !stop;
                if (!__any_sync(0xffffffffu, ok)) {
                    if (__all_sync(0xffffffffu, done)) { warp_done = true; break; }
                    continue;
                }
                const float2 q2 = *reinterpret_cast<const float2*>(&sr[j * 3 + 2]);      // b, 1/depth
                const float a = ok ? alpha : 0.f;
                C[0] += q1.z * a * T;
                C[1] += q1.w * a * T;
                C[2] += q2.x * a * T;
                expected_invdepth += q2.y * a * T;
                T = ok ? test_T : T;
                last_contributor = ok ? base + j + 1 : last_contributor;
            }
        }
    }

    if (inside) {
        final_T[pix_id] = T;
        n_contrib[pix_id] = last_contributor;
        const size_t HW = (size_t)H * W;
        out_color[0 * HW + pix_id] = C[0] + T * bg_color[0];
        out_color[1 * HW + pix_id] = C[1] + T * bg_color[1];
        out_color[2 * HW + pix_id] = C[2] + T * bg_color[2];
        out_invdepth[pix_id] = expected_invdepth;
    }
    // backward starts its reverse traversal at the deepest contributor of the tile
    uint32_t m = last_contributor;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if (lane == 0 && m > 0) atomicMax(&s_max, m);
    __syncthreads();
    if (rank == 0) tile_max_contrib[tile] = s_max;
}

// Work-descriptor counting pass: independent of the kernel under test, one
// thread per pixel reading the records straight from global memory.  counts[0] += entries
// visited by the forward blend (Wf), counts[1] += blended (pixel, Gaussian) pairs (Bp),
// counts[2] += sum of n_contrib (Wb).
__global__ void __launch_bounds__(TILE_PIX)
count_work_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W, int H,
                  const GsRecord* __restrict__ rec, unsigned long long* __restrict__ counts) {
    const uint32_t horizontal_blocks = (W + TILE_X - 1) / TILE_X;
    const uint32_t tile = blockIdx.y * horizontal_blocks + blockIdx.x;
    const uint2 pix = {blockIdx.x * TILE_X + threadIdx.x, blockIdx.y * TILE_Y + threadIdx.y};
    const float2 pixf = {(float)pix.x, (float)pix.y};
    const bool inside = pix.x < (uint32_t)W && pix.y < (uint32_t)H;
    const uint2 range = ranges[tile];
    unsigned long long visited = 0, blended = 0, last = 0;
    float T = 1.0f;
    if (inside) {
        for (uint32_t i = range.x; i < range.y; i++) {
            visited++;
            const GsRecord r = rec[point_list[i]];
            const float2 d = {r.q0.x - pixf.x, r.q0.y - pixf.y};
            const float power = -0.5f * (r.q0.z * d.x * d.x + r.q1.x * d.y * d.y) - r.q0.w * d.x * d.y;
            if (power > 0.0f) continue;
            const float alpha = fminf(0.99f, r.q1.y * expf(power));
            if (alpha < 1.0f / 255.0f) continue;
            const float test_T = T * (1 - alpha);
            if (test_T < 0.0001f) break;
            T = test_T;
            blended++;
            last = visited;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        visited += __shfl_xor_sync(0xffffffffu, visited, o);
        blended += __shfl_xor_sync(0xffffffffu, blended, o);
        last += __shfl_xor_sync(0xffffffffu, last, o);
    }
    if (((threadIdx.y * TILE_X + threadIdx.x) & 31) == 0) {
        atomicAdd(&counts[0], visited);
        atomicAdd(&counts[1], blended);
        atomicAdd(&counts[2], last);
    }
}

int launch_count_work(const GsrParams& p, GeomState g, ImgState im, BinState b, unsigned long long* counts,
                      cudaStream_t st) {
    GSR_CUDA_TRY(cudaMemsetAsync(counts, 0, 3 * sizeof(unsigned long long), st));
    dim3 grid((p.W + TILE_X - 1) / TILE_X, (p.H + TILE_Y - 1) / TILE_Y, 1);
    dim3 block(TILE_X, TILE_Y, 1);
#ifndef GSR_HOST_EMUL
    count_work_kernel<<<grid, block, 0, st>>>(im.ranges, b.point_list, p.W, p.H, g.rec, counts);
#else
    emu::launch(grid, block, [&]() { count_work_kernel(im.ranges, b.point_list, p.W, p.H, g.rec, counts); });
#endif
    return check_stage("count work", true, st);
}

int launch_render_forward(const GsrParams& p, GeomState g, ImgState im, BinState b,
                          float* out_color, float* out_invdepth, float* acc_clear, cudaStream_t st) {
    StageTimer tm(GSR_STAGE_RENDER_FWD, st);
    dim3 grid((p.W + TILE_X - 1) / TILE_X, (p.H + TILE_Y - 1) / TILE_Y, 1);
    dim3 block(TILE_PIX, 1, 1);
    GSR_LAUNCH(render_fwd_kernel, grid, block, 0, st, im.ranges, b.point_list, p.W, p.H, g.rec, p.background,
                                              im.final_T, im.n_contrib, im.tile_max_contrib,
                                              out_color, out_invdepth, reinterpret_cast<float4*>(acc_clear),
                                              (unsigned long long)p.P * (ACC_FLOATS / 4));
    return check_stage("render forward", p.debug, st);
}

}  // namespace gsr
