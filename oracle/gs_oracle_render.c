/*
 * gs_oracle_render.c — CPU restatement of the reference's tile blend, forward and backward.
 * TEST INFRASTRUCTURE ONLY (see the header of gs_oracle.c).
 * Follows DGR/cuda_rasterizer/forward.cu:367-513 and backward.cu:433-612.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include "gso_parallel.h"

#define BLOCK_X 16
#define BLOCK_Y 16
#define GSO_MAX_THREADS 256

typedef struct {
    int W, H, gx;
    const uint32_t *ranges, *point_list;
    const float *means2D, *features, *conic_opacity, *depths, *bg_color;
    float* final_T;
    uint32_t* n_contrib;
    float *out_color, *invdepth;
    int64_t visited[GSO_MAX_THREADS], blended[GSO_MAX_THREADS];
} fwd_ctx;

static void fwd_tiles(int lo, int hi, int tid, void* vp) {
    fwd_ctx* c = (fwd_ctx*)vp;
    const int W = c->W, H = c->H;
    int64_t visited = 0, blended = 0;
    for (int tile = lo; tile < hi; tile++) {
        const int bx = tile % c->gx, by = tile / c->gx;
        const uint32_t r0 = c->ranges[2 * tile], r1 = c->ranges[2 * tile + 1];
        for (int ty = 0; ty < BLOCK_Y; ty++)
            for (int tx = 0; tx < BLOCK_X; tx++) {
                const uint32_t px = bx * BLOCK_X + tx, py = by * BLOCK_Y + ty;
                if (!(px < (uint32_t)W && py < (uint32_t)H)) continue; /* inside, forward.cu:393 */
                const uint32_t pix_id = W * py + px;
                const float pixfx = (float)px, pixfy = (float)py;
                float T = 1.0f;
                uint32_t contributor = 0, last_contributor = 0;
                float C[3] = {0, 0, 0};
                float expected_invdepth = 0.0f;
                for (uint32_t i = r0; i < r1; i++) { /* forward.cu:458-495 */
                    contributor++;
                    const uint32_t id = c->point_list[i];
                    const float dx = c->means2D[2 * id] - pixfx, dy = c->means2D[2 * id + 1] - pixfy;
                    const float* con_o = c->conic_opacity + 4 * id;
                    const float power = -0.5f * (con_o[0] * dx * dx + con_o[2] * dy * dy) - con_o[1] * dx * dy;
                    if (power > 0.0f) continue;
                    float alpha = fminf(0.99f, con_o[3] * expf(power));
                    if (alpha < 1.0f / 255.0f) continue;
                    float test_T = T * (1 - alpha);
                    if (test_T < 0.0001f) break; /* done = true */
                    for (int ch = 0; ch < 3; ch++) C[ch] += c->features[id * 3 + ch] * alpha * T;
                    expected_invdepth += (1 / c->depths[id]) * alpha * T;
                    T = test_T;
                    last_contributor = contributor;
                    blended++;
                }
                visited += contributor;
                c->final_T[pix_id] = T;
                c->n_contrib[pix_id] = last_contributor;
                for (int ch = 0; ch < 3; ch++)
                    c->out_color[(size_t)ch * H * W + pix_id] = C[ch] + T * c->bg_color[ch];
                c->invdepth[pix_id] = expected_invdepth;
            }
    }
    c->visited[tid] += visited;
    c->blended[tid] += blended;
}

/*
 * renderCUDA forward.  The tile loop is the parallel axis.  tile_lo/tile_hi restrict the
 * work to tiles [tile_lo, tile_hi) so a bounded sample of a large frame can be timed
 * (bench.py cpu_baseline); tile_hi < 0 means all tiles.
 * Work counters: stats[0] += entries visited (Wf), stats[1] += blended pairs (Bp).
 */
void gso_render_forward(int W, int H, const uint32_t* ranges, const uint32_t* point_list,
                        const float* means2D, const float* features, const float* conic_opacity,
                        const float* depths, const float* bg_color, float* final_T, uint32_t* n_contrib,
                        float* out_color, float* invdepth, int tile_lo, int tile_hi, int64_t* stats) {
    fwd_ctx* c = (fwd_ctx*)calloc(1, sizeof(fwd_ctx));
    c->W = W; c->H = H; c->gx = (W + BLOCK_X - 1) / BLOCK_X;
    const int gy = (H + BLOCK_Y - 1) / BLOCK_Y;
    if (tile_hi < 0 || tile_hi > c->gx * gy) tile_hi = c->gx * gy;
    c->ranges = ranges; c->point_list = point_list; c->means2D = means2D; c->features = features;
    c->conic_opacity = conic_opacity; c->depths = depths; c->bg_color = bg_color;
    c->final_T = final_T; c->n_contrib = n_contrib; c->out_color = out_color; c->invdepth = invdepth;
    gso_parallel_for(tile_lo, tile_hi, 4, fwd_tiles, c);
    if (stats)
        for (int t = 0; t < GSO_MAX_THREADS; t++) { stats[0] += c->visited[t]; stats[1] += c->blended[t]; }
    free(c);
}

typedef struct {
    int P, W, H, gx;
    const uint32_t *ranges, *point_list, *n_contrib;
    const float *bg_color, *means2D, *conic_opacity, *colors, *depths, *final_Ts, *dL_dpixels, *dL_invdepths;
    double* acc;                       /* P x 10 global sums (fp64) */
    double* local[GSO_MAX_THREADS];    /* per-thread scratch: one tile's range x 10 */
    size_t local_cap[GSO_MAX_THREADS];
} bwd_ctx;

/* fp64 atomic add (CAS loop): tiles processed by different threads share Gaussians. */
static inline void atomic_add_f64(double* addr, double v) {
    uint64_t* p = (uint64_t*)addr;
    uint64_t old = __atomic_load_n(p, __ATOMIC_RELAXED), nw;
    do {
        double d;
        memcpy(&d, &old, 8);
        d += v;
        memcpy(&nw, &d, 8);
    } while (!__atomic_compare_exchange_n(p, &old, nw, 1, __ATOMIC_RELAXED, __ATOMIC_RELAXED));
}

static void bwd_tiles(int lo, int hi, int tid, void* vp) {
    bwd_ctx* c = (bwd_ctx*)vp;
    const int W = c->W, H = c->H;
    const float ddelx_dx = 0.5 * W; /* backward.cu:529-530 */
    const float ddely_dy = 0.5 * H;
    for (int tile = lo; tile < hi; tile++) {
        const int bx = tile % c->gx, by = tile / c->gx;
        const uint32_t r0 = c->ranges[2 * tile], r1 = c->ranges[2 * tile + 1];
        /* per-tile sums indexed by position in the tile's list, flushed once per tile */
        const size_t need = (size_t)(r1 - r0) * 10;
        if (need > c->local_cap[tid]) {
            free(c->local[tid]);
            c->local_cap[tid] = need * 2;
            c->local[tid] = (double*)malloc(c->local_cap[tid] * sizeof(double));
        }
        double* loc = c->local[tid];
        if (need) memset(loc, 0, need * sizeof(double));
        for (int ty = 0; ty < BLOCK_Y; ty++)
            for (int tx = 0; tx < BLOCK_X; tx++) {
                const uint32_t px = bx * BLOCK_X + tx, py = by * BLOCK_Y + ty;
                if (!(px < (uint32_t)W && py < (uint32_t)H)) continue;
                const uint32_t pix_id = W * py + px;
                const float pixfx = (float)px, pixfy = (float)py;
                const float T_final = c->final_Ts[pix_id];
                float T = T_final;
                uint32_t contributor = r1 - r0;
                const int last_contributor = (int)c->n_contrib[pix_id];
                float accum_rec[3] = {0, 0, 0};
                float dL_dpixel[3];
                float dL_invdepth = 0;
                float accum_invdepth_rec = 0;
                for (int i = 0; i < 3; i++) dL_dpixel[i] = c->dL_dpixels[(size_t)i * H * W + pix_id];
                if (c->dL_invdepths) dL_invdepth = c->dL_invdepths[pix_id];
                float last_alpha = 0;
                float last_color[3] = {0, 0, 0};
                float last_invdepth = 0;
                for (uint32_t k = 0; k < r1 - r0; k++) { /* back to front, backward.cu:520 */
                    const uint32_t id = c->point_list[r1 - k - 1];
                    contributor--;
                    if ((int)contributor >= last_contributor) continue;
                    const float dx = c->means2D[2 * id] - pixfx, dy = c->means2D[2 * id + 1] - pixfy;
                    const float* con_o = c->conic_opacity + 4 * id;
                    const float power = -0.5f * (con_o[0] * dx * dx + con_o[2] * dy * dy) - con_o[1] * dx * dy;
                    if (power > 0.0f) continue;
                    const float G = expf(power);
                    const float alpha = fminf(0.99f, con_o[3] * G);
                    if (alpha < 1.0f / 255.0f) continue;
                    T = T / (1.f - alpha);
                    const float dchannel_dcolor = alpha * T;
                    float dL_dalpha = 0.0f;
                    double* a = loc + (size_t)(r1 - r0 - k - 1) * 10;
                    for (int ch = 0; ch < 3; ch++) {
                        const float col = c->colors[id * 3 + ch];
                        accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                        last_color[ch] = col;
                        const float dL_dchannel = dL_dpixel[ch];
                        dL_dalpha += (col - accum_rec[ch]) * dL_dchannel;
                        a[ch] += (double)(dchannel_dcolor * dL_dchannel);
                    }
                    if (c->dL_invdepths) {
                        const float invd = 1.f / c->depths[id];
                        accum_invdepth_rec = last_alpha * last_invdepth + (1.f - last_alpha) * accum_invdepth_rec;
                        last_invdepth = invd;
                        dL_dalpha += (invd - accum_invdepth_rec) * dL_invdepth;
                        a[3] += (double)(dchannel_dcolor * dL_invdepth);
                    }
                    dL_dalpha *= T;
                    last_alpha = alpha;
                    float bg_dot_dpixel = 0;
                    for (int i = 0; i < 3; i++) bg_dot_dpixel += c->bg_color[i] * dL_dpixel[i];
                    dL_dalpha += (-T_final / (1.f - alpha)) * bg_dot_dpixel;
                    const float dL_dG = con_o[3] * dL_dalpha;
                    const float gdx = G * dx;
                    const float gdy = G * dy;
                    const float dG_ddelx = -gdx * con_o[0] - gdy * con_o[1];
                    const float dG_ddely = -gdy * con_o[2] - gdx * con_o[1];
                    a[4] += (double)(dL_dG * dG_ddelx * ddelx_dx);
                    a[5] += (double)(dL_dG * dG_ddely * ddely_dy);
                    a[6] += (double)(-0.5f * gdx * dx * dL_dG);
                    a[7] += (double)(-0.5f * gdx * dy * dL_dG);
                    a[8] += (double)(-0.5f * gdy * dy * dL_dG);
                    a[9] += (double)(G * dL_dalpha);
                }
            }
        for (uint32_t k = 0; k < r1 - r0; k++) {
            const double* a = loc + (size_t)k * 10;
            double* g = c->acc + (size_t)c->point_list[r0 + k] * 10;
            for (int j = 0; j < 10; j++)
                if (a[j] != 0.0) atomic_add_f64(g + j, a[j]);
        }
    }
}

/*
 * renderCUDA backward (backward.cu:433-612).  Per-pair terms are computed in fp32 exactly
 * as the kernel does; the per-Gaussian sums, which the kernel forms with fp32 atomicAdd in
 * nondeterministic order, are accumulated here in fp64 and rounded once.
 * Outputs (fully written): dL_dmean2D (P,3), dL_dconic2D (P,4), dL_dopacity (P),
 * dL_dcolors (P,3), dL_dinvdepths (P) (only if non-NULL; requires dL_invdepths).
 */
void gso_render_backward(int P, int W, int H, const uint32_t* ranges, const uint32_t* point_list,
                         const float* bg_color, const float* means2D, const float* conic_opacity,
                         const float* colors, const float* depths, const float* final_Ts,
                         const uint32_t* n_contrib, const float* dL_dpixels, const float* dL_invdepths,
                         float* dL_dmean2D, float* dL_dconic2D, float* dL_dopacity, float* dL_dcolors,
                         float* dL_dinvdepths, int tile_lo, int tile_hi) {
    bwd_ctx* c = (bwd_ctx*)calloc(1, sizeof(bwd_ctx));
    c->P = P; c->W = W; c->H = H; c->gx = (W + BLOCK_X - 1) / BLOCK_X;
    const int gy = (H + BLOCK_Y - 1) / BLOCK_Y;
    if (tile_hi < 0 || tile_hi > c->gx * gy) tile_hi = c->gx * gy;
    c->ranges = ranges; c->point_list = point_list; c->n_contrib = n_contrib; c->bg_color = bg_color;
    c->means2D = means2D; c->conic_opacity = conic_opacity; c->colors = colors; c->depths = depths;
    c->final_Ts = final_Ts; c->dL_dpixels = dL_dpixels; c->dL_invdepths = dL_invdepths;
    c->acc = (double*)calloc((size_t)(P > 0 ? P : 1) * 10, sizeof(double));
    gso_parallel_for(tile_lo, tile_hi, 4, bwd_tiles, c);
    double* acc = c->acc;
    for (int t = 0; t < GSO_MAX_THREADS; t++) free(c->local[t]);
    for (int id = 0; id < P; id++) {
        static const double zero[10] = {0};
        const double* a = acc ? acc + (size_t)id * 10 : zero;
        dL_dcolors[3 * id + 0] = (float)a[0];
        dL_dcolors[3 * id + 1] = (float)a[1];
        dL_dcolors[3 * id + 2] = (float)a[2];
        if (dL_dinvdepths) dL_dinvdepths[id] = (float)a[3];
        dL_dmean2D[3 * id + 0] = (float)a[4];
        dL_dmean2D[3 * id + 1] = (float)a[5];
        dL_dmean2D[3 * id + 2] = 0.f;
        dL_dconic2D[4 * id + 0] = (float)a[6];
        dL_dconic2D[4 * id + 1] = (float)a[7];
        dL_dconic2D[4 * id + 2] = 0.f;
        dL_dconic2D[4 * id + 3] = (float)a[8];
        dL_dopacity[id] = (float)a[9];
    }
    free(acc);
    free(c);
}
