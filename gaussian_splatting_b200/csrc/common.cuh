// common.cuh — shared types and helpers for the sm_100a Gaussian rasterizer kernels.
//
// Data layout in HBM:
//   GsRecord[P]   64 B per Gaussian, written by preprocess, gathered by the
//                 render kernels with one 48 B (fwd) / 64 B (bwd) bulk copy.
//   rects[P]      packed tile rectangle (4 x u16), consumed by the tile-list construction.
//   point_list[R] Gaussian ids ordered by (tile, depth, id)  == the reference's
//                 sorted value list (rasterizer_impl.cu:335-340).
//   ranges[T]     per-tile [start,end) into point_list (rasterizer_impl.cu:132-164).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/gsrast.h"

namespace gsr {

constexpr int TILE_X = GSR_TILE_X;
constexpr int TILE_Y = GSR_TILE_Y;
constexpr int TILE_PIX = TILE_X * TILE_Y;

// One Gaussian after projection.  16 floats, 64 B, 64 B aligned.
//   q0 = {x, y, conic.x, conic.y}            pixel centre + conic a,b
//   q1 = {conic.z, opacity*aa, r, g}
//   q2 = {b, 1/depth, cull half-extent x, cull half-extent y}     (render_common.cuh: cull_extent)
//   q3 = {clamped bits (int), tiles_touched (int), depth, radius (int bits)}
// The first 48 B are what the blend kernels stage per tile entry.
struct __align__(16) GsRecord {
    float4 q0, q1, q2, q3;
};
static_assert(sizeof(GsRecord) == GSR_RECORD_FLOATS * 4, "record size");

// Gradient accumulator per Gaussian written by render backward (atomics) and
// consumed by preprocess backward.  12 floats, 48 B:
//   {dcolor r,g,b, dinvdepth | S h, S h dx, S h dy, S h dx^2 | S h dx dy, S h dy^2, pad, pad}
// h = G * dL_dalpha, d = centre - pixel, sums over the blended pixels; preprocess_bwd turns the six moments into
// dL_dopacity, dL_dmean2D and dL_dconic with the Gaussian's own conic / opacity.
constexpr int ACC_FLOATS = 12;

__host__ __device__ inline uint32_t div_up(uint32_t a, uint32_t b) { return (a + b - 1) / b; }

// getHigherMsb (rasterizer_impl.cu:36-51): floor(log2 n) + 1 for n > 0.
inline uint32_t higher_msb(uint32_t n) {
    uint32_t msb = sizeof(n) * 4, step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step; else msb -= step;
    }
    if (n >> msb) msb++;
    return msb;
}

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// Bump allocator over a caller-owned chunk (role of `obtain`, rasterizer_impl.h:25-35).
struct Carver {
    char* base; size_t off;
    explicit Carver(void* p) : base(reinterpret_cast<char*>(p)), off(0) {}
    template <typename T> T* take(size_t count) {
        off = align_up(off, 256);
        T* r = reinterpret_cast<T*>(base + off);
        off += count * sizeof(T);
        return r;
    }
    size_t used() const { return align_up(off, 256); }
};

// ---- state layouts -------------------------------------------------------
struct GeomState {          // persists forward -> backward
    GsRecord* rec;          // P
    static GeomState carve(void* chunk, size_t P, size_t* bytes) {
        Carver c(chunk); GeomState g;
        g.rec = c.take<GsRecord>(P);
        if (bytes) *bytes = c.used();
        return g;
    }
};

struct ImgState {           // persists forward -> backward
    float*    final_T;      // N
    uint32_t* n_contrib;    // N
    uint2*    ranges;       // T
    uint32_t* tile_max_contrib; // T : max n_contrib over the tile (backward start point)
    static ImgState carve(void* chunk, size_t N, size_t T, size_t* bytes) {
        Carver c(chunk); ImgState s;
        s.final_T = c.take<float>(N);
        s.n_contrib = c.take<uint32_t>(N);
        s.ranges = c.take<uint2>(T);
        s.tile_max_contrib = c.take<uint32_t>(T);
        if (bytes) *bytes = c.used();
        return s;
    }
};

struct BinState {           // persists forward -> backward
    uint32_t* point_list;   // R
    static BinState carve(void* chunk, size_t R, size_t* bytes) {
        Carver c(chunk); BinState s;
        s.point_list = c.take<uint32_t>(R > 0 ? R : 1);
        if (bytes) *bytes = c.used();
        return s;
    }
};

constexpr int SORT_THREADS = 256;
constexpr int SORT_ITEMS = 16;
constexpr int SORT_CHUNK = SORT_THREADS * SORT_ITEMS;   // 4096 items per block

// Control block of one radix sort (up to 4 digit passes):
//   [0, 1024)            digit histograms of all passes (256 each)
//   [1024, 2048)         exclusive digit bases of all passes
//   [2048, 2048+64)      tile tickets, one per pass (padded)
//   [2112, ...)          look-back status words, pass-major: pass * 256 * nb + tile * 256 + digit
constexpr int SORT_MAX_PASSES = 4;
constexpr size_t SORT_CTL_HIST = 0, SORT_CTL_BASE = 1024, SORT_CTL_TICKET = 2048, SORT_CTL_STATUS = 2112;
inline size_t sort_ctl_words(size_t nb) { return SORT_CTL_STATUS + (size_t)SORT_MAX_PASSES * 256 * nb; }

// Per-tile list construction (binning.cu): a unit of work is (chunk of depth ranks, band of tile rows).
#ifndef GSR_TB_MAX_CELLS
#define GSR_TB_MAX_CELLS 8192                            // (tests lower it to reach several bands at small sizes)
#endif
constexpr int TB_MAX_CELLS = GSR_TB_MAX_CELLS;           // tile counters of one unit in shared memory (32 KB)
constexpr int TB_MAX_BAND_ROWS = 512;
constexpr int TB_MAX_GX = 2048;                          // widest tile grid (32 768 px)
constexpr int TB_COUNT_CELLS = 10304;                    // difference grid: ((gx + 1) | 1) x (band_rows + 1) <= this
constexpr int TB_COUNT_THREADS = 256;
#ifndef GSR_TB_TARGET_UNITS
#define GSR_TB_TARGET_UNITS 888                          // 148 SMs x 6 resident CTAs of the fill kernel (37 KB of shared memory each)
#endif
constexpr int TB_TARGET_UNITS = GSR_TB_TARGET_UNITS;
constexpr int TB_MIN_CHUNK = 64;                         // Gaussians
constexpr unsigned long long TB_MAX_MATRIX_WORDS = 32ull << 20;   // 128 MB of (chunk, tile) counts at most
constexpr int TB_MAX_WBLOCKS = 4096;                     // weight blocks (multiples of 256 depth ranks) used to balance the chunks

struct TileBinPlan { uint32_t gx, gy, T, band_rows, n_bands, n_chunks, stride, wblock, n_wblocks; };
inline TileBinPlan tile_bin_plan(int P, int W, int H) {
    TileBinPlan pl;
    pl.gx = (uint32_t)(W + TILE_X - 1) / TILE_X;
    pl.gy = (uint32_t)(H + TILE_Y - 1) / TILE_Y;
    pl.T = pl.gx * pl.gy;
    uint32_t rows = (uint32_t)TB_MAX_CELLS / (pl.gx ? pl.gx : 1u);
    if (rows < 1) rows = 1;
    if (rows > (uint32_t)TB_MAX_BAND_ROWS) rows = TB_MAX_BAND_ROWS;
    if (rows > pl.gy) rows = pl.gy ? pl.gy : 1u;
    pl.band_rows = rows;
    pl.n_bands = (pl.gy + rows - 1) / rows;
    pl.stride = (pl.gx + 1) | 1u;
    // The list-filling kernel runs band after band (one launch each), so every band gets the full number of chunks;
    // the (chunk, tile) count matrix is capped at TB_MAX_MATRIX_WORDS.
    uint32_t chunks = (uint32_t)TB_TARGET_UNITS;
    if ((unsigned long long)chunks * pl.T > TB_MAX_MATRIX_WORDS) chunks = (uint32_t)(TB_MAX_MATRIX_WORDS / (pl.T ? pl.T : 1u));
    const uint32_t most = (uint32_t)(P > 0 ? P : 1) / TB_MIN_CHUNK;
    if (chunks > most) chunks = most;
    if (chunks < 1) chunks = 1;
    pl.n_chunks = chunks;
    const uint32_t Pn = (uint32_t)(P > 0 ? P : 1);
    pl.wblock = ((Pn + TB_MAX_WBLOCKS - 1) / TB_MAX_WBLOCKS + 255u) / 256u * 256u;
    pl.n_wblocks = (Pn + pl.wblock - 1) / pl.wblock;
    return pl;
}

// device counters (int32[8]): [0..1] num_rendered as one 64-bit sum, [2] visible Gaussians (both accumulated by
// the preprocess kernel)
constexpr int CNT_RENDERED = 0, CNT_VISIBLE = 2;

// band_ctl layout (u32 words): per-band counts, exclusive offsets, fill cursors, exclusive upper tile row
constexpr int MAX_BANDS = 16;
constexpr int BAND_CNT = 0, BAND_OFF = 16, BAND_CUR = 32, BAND_HI = 48, BAND_CTL_WORDS = 64;

struct Scratch {            // only live inside one forward or backward call; does not depend on num_rendered
    // forward
    uint2*    rects;        // P   packed rect {min.x | min.y<<16, max.x | max.y<<16}
    uint64_t* depth_a;      // P   depth_bits<<32 | id   (ping; holds the depth order after the sort)
    uint64_t* depth_b;      // P   (pong)
    uint4*    srects;       // P   {rect.x, rect.y, div_magic(width), id} in depth order
    uint32_t* sort_ctl;     // onesweep control words, see SORT_CTL_* (zeroed per sort)
    uint32_t* hist;         // 256 * blocks(P)   (3-kernel radix pass, P >= 2^30)
    uint32_t* digit_tot;    // 256
    uint32_t* tile_chunk;   // n_chunks * T   per (chunk, tile): count, then first slot relative to the tile's start
    uint32_t* tile_total;   // T
    uint32_t* tile_start;   // T + 1                 first slot of every tile's list (empty tiles included)
    uint32_t* tile_stripes; // n_bands * 32          tile rows at which the filling kernel cuts a band between its warps
    uint32_t* tile_weights; // TB_MAX_WBLOCKS        list-filling steps per weight block of depth ranks
    uint32_t* tile_bounds;  // n_chunks + 1          first depth rank of every chunk
    int32_t*  counters;     // 8   see CNT_*
    // backward
    float*    acc;          // P * ACC_FLOATS
    uint32_t* band_list;    // P   visible Gaussian ids grouped by tile-row band (multimem mode)
    uint32_t* band_ctl;     // BAND_CTL_WORDS: counts, offsets, cursors, band limits
    static Scratch carve(void* chunk, size_t P, int W, int H, size_t* bytes) {
        Carver c(chunk); Scratch s;
        size_t Pn = P > 0 ? P : 1;
        size_t nb = Pn / SORT_CHUNK + 2;
        const TileBinPlan pl = tile_bin_plan((int)(Pn > 0x7fffffff ? 0x7fffffff : Pn), W, H);
        s.counters = c.take<int32_t>(8);
        s.rects = c.take<uint2>(Pn);
        s.depth_a = c.take<uint64_t>(Pn);
        s.depth_b = c.take<uint64_t>(Pn);
        s.srects = c.take<uint4>(Pn);
        s.sort_ctl = c.take<uint32_t>(sort_ctl_words(nb));
        s.hist = c.take<uint32_t>(256 * nb);
        s.digit_tot = c.take<uint32_t>(256);
        s.tile_chunk = c.take<uint32_t>((size_t)pl.n_chunks * pl.T);
        s.tile_total = c.take<uint32_t>(pl.T);
        s.tile_start = c.take<uint32_t>((size_t)pl.T + 1);
        s.tile_stripes = c.take<uint32_t>((size_t)pl.n_bands * 32);
        s.tile_weights = c.take<uint32_t>(TB_MAX_WBLOCKS);
        s.tile_bounds = c.take<uint32_t>((size_t)pl.n_chunks + 1);
        s.acc = c.take<float>(Pn * ACC_FLOATS);
        s.band_list = c.take<uint32_t>(Pn);
        s.band_ctl = c.take<uint32_t>(BAND_CTL_WORDS);
        if (bytes) *bytes = c.used();
        return s;
    }
};

// ---- error plumbing ------------------------------------------------------
void set_error(const char* fmt, ...);
int  check_stage(const char* stage, bool debug, cudaStream_t stream);

#define GSR_CUDA_TRY(expr)                                                        \
    do {                                                                          \
        cudaError_t _e = (expr);                                                  \
        if (_e != cudaSuccess) {                                                  \
            gsr::set_error("%s failed: %s", #expr, cudaGetErrorString(_e));       \
            return GSR_ERR_CUDA;                                                  \
        }                                                                         \
    } while (0)

// ---- measurement hooks (api.cu) ---------------------------------------------
extern unsigned long long g_launches;
#ifndef GSR_HOST_EMUL
#define GSR_LAUNCH(kernel, grid, block, smem, st, ...)              \
    do {                                                            \
        kernel<<<(grid), (block), (smem), (st)>>>(__VA_ARGS__);     \
        ++gsr::g_launches;                                          \
    } while (0)
#else   // tests/host_emul: the same kernel functions run as fibers on the CPU (cuda_emul/cuda_runtime.h)
#define GSR_LAUNCH(kernel, grid, block, smem, st, ...)                          \
    do {                                                                        \
        emu::launch((grid), (block), [&]() { kernel(__VA_ARGS__); });           \
        ++gsr::g_launches;                                                      \
    } while (0)
#endif

void profile_begin(int stage, cudaStream_t st);
void profile_end(int stage, cudaStream_t st);
struct StageTimer {     // brackets a launcher's work with events when profiling is enabled
    int stage; cudaStream_t st;
    StageTimer(int s, cudaStream_t t) : stage(s), st(t) { profile_begin(stage, st); }
    ~StageTimer() { profile_end(stage, st); }
};

// ---- stage launchers (defined in the per-stage .cu files) ------------------
int launch_preprocess(const GsrParams& p, GeomState g, Scratch s, int32_t* radii, cudaStream_t st);
int launch_mark_visible(int P, const float* means, const float* view, uint8_t* present, cudaStream_t st);
int launch_depth_order_and_count(const GsrParams& p, GeomState g, ImgState im, Scratch s, int32_t* num_rendered_host,
                                 void* num_rendered_event, cudaStream_t st);
int launch_tile_lists(const GsrParams& p, GeomState g, ImgState im, BinState b, Scratch s, int64_t R, cudaStream_t st);
// acc_clear: the backward's accumulator (Scratch::acc) to zero along the way, or nullptr
int launch_render_forward(const GsrParams& p, GeomState g, ImgState im, BinState b,
                          float* out_color, float* out_invdepth, float* acc_clear, cudaStream_t st);
int launch_render_backward(const GsrParams& p, GeomState g, ImgState im, BinState b, Scratch s,
                           const float* dL_dpix, const float* dL_dinvdepth, int row_lo, int row_hi, bool zero_acc,
                           cudaStream_t st);
// band < 0: all Gaussians, dense outputs.  band >= 0 (multimem mode): the Gaussians of that band list.
int launch_preprocess_backward(const GsrParams& p, GeomState g, Scratch s, const int32_t* radii,
                               bool have_invdepth, const GsrBackwardIO& io, int band, cudaStream_t st);
// Groups the visible Gaussians by the tile-row band of their last tile row (band_hi[k] = exclusive
// upper row of band k) into s.band_list / s.band_ctl; zeroes dL_dmeans2D of the culled ones.
int launch_band_lists(const GsrParams& p, GeomState g, Scratch s, const int32_t* radii, float* dL_dmeans2D,
                      int bands, const int* band_hi, cudaStream_t st);
int launch_count_work(const GsrParams& p, GeomState g, ImgState im, BinState b, unsigned long long* counts,
                      cudaStream_t st);
int launch_rebuild_keys(const GsrParams& p, GeomState g, ImgState im, BinState b, int64_t R,
                        uint64_t* keys_out, cudaStream_t st);

}  // namespace gsr
