// api.cu — the C-ABI of libgsrast.so (declared in include/gsrast.h): argument validation,
// workspace carving, stage orchestration, error strings.  No torch, no allocation.
#include <stdarg.h>
#include <stdio.h>
#include <string.h>
#include <mutex>
#include <vector>
#include "common.cuh"

namespace gsr {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

// After each stage: always pick up launch-configuration errors; in debug mode also
// synchronise and name the stage (role of CHECK_CUDA, auxiliary.h:193-200).
int check_stage(const char* stage, bool debug, cudaStream_t stream) {
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess && debug) e = cudaStreamSynchronize(stream);
    if (e != cudaSuccess) {
        set_error("[CUDA ERROR] in stage '%s': %s", stage, cudaGetErrorString(e));
        return GSR_ERR_CUDA;
    }
    return GSR_OK;
}

// ---- measurement hooks ----------------------------------------------------
unsigned long long g_launches = 0;
static bool g_profile = false;
struct ProfSpan { int stage; cudaEvent_t a, b; };
static std::vector<ProfSpan> g_spans;       // recorded, not yet collected
static std::vector<cudaEvent_t> g_free;     // event pool
static cudaEvent_t g_open[GSR_STAGE_COUNT];

static std::mutex g_prof_mutex;           // the measurement hooks may be reached from several host threads
static cudaEvent_t take_event() {
    cudaEvent_t e = nullptr;
    if (!g_free.empty()) { e = g_free.back(); g_free.pop_back(); }
    else cudaEventCreate(&e);
    return e;
}
void profile_begin(int stage, cudaStream_t st) {
    if (!g_profile) return;
    std::lock_guard<std::mutex> lock(g_prof_mutex);
    g_open[stage] = take_event();
    cudaEventRecord(g_open[stage], st);
}
void profile_end(int stage, cudaStream_t st) {
    if (!g_profile) return;
    std::lock_guard<std::mutex> lock(g_prof_mutex);
    if (!g_open[stage]) return;
    cudaEvent_t b = take_event();
    cudaEventRecord(b, st);
    g_spans.push_back({stage, g_open[stage], b});
    g_open[stage] = nullptr;
}

static int validate(const GsrParams* p, const GsrBuffers* buf) {
    if (!p || !buf) { set_error("null argument"); return GSR_ERR_INVALID_ARG; }
    if (p->P <= 0) { set_error("P must be > 0 (P == 0 is handled by the caller)"); return GSR_ERR_INVALID_ARG; }
    if (p->W <= 0 || p->H <= 0) { set_error("image size must be positive"); return GSR_ERR_INVALID_ARG; }
    if ((p->W + TILE_X - 1) / TILE_X > TB_MAX_GX || (p->H + TILE_Y - 1) / TILE_Y > 65535) {
        set_error("image too large: tile grid exceeds %d columns or 65535 rows", TB_MAX_GX); return GSR_ERR_INVALID_ARG;
    }
    if (!p->means3D || !p->opacities || !p->background || !p->viewmatrix || !p->projmatrix || !p->campos) {
        set_error("means3D, opacities, background, viewmatrix, projmatrix, campos are required");
        return GSR_ERR_INVALID_ARG;
    }
    if ((p->shs == nullptr) == (p->colors_precomp == nullptr)) {
        set_error("Please provide exactly one of either SHs or precomputed colors!");
        return GSR_ERR_INVALID_ARG;
    }
    const bool have_sr = p->scales != nullptr && p->rotations != nullptr;
    if (have_sr == (p->cov3D_precomp != nullptr) || ((p->scales != nullptr) != (p->rotations != nullptr))) {
        set_error("Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!");
        return GSR_ERR_INVALID_ARG;
    }
    if (p->rotations && (reinterpret_cast<uintptr_t>(p->rotations) & 15) != 0) {
        set_error("rotations must be 16-byte aligned (read as float4)");
        return GSR_ERR_INVALID_ARG;
    }
    if (p->sh_dc && (!p->shs || p->M < 2)) {
        set_error("split SH storage (sh_dc) needs shs = coefficients 1..M-1 and M >= 2");
        return GSR_ERR_INVALID_ARG;
    }
    if (p->shs) {
        if (p->D < 0 || p->D > 3) { set_error("sh_degree must be in 0..3"); return GSR_ERR_INVALID_ARG; }
        if (p->M < (p->D + 1) * (p->D + 1) || p->M > 16) {
            set_error("shs must hold between (sh_degree+1)^2 and 16 coefficients per Gaussian (got M=%d, D=%d)", p->M, p->D);
            return GSR_ERR_INVALID_ARG;
        }
    }
    if (!buf->geom || buf->geom_bytes < gsr_geom_bytes(p->P)) {
        set_error("geom buffer too small: %zu < %zu", buf->geom_bytes, gsr_geom_bytes(p->P));
        return GSR_ERR_CAPACITY;
    }
    if (!buf->img || buf->img_bytes < gsr_img_bytes(p->W, p->H)) {
        set_error("img buffer too small: %zu < %zu", buf->img_bytes, gsr_img_bytes(p->W, p->H));
        return GSR_ERR_CAPACITY;
    }
    return GSR_OK;
}

static int validate_R(const GsrParams* p, const GsrBuffers* buf, int64_t R) {
    if (R < 0 || R > 0x7fffffffLL) { set_error("num_rendered out of int32 range"); return GSR_ERR_OVERFLOW; }
    if (!buf->bin || buf->bin_bytes < gsr_bin_bytes(R)) {
        set_error("bin buffer too small: %zu < %zu", buf->bin_bytes, gsr_bin_bytes(R));
        return GSR_ERR_CAPACITY;
    }
    if (!buf->scratch || buf->scratch_bytes < gsr_scratch_bytes(p->P, p->W, p->H, R)) {
        set_error("scratch buffer too small: %zu < %zu", buf->scratch_bytes,
                  gsr_scratch_bytes(p->P, p->W, p->H, R));
        return GSR_ERR_CAPACITY;
    }
    return GSR_OK;
}

// the 64-bit instance count as the two words stage 1 copies to the host
static int read_num_rendered(const int32_t* host, int64_t* R) {
    if (host[1] != 0 || host[0] < 0) { set_error("num_rendered exceeds int32"); return GSR_ERR_OVERFLOW; }
    *R = host[0];
    return GSR_OK;
}

// Host-side resources for the banded backward: one non-blocking side stream and a few events per
// device, created on first use and kept for the life of the process (no device memory).
struct SideStream {
    cudaStream_t stream = nullptr;
    cudaEvent_t band_done[MAX_BANDS] = {};
    cudaEvent_t joined = nullptr;
};
static SideStream g_side[64];
// The banded backward re-records the per-device events above on every call: two host threads (or two caller
// streams) driving the same device are serialised through this section, so that one call's waits can never
// pick up the other's records.
static std::mutex g_side_mutex;
static SideStream* side_stream() {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    SideStream* s = &g_side[dev];
    if (s->stream == nullptr) {
        if (cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking) != cudaSuccess) return nullptr;
        for (int i = 0; i < MAX_BANDS; i++) cudaEventCreateWithFlags(&s->band_done[i], cudaEventDisableTiming);
        cudaEventCreateWithFlags(&s->joined, cudaEventDisableTiming);
    }
    return s;
}

static inline size_t tiles_of(int W, int H) {
    return (size_t)((W + TILE_X - 1) / TILE_X) * (size_t)((H + TILE_Y - 1) / TILE_Y);
}

}  // namespace gsr

using namespace gsr;

extern "C" {

int gsr_version(void) { return GSR_VERSION; }
const char* gsr_last_error(void) { return g_err; }

size_t gsr_geom_bytes(int32_t P) {
    size_t b; GeomState::carve(nullptr, (size_t)(P > 0 ? P : 1), &b); return b;
}
size_t gsr_img_bytes(int32_t W, int32_t H) {
    size_t b; ImgState::carve(nullptr, (size_t)W * H, tiles_of(W, H), &b); return b;
}
size_t gsr_bin_bytes(int64_t R) {
    size_t b; BinState::carve(nullptr, (size_t)(R > 0 ? R : 0), &b); return b;
}
size_t gsr_scratch_bytes(int32_t P, int32_t W, int32_t H, int64_t R) {
    (void)R;    // no R-sized scratch: the (Gaussian, tile) instances are never materialised outside `bin`
    size_t b; Scratch::carve(nullptr, (size_t)(P > 0 ? P : 1), W, H, &b); return b;
}

int gsr_profile_enable(int on) {
    std::lock_guard<std::mutex> lock(g_prof_mutex);
    g_profile = on != 0;
    if (!g_profile) {
        for (auto& sp : g_spans) { g_free.push_back(sp.a); g_free.push_back(sp.b); }
        g_spans.clear();
    }
    return GSR_OK;
}

int gsr_profile_collect(double* ms_sum, int64_t* calls) {
    std::lock_guard<std::mutex> lock(g_prof_mutex);
    for (auto& sp : g_spans) {
        GSR_CUDA_TRY(cudaEventSynchronize(sp.b));
        float ms = 0.f;
        GSR_CUDA_TRY(cudaEventElapsedTime(&ms, sp.a, sp.b));
        if (ms_sum) ms_sum[sp.stage] += ms;
        if (calls) calls[sp.stage] += 1;
        g_free.push_back(sp.a); g_free.push_back(sp.b);
    }
    g_spans.clear();
    return GSR_OK;
}

const char* gsr_stage_name(int stage) {
    static const char* names[GSR_STAGE_COUNT] = {"preprocess", "depth_sort", "tile_count", "tile_fill",
                                                 "render_fwd", "render_bwd", "preprocess_bwd", "exchange"};
    return stage >= 0 && stage < GSR_STAGE_COUNT ? names[stage] : "?";
}

uint64_t gsr_launch_count(void) { return g_launches; }

int32_t gsr_flat_grad_floats(int32_t M) { return 3 + 3 * M + 1 + 3 + 4; }

int gsr_forward_stage1(const GsrParams* p, const GsrBuffers* buf, const GsrForwardOut* out, void* stream) {
    int rc = validate(p, buf);
    if (rc) return rc;
    if (!out || !out->radii) { set_error("radii output is required"); return GSR_ERR_INVALID_ARG; }
    if (!buf->scratch || buf->scratch_bytes < gsr_scratch_bytes(p->P, p->W, p->H, 0)) {
        set_error("scratch buffer too small for stage 1: %zu < %zu", buf->scratch_bytes,
                  gsr_scratch_bytes(p->P, p->W, p->H, 0));
        return GSR_ERR_CAPACITY;
    }
    cudaStream_t st = (cudaStream_t)stream;
    GeomState g = GeomState::carve(buf->geom, p->P, nullptr);
    ImgState im = ImgState::carve(buf->img, (size_t)p->W * p->H, tiles_of(p->W, p->H), nullptr);
    Scratch s = Scratch::carve(buf->scratch, p->P, p->W, p->H, nullptr);
    rc = launch_preprocess(*p, g, s, out->radii, st);
    if (rc) return rc;
    return launch_depth_order_and_count(*p, g, im, s, buf->num_rendered_host, buf->num_rendered_event, st);
}

int gsr_forward_stage2(const GsrParams* p, const GsrBuffers* buf, const GsrForwardOut* out, int64_t R,
                       void* stream) {
    int rc = validate(p, buf);
    if (rc) return rc;
    if (!out || !out->out_color || !out->out_invdepth) {
        set_error("out_color and out_invdepth are required"); return GSR_ERR_INVALID_ARG;
    }
    rc = validate_R(p, buf, R);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    GeomState g = GeomState::carve(buf->geom, p->P, nullptr);
    ImgState im = ImgState::carve(buf->img, (size_t)p->W * p->H, tiles_of(p->W, p->H), nullptr);
    BinState b = BinState::carve(buf->bin, (size_t)R, nullptr);
    Scratch s = Scratch::carve(buf->scratch, p->P, p->W, p->H, nullptr);
    rc = launch_tile_lists(*p, g, im, b, s, R, st);
    if (rc) return rc;
    return launch_render_forward(*p, g, im, b, out->out_color, out->out_invdepth,
                                 (buf->accumulator & GSR_ACC_CLEAR_IN_FORWARD) ? s.acc : nullptr, st);
}

int gsr_forward(const GsrParams* p, const GsrBuffers* buf, const GsrForwardOut* out, int64_t* num_rendered,
                void* stream) {
    if (!buf || !buf->num_rendered_host) { set_error("num_rendered_host (pinned) is required"); return GSR_ERR_INVALID_ARG; }
    int rc = gsr_forward_stage1(p, buf, out, stream);
    if (rc) return rc;
    GSR_CUDA_TRY(cudaStreamSynchronize((cudaStream_t)stream));
    int64_t R = 0;
    rc = read_num_rendered(buf->num_rendered_host, &R);
    if (rc) return rc;
    if (num_rendered) *num_rendered = R;
    return gsr_forward_stage2(p, buf, out, R, stream);
}

int gsr_backward(const GsrParams* p, const GsrBuffers* buf, int64_t R, const int32_t* radii,
                 const GsrBackwardIO* io, void* stream) {
    int rc = validate(p, buf);
    if (rc) return rc;
    if (!io || !io->dL_dout_color || !radii) { set_error("dL_dout_color and radii are required"); return GSR_ERR_INVALID_ARG; }
    rc = validate_R(p, buf, R);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    GeomState g = GeomState::carve(buf->geom, p->P, nullptr);
    ImgState im = ImgState::carve(buf->img, (size_t)p->W * p->H, tiles_of(p->W, p->H), nullptr);
    BinState b = BinState::carve(buf->bin, (size_t)R, nullptr);
    Scratch s = Scratch::carve(buf->scratch, p->P, p->W, p->H, nullptr);
    const int gy = (p->H + TILE_Y - 1) / TILE_Y;
    const bool have_invdepth = io->dL_dout_invdepth != nullptr;
    const bool fact = io->xchg_records != nullptr;
    const bool acc_is_clear = (buf->accumulator & GSR_ACC_IS_CLEAR) != 0;   // zeroed by this view's forward
    if (p->sh_dc && !fact && ((io->dL_dsh == nullptr) != (io->dL_dsh_dc == nullptr))) {
        set_error("split SH storage: pass both dL_dsh (coefficients 1..M-1) and dL_dsh_dc, or neither");
        return GSR_ERR_INVALID_ARG;
    }
    if (fact) {
        const auto misaligned = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) != 0; };
        if (!io->xchg_rgb || !io->xchg_campos || !p->shs || !p->scales || !p->rotations || p->colors_precomp ||
            p->cov3D_precomp || misaligned(io->xchg_records) || misaligned(io->xchg_rgb) || misaligned(io->xchg_campos)) {
            set_error("factored exchange needs SH + scale/rotation inputs and three 16-byte aligned xchg_* buffers");
            return GSR_ERR_INVALID_ARG;
        }
    }
    // the factored exchange moves ~4x fewer NVLink bytes: no band pipelining needed
    int bands = (io->multimem_reduce && !fact) ? (io->reserved1 > 0 ? io->reserved1 : 1) : 1;
    if (bands > gy) bands = gy;
    if (bands > MAX_BANDS) bands = MAX_BANDS;
    SideStream* side = bands > 1 ? side_stream() : nullptr;
    // list mode needs whole-float4, 16 B aligned SH rows (same condition as the bulk-copy path)
    const bool list_ok = p->sh_dc != nullptr ? false : p->shs == nullptr || io->dL_dsh == nullptr ||
                         ((p->M & 3) == 0 && (reinterpret_cast<uintptr_t>(p->shs) & 15) == 0 &&
                          (reinterpret_cast<uintptr_t>(io->dL_dsh) & 15) == 0);
    if (bands <= 1 || side == nullptr || !list_ok) {
        rc = launch_render_backward(*p, g, im, b, s, io->dL_dout_color, io->dL_dout_invdepth, 0, gy, !acc_is_clear, st);
        if (rc) return rc;
        return launch_preprocess_backward(*p, g, s, radii, have_invdepth, *io, -1, st);
    }
    // Multicast-reduce mode with bands: run the blend backward band by band (tile rows, top to
    // bottom) and hand every finished band's Gaussians (those whose last tile row lies in it) to the
    // per-Gaussian kernel on a side stream, so that the gradient traffic of band k crosses NVLink
    // while band k+1 is blended.
    std::lock_guard<std::mutex> side_lock(g_side_mutex);
    int band_hi[MAX_BANDS];
    for (int k = 0; k < bands; k++) band_hi[k] = (int)((long long)gy * (k + 1) / bands);
    GSR_CUDA_TRY(cudaEventRecord(side->joined, st));                       // order the side stream after the caller's work
    GSR_CUDA_TRY(cudaStreamWaitEvent(side->stream, side->joined, 0));
    rc = launch_band_lists(*p, g, s, radii, io->dL_dmeans2D, bands, band_hi, side->stream);
    if (rc) return rc;
    for (int k = 0; k < bands; k++) {
        const int lo = k == 0 ? 0 : band_hi[k - 1];
        rc = launch_render_backward(*p, g, im, b, s, io->dL_dout_color, io->dL_dout_invdepth, lo, band_hi[k],
                                    k == 0 && !acc_is_clear, st);
        if (rc) return rc;
        GSR_CUDA_TRY(cudaEventRecord(side->band_done[k], st));
        GSR_CUDA_TRY(cudaStreamWaitEvent(side->stream, side->band_done[k], 0));
        rc = launch_preprocess_backward(*p, g, s, radii, have_invdepth, *io, k, side->stream);
        if (rc) return rc;
    }
    GSR_CUDA_TRY(cudaEventRecord(side->joined, side->stream));
    GSR_CUDA_TRY(cudaStreamWaitEvent(st, side->joined, 0));
    return GSR_OK;
}

int gsr_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                     uint8_t* present, void* stream) {
    (void)projmatrix;   // read but unused by the reference's test (auxiliary.h:164-190)
    if (P < 0 || (P > 0 && (!means3D || !viewmatrix || !present))) { set_error("bad arguments"); return GSR_ERR_INVALID_ARG; }
    if (P == 0) return GSR_OK;
    return launch_mark_visible(P, means3D, viewmatrix, present, (cudaStream_t)stream);
}

int gsr_debug_sorted_keys(const GsrParams* p, const GsrBuffers* buf, int64_t R, uint64_t* keys_out, void* stream) {
    int rc = validate(p, buf);
    if (rc) return rc;
    GeomState g = GeomState::carve(buf->geom, p->P, nullptr);
    ImgState im = ImgState::carve(buf->img, (size_t)p->W * p->H, tiles_of(p->W, p->H), nullptr);
    BinState b = BinState::carve(buf->bin, (size_t)R, nullptr);
    return launch_rebuild_keys(*p, g, im, b, R, keys_out, (cudaStream_t)stream);
}

int gsr_debug_count_work(const GsrParams* p, const GsrBuffers* buf, int64_t R, uint64_t* counts, void* stream) {
    int rc = validate(p, buf);
    if (rc) return rc;
    if (!counts) { set_error("counts is required"); return GSR_ERR_INVALID_ARG; }
    GeomState g = GeomState::carve(buf->geom, p->P, nullptr);
    ImgState im = ImgState::carve(buf->img, (size_t)p->W * p->H, tiles_of(p->W, p->H), nullptr);
    BinState b = BinState::carve(buf->bin, (size_t)R, nullptr);
    return launch_count_work(*p, g, im, b, reinterpret_cast<unsigned long long*>(counts), (cudaStream_t)stream);
}

int gsr_debug_state_ptrs(const GsrParams* p, const GsrBuffers* buf, int64_t R, const uint32_t** point_list,
                         const uint32_t** ranges, const float** final_T, const uint32_t** n_contrib,
                         const float** records) {
    int rc = validate(p, buf);
    if (rc) return rc;
    GeomState g = GeomState::carve(buf->geom, p->P, nullptr);
    ImgState im = ImgState::carve(buf->img, (size_t)p->W * p->H, tiles_of(p->W, p->H), nullptr);
    BinState b = BinState::carve(buf->bin, (size_t)R, nullptr);
    if (point_list) *point_list = b.point_list;
    if (ranges) *ranges = reinterpret_cast<const uint32_t*>(im.ranges);
    if (final_T) *final_T = im.final_T;
    if (n_contrib) *n_contrib = im.n_contrib;
    if (records) *records = reinterpret_cast<const float*>(g.rec);
    return GSR_OK;
}

}  // extern "C"
