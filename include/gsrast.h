/*
 * gsrast.h — C-ABI of the B200-native differentiable Gaussian rasterizer.
 *
 * This is the drop-in boundary for the reference's native layer
 * (submodules/diff-gaussian-rasterization, "DGR"):
 *   gsr_forward_*      replaces RasterizeGaussiansCUDA          (DGR/rasterize_points.cu:45-146,
 *                      CudaRasterizer::Rasterizer::forward, DGR/cuda_rasterizer/rasterizer_impl.cu:227-370)
 *   gsr_backward       replaces RasterizeGaussiansBackwardCUDA  (DGR/rasterize_points.cu:149-248,
 *                      Rasterizer::backward, rasterizer_impl.cu:374-479)
 *   gsr_mark_visible   replaces markVisible                     (DGR/rasterize_points.cu:250-274,
 *                      rasterizer_impl.cu:169-181)
 *
 * Conventions
 *  - extern "C", plain pointers and sizes; no torch / C++ types.
 *  - The library never allocates or frees device memory that outlives a call:
 *    every output, state and scratch buffer is owned by the caller (Python:
 *    torch tensors).  Sizes come from the gsr_*_bytes queries.
 *  - All device pointers must be 16-byte aligned (torch allocations are 512 B).
 *  - Every entry point takes the CUDA stream to enqueue on (a cudaStream_t
 *    passed as void*; NULL = legacy default stream).  Nothing synchronises the
 *    host except where stated.
 *  - Return value: 0 on success, negative GSR_ERR_* otherwise; text through
 *    gsr_last_error() (thread-local).  Never exit()/abort().
 *  - Optional inputs follow the reference's "empty tensor == nullptr" rule
 *    (rasterize_points.cu:120-139): NULL shs => colors_precomp mode, NULL
 *    scales/rotations => cov3D_precomp mode, NULL dL_dout_invdepth => no depth
 *    gradient.
 */
#ifndef GSRAST_H_INCLUDED
#define GSRAST_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GSR_VERSION 100

#define GSR_OK                 0
#define GSR_ERR_INVALID_ARG   -1
#define GSR_ERR_CUDA          -2
#define GSR_ERR_CAPACITY      -3   /* a caller-provided buffer is too small */
#define GSR_ERR_OVERFLOW      -4   /* num_rendered does not fit in int32 (reference: int num_rendered) */
#define GSR_ERR_NO_DEVICE     -5

#define GSR_TILE_X 16               /* DGR/cuda_rasterizer/config.h:16-17 */
#define GSR_TILE_Y 16
#define GSR_CHANNELS 3              /* config.h:15 */

/* Per-call arguments shared by the forward stages and the backward pass.
 * Field meaning follows GaussianRasterizationSettings
 * (DGR/diff_gaussian_rasterization/__init__.py:151-180) and the argument list
 * of Rasterizer::forward (rasterizer_impl.cu:227-254). */
/* Fused activations.  The reference's model stores log-scales, unnormalised quaternions and
 * opacity logits and activates them with three torch kernels per step before the rasterizer sees them
 * (scene/gaussian_model.py:32-55,130-157: torch.exp, torch.nn.functional.normalize, torch.sigmoid; called from
 * gaussian_renderer/__init__.py:77,89-90), and autograd runs their three backward kernels afterwards.  With a bit
 * set, the corresponding input of GsrParams is the RAW parameter: the per-Gaussian kernels apply exp /
 * x / max(|x|, 1e-12) / 1 / (1 + exp(-x)) in registers, and the backward returns the gradient with respect to the
 * raw parameter (dL_dscales, dL_drotations, dL_dopacity and the exchange records alike). */
#define GSR_RAW_SCALES     1
#define GSR_RAW_ROTATIONS  2
#define GSR_RAW_OPACITIES  4

typedef struct GsrParams {
    int32_t P;                 /* number of Gaussians */
    int32_t D;                 /* active SH degree 0..3 */
    int32_t M;                 /* SH coefficients stored per Gaussian (sh.size(1)); 0 if no shs */
    int32_t W, H;              /* image size in pixels */
    float tanfovx, tanfovy;
    float scale_modifier;
    int32_t prefiltered;
    int32_t antialiasing;
    int32_t debug;             /* 1: synchronise + check after every stage and name the stage */
    int32_t raw_inputs;        /* bit mask GSR_RAW_*: the kernels apply the model's activations themselves (below) */

    /* device inputs (fp32) */
    const float* background;     /* 3 */
    const float* means3D;        /* P*3 */
    const float* shs;            /* P*M*3 or NULL */
    const float* colors_precomp; /* P*3 or NULL */
    const float* opacities;      /* P */
    const float* scales;         /* P*3 or NULL */
    const float* rotations;      /* P*4 (w,x,y,z) or NULL */
    const float* cov3D_precomp;  /* P*6 or NULL */
    const float* viewmatrix;     /* 16, column-major W2V   (auxiliary.h:75-95) */
    const float* projmatrix;     /* 16, column-major Proj*W2V */
    const float* campos;         /* 3 */
    /* Split SH storage (the reference's separate_sh call, gaussian_renderer/__init__.py:115-125: `dc=` and `shs=`
     * are separate tensors, as scene/gaussian_model.py keeps _features_dc / _features_rest).  When sh_dc != NULL,
     * coefficient 0 of every Gaussian is read from sh_dc [P][3] and `shs` holds coefficients 1..M-1 as
     * [P][M-1][3]; M still counts all coefficients (M >= 2).  No concatenated copy is ever made. */
    const float* sh_dc;
} GsrParams;

/* Caller-owned state and scratch.  geom/img/bin are kept alive by the caller
 * between forward and backward (the reference keeps geomBuffer/binningBuffer/
 * imgBuffer on the autograd ctx, __init__.py:93).  scratch is only used inside
 * a call and may be shared across calls on the same stream. */
typedef struct GsrBuffers {
    void*  geom;    size_t geom_bytes;     /* >= gsr_geom_bytes(P) */
    void*  img;     size_t img_bytes;      /* >= gsr_img_bytes(W,H) */
    void*  bin;     size_t bin_bytes;      /* >= gsr_bin_bytes(R) */
    void*  scratch; size_t scratch_bytes;  /* >= gsr_scratch_bytes(P,W,H,R) */
    int32_t* num_rendered_host;            /* pinned host int32[2]: the 64-bit instance count as {low word, high word};
                                            * it overflows the reference's int (rasterizer_impl.cu:312) when
                                            * [1] != 0 || [0] < 0 */
    void*    num_rendered_event;           /* optional cudaEvent_t recorded by stage 1 right after the copy into
                                            * num_rendered_host: waiting on it (instead of the stream) lets the host
                                            * size `bin` while the device is still ordering the Gaussians */
    int32_t  accumulator;                  /* GSR_ACC_* bits, 0 = the backward clears its accumulator itself */
    int32_t  reserved2;
} GsrBuffers;

/* GsrBuffers.accumulator.  The backward sums per-Gaussian gradients into 12*P floats inside `scratch`, which must
 * start from zero (the reference zero-fills its ten gradient tensors per call instead, rasterize_points.cu:186-204).
 * GSR_ACC_CLEAR_IN_FORWARD (gsr_forward_stage2 / gsr_forward): the blend forward, which is bound by instruction
 *   issue and leaves the memory system idle, also zeroes that accumulator.
 * GSR_ACC_IS_CLEAR (gsr_backward): the accumulator is still zero, skip the memset.  Only valid when the LAST call
 *   that used this scratch was a forward with GSR_ACC_CLEAR_IN_FORWARD and the same P, W, H (any other forward may
 *   lay its own buffers over the same bytes; a backward leaves the accumulator dirty). */
#define GSR_ACC_CLEAR_IN_FORWARD 1
#define GSR_ACC_IS_CLEAR         2

typedef struct GsrForwardOut {
    float*   out_color;     /* 3*H*W */
    float*   out_invdepth;  /* H*W   */
    int32_t* radii;         /* P     */
} GsrForwardOut;

typedef struct GsrBackwardIO {
    /* upstream gradients */
    const float* dL_dout_color;     /* 3*H*W */
    const float* dL_dout_invdepth;  /* H*W or NULL */
    /* outputs; every one is fully written (zeros for culled Gaussians), so the
     * caller may pass uninitialised memory.  NULL = not wanted. */
    float* dL_dmeans3D;    /* P*3 */
    float* dL_dmeans2D;    /* P*3 (z stays 0; x,y scaled by 0.5*W, 0.5*H as backward.cu:529-530) */
    float* dL_dopacity;    /* P   */
    float* dL_dcolors;     /* P*3   gradient of colors_precomp; NULL in SH mode */
    float* dL_dcov3D;      /* P*6   gradient of cov3D_precomp;  NULL in scale/rot mode */
    float* dL_dsh;         /* P*M*3 or NULL */
    float* dL_dscales;     /* P*3 or NULL */
    float* dL_drotations;  /* P*4 or NULL */
    /* View-parallel fused reduction.  0: plain stores.  1: every output above except
     * dL_dmeans2D is an NVLink MULTICAST address (NVLS) of a buffer replicated on all ranks; the
     * kernel adds its non-zero values with multimem.red.add.f32, so after all ranks have run, every
     * replica holds the sum over ranks.  The caller zeroes its replica and synchronises the ranks
     * before and after (no NCCL all-reduce pass).  Pointers must be 16 B aligned. */
    int32_t multimem_reduce;
    int32_t reserved1;     /* multimem mode: number of tile-row bands pipelined against the blend backward (0 or 1 = no pipelining) */
    /* Factored exchange (view-parallel, SH + scale/rotation inputs).  dL_dsh is the outer
     * product  basis(mean - campos) x dL_dRGB  (backward.cu:12-146), so a rank only has to publish the 3
     * clamp-masked colour gradients of a Gaussian instead of its 3*M SH gradients; every receiver rebuilds
     * the basis of each sender's camera from its own copy of the means (gsr_exchange_expand).  When
     * xchg_records != NULL the kernel writes ONLY these three buffers and dL_dmeans2D (the five per-input
     * gradient pointers above are ignored); rows of culled Gaussians are not touched, so the caller
     * zeroes the buffers beforehand.  With multimem_reduce = 1 the three pointers are multicast addresses. */
    float* xchg_records;   /* P*12: {dL_dmeans3D xyz, dL_dopacity, dL_dscales xyz, 0, dL_drotations wxyz}, summed over ranks */
    float* xchg_rgb;       /* this rank's slot of [world][P][4]: {dL_dR, dL_dG, dL_dB, 0} */
    float* xchg_campos;    /* this rank's slot of [world][4]:    {campos xyz, 1}; stored, not added */
    float* dL_dsh_dc;      /* split SH storage (GsrParams.sh_dc): P*3; dL_dsh then is P*(M-1)*3 */
} GsrBackwardIO;

int         gsr_version(void);
const char* gsr_last_error(void);

/* Workspace sizing (bytes).  R = number of (Gaussian, tile) instances. */
size_t gsr_geom_bytes(int32_t P);
size_t gsr_img_bytes(int32_t W, int32_t H);
size_t gsr_bin_bytes(int64_t R);
size_t gsr_scratch_bytes(int32_t P, int32_t W, int32_t H, int64_t R);

/* Forward, stage 1: preprocess (projection, EWA + antialiasing, SH->RGB, tile
 * rects, num_rendered), then stable depth ordering, per-tile counts and tile
 * ranges.  The async copy of num_rendered into buf->num_rendered_host is
 * enqueued right after the preprocess kernel (and buf->num_rendered_event
 * recorded, when given); the caller waits for that event or for the stream
 * before reading it and before stage 2.  Needs buf->img (the ranges live
 * there).  Replaces rasterizer_impl.cu:279-313 and 342-350. */
int gsr_forward_stage1(const GsrParams* p, const GsrBuffers* buf,
                       const GsrForwardOut* out, void* stream);

/* Forward, stage 2: fills the per-tile, depth-ordered Gaussian lists
 * (point_list == the reference's sorted value list) and blends.  R must be
 * the value produced by stage 1; buf->bin must hold gsr_bin_bytes(R).
 * Replaces rasterizer_impl.cu:315-340 and 352-368. */
int gsr_forward_stage2(const GsrParams* p, const GsrBuffers* buf,
                       const GsrForwardOut* out, int64_t R, void* stream);

/* Convenience: stage 1, host wait on `stream`, stage 2.  *num_rendered gets R.
 * If buf->bin is too small returns GSR_ERR_CAPACITY with *num_rendered set;
 * the caller regrows bin and calls gsr_forward_stage2. */
int gsr_forward(const GsrParams* p, const GsrBuffers* buf,
                const GsrForwardOut* out, int64_t* num_rendered, void* stream);

/* Backward of the whole path. R as returned by forward. */
int gsr_backward(const GsrParams* p, const GsrBuffers* buf, int64_t R,
                 const int32_t* radii, const GsrBackwardIO* io, void* stream);

/* Second half of the factored exchange: after every rank's gsr_backward has landed (rank barrier), turn this
 * rank's replica of the exchange buffers into the dense gradient tensors, summed over ranks:
 *   dL_dsh[g] = sum_r basis_D(means3D[g] - campos[r]) x rgb[r][g]   (fixed order r = 0..world-1, deterministic),
 * the other four straight from the records.  Uses p->P, D, M, means3D.  Every output row is written.
 * Rows of `records` / `rgb` that held something are reset to zero on the way, so the buffers are clean for the
 * next step without a memset (the camera slots are plain stores and need no reset). */
int gsr_exchange_expand(const GsrParams* p, int32_t world, float* records /* P*12 */,
                        float* rgb /* world*P*4 */, const float* campos /* world*4 */,
                        float* dL_dmeans3D, float* dL_dopacity, float* dL_dsh, float* dL_dscales,
                        float* dL_drotations, float* dL_dsh_dc /* split SH storage: P*3, dL_dsh is P*(M-1)*3; else NULL */,
                        void* stream);

/* present[i] = (W2V * p_i).z > 0.2   (rasterizer_impl.cu:56-73). present is P bytes (bool). */
int gsr_mark_visible(int32_t P, const float* means3D, const float* viewmatrix,
                     const float* projmatrix, uint8_t* present, void* stream);

/* Fused Adam step over one parameter tensor of N Gaussians x M floats (role of the
 * `SparseGaussianAdam.step(visible, N)` that train.py:240-247 calls and the fork does not ship).
 * torch.optim.Adam's update (bias_correction1 = 1 - beta1^t, bias_correction2_sqrt = sqrt(1 - beta2^t)),
 * applied only where visible[g] != 0 (visible == NULL: everywhere).  In place, one pass. */
int gsr_adam_update(float* param, const float* grad, float* exp_avg, float* exp_avg_sq,
                    const uint8_t* visible, double lr, double beta1, double beta2, double eps,
                    double bias_correction1, double bias_correction2_sqrt, int64_t N, int32_t M, void* stream);

/* Fused SSIM: replaces `fusedssim` / `fusedssim_backward` of the reference's
 * fused-ssim submodule (submodules/fused-ssim/ssim.h:7-26, ssim.cu:367-444), the D-SSIM term of the
 * training loss (train.py:155-162).  Images are [B, CH, H, W] fp32, contiguous; 11x11 Gaussian
 * window (sigma 1.5), zero padding.  crop = 0 is the reference's padding="same"; crop = 5 its
 * padding="valid" (fused_ssim/__init__.py:13-14,29-31): the mean covers, and a scalar upstream
 * gradient is spread over, [crop, H-crop) x [crop, W-crop) only. */
typedef struct GsrSsimParams {
    const float* img1;   /* the rendered image (receives the gradient) */
    const float* img2;   /* ground truth */
    int32_t B, CH, H, W;
    float C1, C2;        /* 0.01^2, 0.03^2 in the reference (fused_ssim/__init__.py:36-37) */
    int32_t crop;
} GsrSsimParams;

/* floats of scratch `partials` that gsr_ssim_forward needs when mean_out != NULL */
int64_t gsr_ssim_partials(int32_t B, int32_t CH, int32_t H, int32_t W);
/* ssim_map (nullable): the [B,CH,H,W] map.  dm_*: the three partial-derivative maps the backward
 * consumes (ssim.cu:268-277) — all three or none (none = the reference's train=False).
 * mean_out (nullable, device scalar): mean of the (cropped) map, reduced in the same launch in a
 * fixed order; needs `partials`. */
int gsr_ssim_forward(const GsrSsimParams* p, float* ssim_map, float* dm_dmu1, float* dm_dsigma1_sq,
                     float* dm_dsigma12, float* mean_out, float* partials, void* stream);
/* Allocation-free variant of gsr_ssim_forward (mean_out required).  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA-graph capture, tight serving loops,
 * harnesses that forbid allocation).  The per-tile partial sums go to a static device array inside the library
 * (GSR_SSIM_NOALLOC_MAX_TILES floats = 256 KB, held by every GPU that loads the library until the process exits);
 * the call makes no cudaMalloc / cudaFree / allocating library call.  Inputs and outputs are still the caller's.
 * Limit: B*CH*ceil(H/32)*ceil(W/32) <= GSR_SSIM_NOALLOC_MAX_TILES (e.g. 3 x 2160 x 3840, or 5 x 5 x 1080 x 1920),
 * larger sizes return GSR_ERR_CAPACITY.  All calls on a device share the array: calls that overlap on the GPU
 * (different streams without ordering) race; calls ordered on one stream are safe, and nothing needs clearing
 * between them (every slot read is rewritten by the same call). */
#define GSR_SSIM_NOALLOC_MAX_TILES 65536
int gsr_ssim_forward_noalloc(const GsrSsimParams* p, float* ssim_map, float* dm_dmu1, float* dm_dsigma1_sq,
                             float* dm_dsigma12, float* mean_out, void* stream);
/* dL_dimg1 [B,CH,H,W] from either dL_dmap (full-size map gradient, ssim.cu:283-365) or, when
 * dL_dmap == NULL, the device scalar *dL_dmean = dL/d(mean of the cropped map). */
int gsr_ssim_backward(const GsrSsimParams* p, const float* dL_dmap, const float* dL_dmean,
                      const float* dm_dmu1, const float* dm_dsigma1_sq, const float* dm_dsigma12,
                      float* dL_dimg1, void* stream);

/* ---------------------------------------------------------------------------------------------
 * The training loss in one forward and one backward launch.  Replaces, for the image the rasterizer just produced,
 * l1_loss (utils/loss_utils.py:40), fused_ssim (train.py:155-159), the weighted sum of train.py:162 and the
 * inverse-depth regulariser of train.py:165-174:
 *   loss = (1 - lambda_dssim) * mean|image - gt| + lambda_dssim * (1 - mean SSIM(image, gt))
 *          [+ depth_weight * mean|(invdepth - mono_invdepth) * depth_mask|     when invdepth != NULL]
 * Forward: the SSIM tiles also sum |image - gt| from the pixels they stage; the inverse-depth term rides on one
 * extra plane of the same grid; out4 = {loss, L1, mean SSIM, depth L1} (device), deterministic.  Backward: dL_dimage
 * and dL_dinvdepth (what the rasterizer's backward consumes) from the device scalar *dL_dloss.  `partials`:
 * gsr_loss_partials(CH, H, W) floats of scratch. */
typedef struct GsrLossParams {
    const float* image;          /* CH*H*W  rendered image */
    const float* gt;             /* CH*H*W */
    const float* invdepth;       /* H*W rendered inverse depth, or NULL: no depth term */
    const float* mono_invdepth;  /* H*W */
    const float* depth_mask;     /* H*W or NULL (= ones) */
    int32_t CH, H, W;
    float lambda_dssim;          /* opt.lambda_dssim (arguments/__init__.py: 0.2) */
    float depth_weight;          /* depth_l1_weight(iteration) */
    float C1, C2;                /* 0.01^2, 0.03^2 */
    int32_t reserved;
} GsrLossParams;
int64_t gsr_loss_partials(int32_t CH, int32_t H, int32_t W);
int gsr_loss_forward(const GsrLossParams* p, float* dm_dmu1, float* dm_dsigma1_sq, float* dm_dsigma12,
                     float* out4, float* partials, void* stream);
int gsr_loss_backward(const GsrLossParams* p, const float* dL_dloss, const float* dm_dmu1,
                      const float* dm_dsigma1_sq, const float* dm_dsigma12, float* dL_dimage,
                      float* dL_dinvdepth, void* stream);

/* ---------------------------------------------------------------------------------------------
 * Adaptive density control on the device.  Replaces the tensor surgery of
 * GaussianModel.densify_and_prune / densify_and_clone / densify_and_split / prune_points /
 * cat_tensors_to_optimizer / _prune_optimizer (scene/gaussian_model.py:414-621), reset_opacity
 * (:323-327) and add_densification_stats (:630-633, with train.py:212-213) — in the reference ~150
 * torch kernels with a host synchronisation per boolean-mask index — by a plan/apply pair:
 *   gsr_densify_plan   one pass over the P Gaussians decides keep / clone / split / prune for every
 *                      row the reference would produce and scans the decisions into destination rows;
 *   gsr_densify_apply  ONE launch moves every parameter tensor and both Adam moments to their final
 *                      rows, computing the children's position and scale on the way.
 * Result layout = the reference's: [surviving originals][clones][first children][second children],
 * each in source order (cat at :470-484, prune at :549-550 and :611).
 *
 * `plan` is caller-owned device scratch of gsr_densify_plan_words(P) 32-bit words; `counts_host` is pinned host
 * memory for 4 int32 {A = surviving originals, B = clones, C = split sources whose children survive,
 * P_new = A + B + 2C}, written by an async copy on `stream`: the caller waits for the stream, allocates
 * the P_new-row outputs and calls gsr_densify_apply with the same plan and the counts. */
typedef struct GsrDensifyPlanArgs {
    int32_t P;
    int32_t use_screen_size;         /* `if max_screen_size:` (gaussian_model.py:600) */
    float grad_threshold;            /* max_grad */
    float dense_extent;              /* percent_dense * scene_extent, rounded to fp32 (:515,555) */
    float min_opacity;               /* :598 */
    float max_screen_size;           /* :601: compared with max_radii2D, which :508 has just zeroed */
    float world_extent_limit;        /* 0.1 * extent (:602) */
    int32_t reserved;
    const float* xyz_gradient_accum; /* P   (:589)  NULL with prune_mask */
    const float* denom;              /* P */
    const float* scaling;            /* P*3 raw (pre-exp) */
    const float* opacity;            /* P   raw (pre-sigmoid) */
    const uint8_t* prune_mask;       /* non-NULL: plain prune_points(mask) (:433-451): row i survives iff !mask[i];
                                        every other input is ignored */
} GsrDensifyPlanArgs;

#define GSR_DENSIFY_COPY     0   /* parameter or statistic: copied to every row derived from the source */
#define GSR_DENSIFY_MOMENT   1   /* Adam moment: copied for surviving originals, zero for new rows (:462-466) */
#define GSR_DENSIFY_XYZ      2   /* children: R(rotation) * (exp(scaling) * noise) + xyz   (:519-523) */
#define GSR_DENSIFY_SCALING  3   /* children: log(exp(scaling) / (0.8 * 2))                (:524) */
#define GSR_DENSIFY_MAX_TENSORS 24

typedef struct GsrDensifyTensor {
    const float* src;    /* P x width */
    float* dst;          /* P_new x width */
    int32_t width;       /* floats per Gaussian */
    int32_t kind;        /* GSR_DENSIFY_* */
} GsrDensifyTensor;

typedef struct GsrDensifyApplyArgs {
    int32_t P, A, B, C;              /* the counts gsr_densify_plan reported */
    int32_t n_tensors;
    int32_t reserved;
    const float* scaling;            /* P*3 raw, of the SOURCE model (for GSR_DENSIFY_XYZ rows) */
    const float* rotation;           /* P*4 raw */
    const float* noise;              /* 2 x noise_stride x 3 standard-normal numbers, [copy][source Gaussian][axis]:
                                        the draw of torch.normal (:521) as an input; may be NULL when C == 0 */
    int64_t noise_stride;            /* >= P */
    GsrDensifyTensor tensors[GSR_DENSIFY_MAX_TENSORS];
} GsrDensifyApplyArgs;

int64_t gsr_densify_plan_words(int32_t P);
int gsr_densify_plan(const GsrDensifyPlanArgs* a, uint32_t* plan, int32_t* counts_host, void* stream);
int gsr_densify_apply(const GsrDensifyApplyArgs* a, const uint32_t* plan, void* stream);
/* add_densification_stats (:630-633) fused with train.py:212-213.  The Gaussians of the view are selected by ONE of
 *   visible  P bytes (a boolean mask),
 *   indices  n_indices int64 row numbers, unique (what render() returns: `(radii > 0).nonzero()`,
 *            gaussian_renderer/__init__.py:152),
 *   neither  radii[i] > 0 (the definition itself; needs radii).
 * For each selected i: max_radii2D[i] = max(max_radii2D[i], radii[i]) (skipped when either is NULL);
 * xyz_gradient_accum[i] += ||viewspace_grad[i, 0:2]||; denom[i] += 1.   viewspace_grad is P x 3. */
int gsr_densify_stats(int32_t P, const float* viewspace_grad, const uint8_t* visible, const int64_t* indices,
                      int64_t n_indices, const int32_t* radii, float* xyz_gradient_accum, float* denom,
                      float* max_radii2D, void* stream);
/* reset_opacity (:323-327): opacity = inverse_sigmoid(min(sigmoid(opacity), 0.01)) in place; both moments
 * zeroed (replace_tensor_to_optimizer, :400-412; either may be NULL). */
int gsr_reset_opacity(int32_t P, float* opacity, float* exp_avg, float* exp_avg_sq, void* stream);

/* Gaussian file format: rows of the binary_little_endian PLY that save_ply writes
 * (gaussian_model.py:285-321): x y z nx ny nz f_dc_0..2 f_rest_0..3K-1 opacity scale_0..2 rot_0..3 as
 * fp32, with f_rest channel-major (the transpose(1, 2) at :309).  pack: model tensors -> rows
 * [P][17 + 3K]; unpack: rows -> model tensors (load_ply, :329-380).  All device pointers. */
int gsr_ply_pack(int32_t P, int32_t K_rest, const float* xyz, const float* f_dc, const float* f_rest,
                 const float* opacity, const float* scaling, const float* rotation, float* rows, void* stream);
int gsr_ply_unpack(int32_t P, int32_t K_rest, const float* rows, float* xyz, float* f_dc, float* f_rest,
                   float* opacity, float* scaling, float* rotation, void* stream);

/* Test/inspection hooks (bit-exact parity of the binning stage):
 * rebuild the reference's sorted 64-bit key list (tile<<32 | depth bits) and
 * expose point_list / ranges held in the state buffers.  Device pointers. */
int gsr_debug_sorted_keys(const GsrParams* p, const GsrBuffers* buf, int64_t R,
                          uint64_t* keys_out /* R */, void* stream);
/* Work descriptors of the last forward, by an independent counting pass:
 * counts[0] = entries visited by the forward blend (Wf), counts[1] = blended (pixel, Gaussian)
 * pairs (Bp), counts[2] = sum of n_contrib (Wb).  counts: device uint64[3]. Synchronises. */
int gsr_debug_count_work(const GsrParams* p, const GsrBuffers* buf, int64_t R,
                         uint64_t* counts, void* stream);
int gsr_debug_state_ptrs(const GsrParams* p, const GsrBuffers* buf, int64_t R,
                         const uint32_t** point_list /* R */,
                         const uint32_t** ranges /* T*2 */,
                         const float** final_T /* H*W */,
                         const uint32_t** n_contrib /* H*W */,
                         const float** records /* P*GSR_RECORD_FLOATS */);

#define GSR_RECORD_FLOATS 16

/* Measurement hook (no counterpart in the reference): launches `blocks` CTAs of 256 threads that do nothing but
 * 8 independent scalar FFMA chains for 4*iters steps each; *flops gets the floating-point operations of the launch.
 * The caller times it with events on `stream`: the FP32 rate the blend kernels' roofline is quoted against
 * (SURVEY.md 8d asks for the peak at run time, not a nominal one).  sink: one device float, never written. */
int gsr_debug_fp32_peak(int32_t blocks, int32_t iters, float* sink, double* flops, void* stream);

/* Measurement hooks used by bench.py (no effect on results).
 * gsr_profile_enable(1): every stage launcher brackets its launches with CUDA events recorded
 * on the launching stream (cudaEventRecord only; no synchronisation).  gsr_profile_collect
 * waits for the recorded events, adds each stage's elapsed milliseconds / call counts into the
 * two GSR_STAGE_COUNT-sized arrays (either may be NULL) and clears the record.
 * gsr_launch_count: kernels launched by this library in this process so far. */
#define GSR_STAGE_PREPROCESS      0
#define GSR_STAGE_DEPTH_SORT      1
#define GSR_STAGE_TILE_COUNT      2
#define GSR_STAGE_TILE_FILL       3
#define GSR_STAGE_RENDER_FWD      4
#define GSR_STAGE_RENDER_BWD      5
#define GSR_STAGE_PREPROCESS_BWD  6
#define GSR_STAGE_EXCHANGE        7
#define GSR_STAGE_COUNT           8
int         gsr_profile_enable(int on);
int         gsr_profile_collect(double* ms_sum, int64_t* calls);
const char* gsr_stage_name(int stage);
uint64_t    gsr_launch_count(void);

/* View-parallel helper: number of floats per Gaussian in the
 * flat parameter-gradient buffer that is all-reduced (xyz 3, sh 3*M, opacity 1,
 * scale 3, rotation 4). */
int32_t gsr_flat_grad_floats(int32_t M);

#ifdef __cplusplus
}
#endif
#endif /* GSRAST_H_INCLUDED */
