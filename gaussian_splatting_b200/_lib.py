"""ctypes binding of libgsrast.so (the C-ABI declared in include/gsrast.h).

There is no fallback: if the shared library is missing or a call fails, a RuntimeError is
raised.  Build it with `python -m gaussian_splatting_b200.build` (nvcc, sm_100a).
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# GSR_LIB: alternative build of the same library (A/B measurements of kernel variants)
LIB_PATH = os.environ.get("GSR_LIB") or os.path.join(_HERE, "libgsrast.so")

GSR_OK = 0
RAW_SCALES, RAW_ROTATIONS, RAW_OPACITIES = 1, 2, 4
ACC_CLEAR_IN_FORWARD, ACC_IS_CLEAR = 1, 2      # GsrBuffers.accumulator (GSR_ACC_*)
GSR_ERR_INVALID_ARG = -1
GSR_ERR_CUDA = -2
GSR_ERR_CAPACITY = -3
GSR_ERR_OVERFLOW = -4

_fp = C.c_void_p   # device pointers are passed as integers


class GsrParams(C.Structure):
    _fields_ = [
        ("P", C.c_int32), ("D", C.c_int32), ("M", C.c_int32), ("W", C.c_int32), ("H", C.c_int32),
        ("tanfovx", C.c_float), ("tanfovy", C.c_float), ("scale_modifier", C.c_float),
        ("prefiltered", C.c_int32), ("antialiasing", C.c_int32), ("debug", C.c_int32),
        ("raw_inputs", C.c_int32),
        ("background", _fp), ("means3D", _fp), ("shs", _fp), ("colors_precomp", _fp),
        ("opacities", _fp), ("scales", _fp), ("rotations", _fp), ("cov3D_precomp", _fp),
        ("viewmatrix", _fp), ("projmatrix", _fp), ("campos", _fp), ("sh_dc", _fp),
    ]


class GsrBuffers(C.Structure):
    _fields_ = [
        ("geom", _fp), ("geom_bytes", C.c_size_t),
        ("img", _fp), ("img_bytes", C.c_size_t),
        ("bin", _fp), ("bin_bytes", C.c_size_t),
        ("scratch", _fp), ("scratch_bytes", C.c_size_t),
        ("num_rendered_host", _fp), ("num_rendered_event", _fp),
        ("accumulator", C.c_int32), ("reserved2", C.c_int32),
    ]


class GsrForwardOut(C.Structure):
    _fields_ = [("out_color", _fp), ("out_invdepth", _fp), ("radii", _fp)]


class GsrBackwardIO(C.Structure):
    _fields_ = [
        ("dL_dout_color", _fp), ("dL_dout_invdepth", _fp),
        ("dL_dmeans3D", _fp), ("dL_dmeans2D", _fp), ("dL_dopacity", _fp), ("dL_dcolors", _fp),
        ("dL_dcov3D", _fp), ("dL_dsh", _fp), ("dL_dscales", _fp), ("dL_drotations", _fp),
        ("multimem_reduce", C.c_int32), ("reserved1", C.c_int32),
        ("xchg_records", _fp), ("xchg_rgb", _fp), ("xchg_campos", _fp), ("dL_dsh_dc", _fp),
    ]


class GsrSsimParams(C.Structure):
    _fields_ = [
        ("img1", _fp), ("img2", _fp),
        ("B", C.c_int32), ("CH", C.c_int32), ("H", C.c_int32), ("W", C.c_int32),
        ("C1", C.c_float), ("C2", C.c_float), ("crop", C.c_int32),
    ]


class GsrLossParams(C.Structure):
    _fields_ = [
        ("image", _fp), ("gt", _fp), ("invdepth", _fp), ("mono_invdepth", _fp), ("depth_mask", _fp),
        ("CH", C.c_int32), ("H", C.c_int32), ("W", C.c_int32),
        ("lambda_dssim", C.c_float), ("depth_weight", C.c_float), ("C1", C.c_float), ("C2", C.c_float),
        ("reserved", C.c_int32),
    ]


DENSIFY_COPY, DENSIFY_MOMENT, DENSIFY_XYZ, DENSIFY_SCALING = 0, 1, 2, 3
DENSIFY_MAX_TENSORS = 24


class GsrDensifyPlanArgs(C.Structure):
    _fields_ = [
        ("P", C.c_int32), ("use_screen_size", C.c_int32),
        ("grad_threshold", C.c_float), ("dense_extent", C.c_float), ("min_opacity", C.c_float),
        ("max_screen_size", C.c_float), ("world_extent_limit", C.c_float), ("reserved", C.c_int32),
        ("xyz_gradient_accum", _fp), ("denom", _fp), ("scaling", _fp), ("opacity", _fp), ("prune_mask", _fp),
    ]


class GsrDensifyTensor(C.Structure):
    _fields_ = [("src", _fp), ("dst", _fp), ("width", C.c_int32), ("kind", C.c_int32)]


class GsrDensifyApplyArgs(C.Structure):
    _fields_ = [
        ("P", C.c_int32), ("A", C.c_int32), ("B", C.c_int32), ("C", C.c_int32),
        ("n_tensors", C.c_int32), ("reserved", C.c_int32),
        ("scaling", _fp), ("rotation", _fp), ("noise", _fp), ("noise_stride", C.c_int64),
        ("tensors", GsrDensifyTensor * DENSIFY_MAX_TENSORS),
    ]


# every symbol include/gsrast.h declares (tests check that the library exports all of them)
EXPORTS = (
    "gsr_version", "gsr_last_error", "gsr_geom_bytes", "gsr_img_bytes", "gsr_bin_bytes",
    "gsr_scratch_bytes", "gsr_forward_stage1", "gsr_forward_stage2", "gsr_forward", "gsr_backward",
    "gsr_mark_visible", "gsr_debug_sorted_keys", "gsr_debug_state_ptrs", "gsr_flat_grad_floats",
    "gsr_debug_count_work", "gsr_debug_fp32_peak",
    "gsr_profile_enable", "gsr_profile_collect", "gsr_stage_name", "gsr_launch_count", "gsr_adam_update",
    "gsr_ssim_partials", "gsr_ssim_forward", "gsr_ssim_forward_noalloc", "gsr_ssim_backward", "gsr_exchange_expand",
    "gsr_densify_plan_words", "gsr_densify_plan", "gsr_densify_apply", "gsr_densify_stats", "gsr_reset_opacity",
    "gsr_ply_pack", "gsr_ply_unpack", "gsr_loss_partials", "gsr_loss_forward", "gsr_loss_backward",
)
STAGE_COUNT = 8

_lib = None


def bind_densify(lib):
    """Signatures of the density-control and PLY-row entry points."""
    lib.gsr_densify_plan_words.restype = C.c_int64
    lib.gsr_densify_plan_words.argtypes = [C.c_int32]
    lib.gsr_densify_plan.restype = C.c_int
    lib.gsr_densify_plan.argtypes = [C.POINTER(GsrDensifyPlanArgs), C.c_void_p, C.c_void_p, C.c_void_p]
    lib.gsr_densify_apply.restype = C.c_int
    lib.gsr_densify_apply.argtypes = [C.POINTER(GsrDensifyApplyArgs), C.c_void_p, C.c_void_p]
    lib.gsr_densify_stats.restype = C.c_int
    lib.gsr_densify_stats.argtypes = [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64] + [C.c_void_p] * 5
    lib.gsr_reset_opacity.restype = C.c_int
    lib.gsr_reset_opacity.argtypes = [C.c_int32] + [C.c_void_p] * 4
    lib.gsr_ply_pack.restype = C.c_int
    lib.gsr_ply_pack.argtypes = [C.c_int32, C.c_int32] + [C.c_void_p] * 8
    lib.gsr_ply_unpack.restype = C.c_int
    lib.gsr_ply_unpack.argtypes = [C.c_int32, C.c_int32] + [C.c_void_p] * 8
    return lib


def load():
    """Load libgsrast.so once; raises RuntimeError (never falls back) if it is unavailable."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: the CUDA extension is required (no CPU fallback). "
            "Build it with `python -m gaussian_splatting_b200.build`.")
    lib = C.CDLL(LIB_PATH)
    lib.gsr_version.restype = C.c_int
    lib.gsr_last_error.restype = C.c_char_p
    lib.gsr_geom_bytes.restype = C.c_size_t
    lib.gsr_geom_bytes.argtypes = [C.c_int32]
    lib.gsr_img_bytes.restype = C.c_size_t
    lib.gsr_img_bytes.argtypes = [C.c_int32, C.c_int32]
    lib.gsr_bin_bytes.restype = C.c_size_t
    lib.gsr_bin_bytes.argtypes = [C.c_int64]
    lib.gsr_scratch_bytes.restype = C.c_size_t
    lib.gsr_scratch_bytes.argtypes = [C.c_int32, C.c_int32, C.c_int32, C.c_int64]
    lib.gsr_flat_grad_floats.restype = C.c_int32
    lib.gsr_flat_grad_floats.argtypes = [C.c_int32]
    PP, PB, PO = C.POINTER(GsrParams), C.POINTER(GsrBuffers), C.POINTER(GsrForwardOut)
    lib.gsr_forward_stage1.restype = C.c_int
    lib.gsr_forward_stage1.argtypes = [PP, PB, PO, C.c_void_p]
    lib.gsr_forward_stage2.restype = C.c_int
    lib.gsr_forward_stage2.argtypes = [PP, PB, PO, C.c_int64, C.c_void_p]
    lib.gsr_forward.restype = C.c_int
    lib.gsr_forward.argtypes = [PP, PB, PO, C.POINTER(C.c_int64), C.c_void_p]
    lib.gsr_backward.restype = C.c_int
    lib.gsr_backward.argtypes = [PP, PB, C.c_int64, C.c_void_p, C.POINTER(GsrBackwardIO), C.c_void_p]
    lib.gsr_mark_visible.restype = C.c_int
    lib.gsr_mark_visible.argtypes = [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.gsr_debug_sorted_keys.restype = C.c_int
    lib.gsr_debug_sorted_keys.argtypes = [PP, PB, C.c_int64, C.c_void_p, C.c_void_p]
    lib.gsr_debug_state_ptrs.restype = C.c_int
    lib.gsr_debug_state_ptrs.argtypes = [PP, PB, C.c_int64] + [C.POINTER(C.c_void_p)] * 5
    lib.gsr_debug_count_work.restype = C.c_int
    lib.gsr_debug_count_work.argtypes = [PP, PB, C.c_int64, C.c_void_p, C.c_void_p]
    lib.gsr_debug_fp32_peak.restype = C.c_int
    lib.gsr_debug_fp32_peak.argtypes = [C.c_int32, C.c_int32, C.c_void_p, C.POINTER(C.c_double), C.c_void_p]
    lib.gsr_profile_enable.restype = C.c_int
    lib.gsr_profile_enable.argtypes = [C.c_int]
    lib.gsr_profile_collect.restype = C.c_int
    lib.gsr_profile_collect.argtypes = [C.POINTER(C.c_double), C.POINTER(C.c_int64)]
    lib.gsr_stage_name.restype = C.c_char_p
    lib.gsr_stage_name.argtypes = [C.c_int]
    lib.gsr_launch_count.restype = C.c_uint64
    lib.gsr_adam_update.restype = C.c_int
    lib.gsr_adam_update.argtypes = [C.c_void_p] * 5 + [C.c_double] * 6 + [C.c_int64, C.c_int32, C.c_void_p]
    lib.gsr_exchange_expand.restype = C.c_int
    lib.gsr_exchange_expand.argtypes = [PP, C.c_int32] + [C.c_void_p] * 10
    PS = C.POINTER(GsrSsimParams)
    lib.gsr_ssim_partials.restype = C.c_int64
    lib.gsr_ssim_partials.argtypes = [C.c_int32] * 4
    lib.gsr_ssim_forward.restype = C.c_int
    lib.gsr_ssim_forward.argtypes = [PS] + [C.c_void_p] * 7
    lib.gsr_ssim_forward_noalloc.restype = C.c_int
    lib.gsr_ssim_forward_noalloc.argtypes = [PS] + [C.c_void_p] * 6
    lib.gsr_ssim_backward.restype = C.c_int
    lib.gsr_ssim_backward.argtypes = [PS] + [C.c_void_p] * 7
    PL = C.POINTER(GsrLossParams)
    lib.gsr_loss_partials.restype = C.c_int64
    lib.gsr_loss_partials.argtypes = [C.c_int32] * 3
    lib.gsr_loss_forward.restype = C.c_int
    lib.gsr_loss_forward.argtypes = [PL] + [C.c_void_p] * 6
    lib.gsr_loss_backward.restype = C.c_int
    lib.gsr_loss_backward.argtypes = [PL] + [C.c_void_p] * 7
    bind_densify(lib)
    _lib = lib
    return lib


def profile_enable(on=True):
    check(load().gsr_profile_enable(int(bool(on))))


def profile_collect():
    """{stage name: (total ms, calls)} of the spans recorded since the last collect."""
    lib = load()
    ms = (C.c_double * STAGE_COUNT)()
    n = (C.c_int64 * STAGE_COUNT)()
    check(lib.gsr_profile_collect(ms, n))
    return {lib.gsr_stage_name(i).decode(): (ms[i], n[i]) for i in range(STAGE_COUNT)}


def launch_count():
    return int(load().gsr_launch_count())


def check(rc):
    if rc != GSR_OK:
        msg = load().gsr_last_error().decode("utf-8", "replace")
        raise RuntimeError(msg or f"libgsrast error {rc}")
