"""Builds the reference's own CUDA rasterizer for sm_100 into oracle/_ref/ (git-ignored).

TEST / BASELINE INFRASTRUCTURE ONLY.  Runs where /root/reference exists (the authoring
container); the resulting .so files travel to the GPU box with the repo snapshot.  Reference
sources are compiled from a scratch copy under /tmp — none are copied into the repository
(oracle/_ref/ holds only build outputs and the reference's own Python wrapper, and is ignored
by git).

The fork cannot be imported as shipped, so the build applies the minimal
repairs, all of them additive:
  F3  `-include cstdint`                      (rasterizer_impl.h:29,48-49 miss the header)
  F1  append the missing BACKWARD::render launcher (declared backward.h:25-43, called
      rasterizer_impl.cu:428, never defined) — a plain <<<grid, block>>> launch of the
      reference's own renderCUDA<NUM_CHANNELS> kernel (backward.cu:433-612);
  F2  variant `restored` only: re-insert the tail of computeCov2DCUDA that the fork replaced
      by a comment (backward.cu:285-290), written from the chain rule of
      cov2D = T^T Vrk^T T.  Variant `asis` leaves the truncation.

The reference's fused-ssim submodule is built unmodified the same way
(`ssim`): submodules/fused-ssim/{ssim.cu, ext.cpp} -> oracle/_ref/fused_ssim/fused_ssim_cuda.so plus its
own Python wrapper.

Usage:  python oracle/build_ref.py [asis] [restored] [ssim] [--force]
"""
import os
import shutil
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DGR = "/root/reference/submodules/diff-gaussian-rasterization"
OUT_ROOT = os.path.join(HERE, "_ref")
TMP_ROOT = "/tmp/gsref_build"

LAUNCHER = r'''

// ---- repair F1 (added by oracle/build_ref.py): the launcher the fork declares but omits ----
void BACKWARD::render(
    const dim3 grid, dim3 block, const uint2* ranges, const uint32_t* point_list, int W, int H,
    const float* bg_color, const float2* means2D, const float4* conic_opacity, const float* colors,
    const float* depths, const float* final_Ts, const uint32_t* n_contrib, const float* dL_dpixels,
    const float* dL_invdepths, float3* dL_dmean2D, float4* dL_dconic2D, float* dL_dopacity,
    float* dL_dcolors, float* dL_dinvdepths)
{
    renderCUDA<NUM_CHANNELS><<<grid, block>>>(
        ranges, point_list, W, H, bg_color, means2D, conic_opacity, colors, depths, final_Ts,
        n_contrib, dL_dpixels, dL_invdepths, dL_dmean2D, dL_dconic2D, dL_dopacity, dL_dcolors,
        dL_dinvdepths);
}
'''

# Tail of computeCov2DCUDA (variant `restored`).  Uses the kernel's own locals:
# T, Vrk, W (glm::mat3), t, h_x, h_y, x_grad_mul, y_grad_mul, dL_dc_xx/xy/yy, view_matrix.
TAIL = r'''
	// ---- repair F2 (added by oracle/build_ref.py): chain rule of cov2D = T^T Vrk^T T ----
	{
		const float dL_da = dL_dc_xx, dL_db = dL_dc_xy, dL_dc = dL_dc_yy;
		dL_dcov[6 * idx + 0] = (T[0][0] * T[0][0] * dL_da + T[0][0] * T[1][0] * dL_db + T[1][0] * T[1][0] * dL_dc);
		dL_dcov[6 * idx + 3] = (T[0][1] * T[0][1] * dL_da + T[0][1] * T[1][1] * dL_db + T[1][1] * T[1][1] * dL_dc);
		dL_dcov[6 * idx + 5] = (T[0][2] * T[0][2] * dL_da + T[0][2] * T[1][2] * dL_db + T[1][2] * T[1][2] * dL_dc);
		dL_dcov[6 * idx + 1] = 2 * T[0][0] * T[0][1] * dL_da + (T[0][0] * T[1][1] + T[0][1] * T[1][0]) * dL_db + 2 * T[1][0] * T[1][1] * dL_dc;
		dL_dcov[6 * idx + 2] = 2 * T[0][0] * T[0][2] * dL_da + (T[0][0] * T[1][2] + T[0][2] * T[1][0]) * dL_db + 2 * T[1][0] * T[1][2] * dL_dc;
		dL_dcov[6 * idx + 4] = 2 * T[0][2] * T[0][1] * dL_da + (T[0][1] * T[1][2] + T[0][2] * T[1][1]) * dL_db + 2 * T[1][1] * T[1][2] * dL_dc;

		float dL_dT00 = 2 * (T[0][0] * Vrk[0][0] + T[0][1] * Vrk[0][1] + T[0][2] * Vrk[0][2]) * dL_da +
			(T[1][0] * Vrk[0][0] + T[1][1] * Vrk[0][1] + T[1][2] * Vrk[0][2]) * dL_db;
		float dL_dT01 = 2 * (T[0][0] * Vrk[1][0] + T[0][1] * Vrk[1][1] + T[0][2] * Vrk[1][2]) * dL_da +
			(T[1][0] * Vrk[1][0] + T[1][1] * Vrk[1][1] + T[1][2] * Vrk[1][2]) * dL_db;
		float dL_dT02 = 2 * (T[0][0] * Vrk[2][0] + T[0][1] * Vrk[2][1] + T[0][2] * Vrk[2][2]) * dL_da +
			(T[1][0] * Vrk[2][0] + T[1][1] * Vrk[2][1] + T[1][2] * Vrk[2][2]) * dL_db;
		float dL_dT10 = 2 * (T[1][0] * Vrk[0][0] + T[1][1] * Vrk[0][1] + T[1][2] * Vrk[0][2]) * dL_dc +
			(T[0][0] * Vrk[0][0] + T[0][1] * Vrk[0][1] + T[0][2] * Vrk[0][2]) * dL_db;
		float dL_dT11 = 2 * (T[1][0] * Vrk[1][0] + T[1][1] * Vrk[1][1] + T[1][2] * Vrk[1][2]) * dL_dc +
			(T[0][0] * Vrk[1][0] + T[0][1] * Vrk[1][1] + T[0][2] * Vrk[1][2]) * dL_db;
		float dL_dT12 = 2 * (T[1][0] * Vrk[2][0] + T[1][1] * Vrk[2][1] + T[1][2] * Vrk[2][2]) * dL_dc +
			(T[0][0] * Vrk[2][0] + T[0][1] * Vrk[2][1] + T[0][2] * Vrk[2][2]) * dL_db;

		float dL_dJ00 = W[0][0] * dL_dT00 + W[0][1] * dL_dT01 + W[0][2] * dL_dT02;
		float dL_dJ02 = W[2][0] * dL_dT00 + W[2][1] * dL_dT01 + W[2][2] * dL_dT02;
		float dL_dJ11 = W[1][0] * dL_dT10 + W[1][1] * dL_dT11 + W[1][2] * dL_dT12;
		float dL_dJ12 = W[2][0] * dL_dT10 + W[2][1] * dL_dT11 + W[2][2] * dL_dT12;

		float tz = 1.f / t.z;
		float tz2 = tz * tz;
		float tz3 = tz2 * tz;
		float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
		float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
		float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * t.x) * tz3 * dL_dJ02 + (2 * h_y * t.y) * tz3 * dL_dJ12;
		if (dL_dinvdepth)
			dL_dtz -= dL_dinvdepth[idx] / (t.z * t.z);
		float3 dL_dmean = transformVec4x3Transpose({ dL_dtx, dL_dty, dL_dtz }, view_matrix);
		dL_dmeans[idx] = dL_dmean;
	}
'''

MARKER = "(更多的反向传播计算)"


def _torch_paths():
    import torch
    from torch.utils import cpp_extension
    inc = cpp_extension.include_paths()
    libdir = os.path.join(os.path.dirname(torch.__file__), "lib")
    abi = int(torch._C._GLIBCXX_USE_CXX11_ABI)
    return inc, libdir, abi


def build_variant(variant, force=False):
    assert variant in ("asis", "restored")
    out_pkg = os.path.join(OUT_ROOT, variant, "diff_gaussian_rasterization")
    out_so = os.path.join(out_pkg, "_C.so")
    if os.path.exists(out_so) and not force:
        return out_so
    if not os.path.isdir(REF_DGR):
        raise RuntimeError(f"{REF_DGR} not present: the reference can only be built in the authoring container")
    tmp = os.path.join(TMP_ROOT, variant)
    shutil.rmtree(tmp, ignore_errors=True)
    os.makedirs(tmp)
    shutil.copytree(os.path.join(REF_DGR, "cuda_rasterizer"), os.path.join(tmp, "cuda_rasterizer"))
    for f in ("rasterize_points.cu", "rasterize_points.h", "ext.cpp"):
        shutil.copy(os.path.join(REF_DGR, f), tmp)
    bpath = os.path.join(tmp, "cuda_rasterizer", "backward.cu")
    src = open(bpath, encoding="utf-8").read()
    if variant == "restored":
        lines = src.split("\n")
        hit = [i for i, l in enumerate(lines) if MARKER in l]
        assert len(hit) == 1, "truncation marker not found exactly once in backward.cu"
        lines.insert(hit[0] + 1, TAIL)
        src = "\n".join(lines)
    src += LAUNCHER
    open(bpath, "w", encoding="utf-8").write(src)

    inc, libdir, abi = _torch_paths()
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    common = [nvcc, "-std=c++17", "-O3", "-gencode", "arch=compute_100,code=sm_100",
              "-Xcompiler", "-fPIC", "-include", "cstdint",
              "-DTORCH_EXTENSION_NAME=_C", "-DTORCH_API_INCLUDE_EXTENSION_H",
              f"-D_GLIBCXX_USE_CXX11_ABI={abi}", "--expt-relaxed-constexpr", "-w",
              "-I" + os.path.join(REF_DGR, "third_party", "glm"), "-I" + sysconfig.get_paths()["include"]]
    for i in inc:
        common += ["-I" + i]
    objs = []
    procs = []
    for s in ("cuda_rasterizer/rasterizer_impl.cu", "cuda_rasterizer/forward.cu", "cuda_rasterizer/backward.cu",
              "rasterize_points.cu", "ext.cpp"):
        o = os.path.join(tmp, os.path.basename(s) + ".o")
        objs.append(o)
        extra = ["-x", "cu"] if s.endswith(".cpp") else []
        procs.append(subprocess.Popen(common + extra + ["-c", os.path.join(tmp, s), "-o", o], cwd=tmp))
    for p in procs:
        if p.wait() != 0:
            raise RuntimeError("reference compilation failed")
    os.makedirs(out_pkg, exist_ok=True)
    link = [nvcc, "-shared", "-cudart", "static", "-o", out_so] + objs + \
           ["-L" + libdir, "-lc10", "-ltorch", "-ltorch_cpu", "-ltorch_python", "-lc10_cuda", "-ltorch_cuda",
            "-Xlinker", "-rpath," + libdir]
    subprocess.check_call(link, cwd=tmp)
    # the reference's own Python wrapper, unmodified (kept out of git: oracle/_ref/ is ignored)
    shutil.copy(os.path.join(REF_DGR, "diff_gaussian_rasterization", "__init__.py"),
                os.path.join(out_pkg, "__init__.py"))
    return out_so


REF_SSIM = "/root/reference/submodules/fused-ssim"


def build_fused_ssim(force=False):
    """The reference's fused-ssim extension, unmodified, for sm_100."""
    out_dir = os.path.join(OUT_ROOT, "fused_ssim")
    out_so = os.path.join(out_dir, "fused_ssim_cuda.so")
    if os.path.exists(out_so) and not force:
        return out_so
    if not os.path.isdir(REF_SSIM):
        raise RuntimeError(f"{REF_SSIM} not present: the reference can only be built in the authoring container")
    tmp = os.path.join(TMP_ROOT, "fused_ssim")
    shutil.rmtree(tmp, ignore_errors=True)
    os.makedirs(tmp)
    inc, libdir, abi = _torch_paths()
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    common = [nvcc, "-std=c++17", "-O3", "-gencode", "arch=compute_100,code=sm_100", "-Xcompiler", "-fPIC",
              "-DTORCH_EXTENSION_NAME=fused_ssim_cuda", "-DTORCH_API_INCLUDE_EXTENSION_H",
              f"-D_GLIBCXX_USE_CXX11_ABI={abi}", "--expt-relaxed-constexpr", "-w",
              "-I" + REF_SSIM, "-I" + sysconfig.get_paths()["include"]]
    for i in inc:
        common += ["-I" + i]
    objs, procs = [], []
    for s in ("ssim.cu", "ext.cpp"):                      # compiled where they lie; only objects go to /tmp
        o = os.path.join(tmp, s + ".o")
        objs.append(o)
        extra = ["-x", "cu"] if s.endswith(".cpp") else []
        procs.append(subprocess.Popen(common + extra + ["-c", os.path.join(REF_SSIM, s), "-o", o], cwd=tmp))
    for p in procs:
        if p.wait() != 0:
            raise RuntimeError("fused-ssim reference compilation failed")
    os.makedirs(os.path.join(out_dir, "fused_ssim"), exist_ok=True)
    link = [nvcc, "-shared", "-cudart", "static", "-o", out_so] + objs + \
           ["-L" + libdir, "-lc10", "-ltorch", "-ltorch_cpu", "-ltorch_python", "-lc10_cuda", "-ltorch_cuda",
            "-Xlinker", "-rpath," + libdir]
    subprocess.check_call(link, cwd=tmp)
    shutil.copy(os.path.join(REF_SSIM, "fused_ssim", "__init__.py"), os.path.join(out_dir, "fused_ssim", "__init__.py"))
    return out_so


if __name__ == "__main__":
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    variants = [a for a in args if a in ("asis", "restored")] or ([] if args else ["restored", "asis"])
    for v in variants:
        print(build_variant(v, force="--force" in sys.argv))
    if "ssim" in args or not args:
        print(build_fused_ssim(force="--force" in sys.argv))
