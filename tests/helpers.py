"""Shared test helpers: small seeded scenes and comparison metrics."""
import numpy as np
import torch

from gaussian_splatting_b200 import synthetic


def small_scene(P=300, W=48, H=40, seed=7, sh_degree=3, scale_mult=6.0, fovx=0.9, antialiasing=False,
                bg=(0.0, 0.0, 0.0), yaw_deg=0.0, behind_fraction=0.1):
    """Gaussians + settings dict (CPU tensors). A fraction of the Gaussians is pushed behind the
    near plane so that culling is exercised; sizes are inflated so tiles overlap heavily."""
    gs = synthetic.make_gaussians(P, W, H, seed, sh_degree=sh_degree, scale_mult=scale_mult)
    cam = synthetic.make_camera(W, H, fovx, yaw_deg)
    nb = int(P * behind_fraction)
    if nb:
        g = torch.Generator().manual_seed(seed + 99)
        idx = torch.randperm(P, generator=g)[:nb]
        gs["means3D"][idx] = gs["means3D"][idx] * 0.2 + cam["campos"][None, :] * 1.02
    s = dict(cam)
    s.update(bg=torch.tensor(bg, dtype=torch.float32), scale_modifier=1.0, sh_degree=sh_degree,
             antialiasing=antialiasing, prefiltered=False, debug=False)
    return gs, s


def np_inputs(gs):
    return {k: (v.numpy() if isinstance(v, torch.Tensor) else v) for k, v in gs.items()}


def np_settings(s):
    return {k: (v.numpy() if isinstance(v, torch.Tensor) else v) for k, v in s.items()}


def rel_err(a, b, eps=1e-12):
    """Norm-wise relative error  ||a-b||_inf / max(||b||_inf, eps)  (SURVEY.md §8c)."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), eps)) if a.size else 0.0


def elem_rel_err(a, b, floor_frac=1e-3):
    """Max per-element relative error over elements with |b| > floor_frac * ||b||_inf."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    if not a.size:
        return 0.0
    m = np.abs(b) > floor_frac * np.abs(b).max()
    if not m.any():
        return 0.0
    return float((np.abs(a - b)[m] / np.abs(b)[m]).max())


# ------------------------------------------------------------------ GPU-side helpers
def settings_for(module, s, device="cuda"):
    """GaussianRasterizationSettings of `module` (ours or the reference build) from a settings dict."""
    return module.GaussianRasterizationSettings(
        image_height=int(s["image_height"]), image_width=int(s["image_width"]),
        tanfovx=float(s["tanfovx"]), tanfovy=float(s["tanfovy"]), bg=s["bg"].to(device),
        scale_modifier=float(s["scale_modifier"]), viewmatrix=s["viewmatrix"].to(device),
        projmatrix=s["projmatrix"].to(device), sh_degree=int(s["sh_degree"]), campos=s["campos"].to(device),
        prefiltered=bool(s["prefiltered"]), debug=bool(s["debug"]), antialiasing=bool(s["antialiasing"]))


GRAD_KEYS = ("means3D", "means2D", "opacities", "shs", "colors_precomp", "scales", "rotations", "cov3D_precomp")


def run_module(module, gs, s, dc=None, dd=None, device="cuda"):
    """Forward (+ backward when dc is given) through a GaussianRasterizer of `module`.

    Returns dict(color, invdepth, radii, grads{name: ndarray}) as numpy arrays."""
    rs = settings_for(module, s, device)
    leaves = {k: v.to(device).clone().requires_grad_(dc is not None) for k, v in gs.items()}
    P = gs["means3D"].shape[0]
    means2D = torch.zeros(P, 3, device=device, requires_grad=dc is not None)
    rast = module.GaussianRasterizer(rs)
    color, radii, invd = rast(
        means3D=leaves["means3D"], means2D=means2D, opacities=leaves["opacities"],
        shs=leaves.get("shs"), colors_precomp=leaves.get("colors_precomp"), scales=leaves.get("scales"),
        rotations=leaves.get("rotations"), cov3D_precomp=leaves.get("cov3D_precomp"))
    out = dict(color=color.detach().cpu().numpy(), invdepth=invd.detach().cpu().numpy(),
               radii=radii.cpu().numpy(), grads={})
    if dc is not None:
        loss = (color * dc.to(device)).sum()
        if dd is not None:
            loss = loss + (invd * dd.to(device)).sum()
        loss.backward()
        for k, v in leaves.items():
            if v.grad is not None:
                out["grads"][k] = v.grad.cpu().numpy()
        out["grads"]["means2D"] = means2D.grad.cpu().numpy()
    return out


def run_oracle(oracle, gs, s, dc=None, dd=None, restored=True):
    fw = oracle.forward(np_inputs(gs), np_settings(s))
    out = dict(color=fw["color"], invdepth=fw["invdepth"], radii=fw["radii"], fw=fw, grads={})
    if dc is not None:
        bw = oracle.backward(fw, np_settings(s), dc.numpy(), None if dd is None else dd.numpy().reshape(-1),
                             restored=restored)
        P = gs["means3D"].shape[0]
        out["grads"] = dict(means3D=bw["dL_dmeans3D"], means2D=bw["dL_dmeans2D"],
                            opacities=bw["dL_dopacity"].reshape(P, 1), shs=bw["dL_dsh"],
                            colors_precomp=bw["dL_dcolors"], scales=bw["dL_dscales"],
                            rotations=bw["dL_drotations"], cov3D_precomp=bw["dL_dcov3D"])
    return out


def ours_forward_state(gs, s, device="cuda"):
    """Our forward with the binning / image state exposed (numpy)."""
    import diff_gaussian_rasterization as dgr
    from gaussian_splatting_b200 import rasterizer
    rs = settings_for(dgr, s, device)
    t = {k: v.to(device) for k, v in gs.items()}
    e = torch.Tensor([])
    st = {}
    R, color, radii, geom, binb, img, invd = rasterizer._forward_native(
        rs, t["means3D"], t.get("shs", e), t.get("colors_precomp", e), t["opacities"], t.get("scales", e),
        t.get("rotations", e), t.get("cov3D_precomp", e), state_out=st)
    torch.cuda.synchronize()
    out = {k: v.cpu().numpy() for k, v in st.items()}
    out.update(num_rendered=R, color=color.cpu().numpy(), radii=radii.cpu().numpy(), invdepth=invd.cpu().numpy())
    return out


def ref_forward_state(ref, gs, s, device="cuda"):
    from oracle import ref_loader
    rs = settings_for(ref, s, device)
    t = {k: v.to(device) for k, v in gs.items()}
    st = ref_loader.forward_state(ref, rs, t["means3D"], t.get("shs"), t.get("colors_precomp"), t["opacities"],
                                  t.get("scales"), t.get("rotations"), t.get("cov3D_precomp"))
    torch.cuda.synchronize()
    return {k: (v.cpu().numpy() if isinstance(v, torch.Tensor) else v) for k, v in st.items()
            if not k.endswith("Buffer")}
