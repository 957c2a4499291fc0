"""fp64 PyTorch-autograd restatement of the rasterizer forward (mathematical ground truth).

TEST INFRASTRUCTURE ONLY.  Used to validate the gradients of the CPU oracle and of the CUDA
path, in particular the tail of computeCov2DCUDA that the fork elided
(submodules/diff-gaussian-rasterization/cuda_rasterizer/backward.cu:285-290):
gradients come from torch.autograd of the forward math (forward.cu:89-190, 222-351, 367-513),
not from a hand-written backward.

Conventions kept from the reference so that the comparison is like for like:
  * the 1.3*tanfov clamp of t.x, t.y passes no gradient when active and the clamped value is
    treated as independent of t.z (x_grad_mul / y_grad_mul, backward.cu:186-187);
  * min(0.99, alpha) passes the gradient through as if unclamped (backward.cu:551-595);
  * means2D gradient is reported in the reference's units: d/d(pixel) * 0.5*W (0.5*H);
  * discrete decisions (tile rect membership, alpha < 1/255 skip, T < 1e-4 stop) are taken
    from the forward values and are not differentiated.
Dense O(P * pixels): only for small cases.
"""
import math

import torch

SH_C0 = 0.28209479177387814
SH_C1 = 0.4886025119029199
SH_C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396]
SH_C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154,
         -0.4570457994644658, 1.445305721320277, -0.5900435899266435]


def eval_sh(deg, sh, dirs):
    x, y, z = dirs[:, 0:1], dirs[:, 1:2], dirs[:, 2:3]
    res = SH_C0 * sh[:, 0]
    if deg > 0:
        res = res - SH_C1 * y * sh[:, 1] + SH_C1 * z * sh[:, 2] - SH_C1 * x * sh[:, 3]
        if deg > 1:
            xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
            res = (res + SH_C2[0] * xy * sh[:, 4] + SH_C2[1] * yz * sh[:, 5] +
                   SH_C2[2] * (2.0 * zz - xx - yy) * sh[:, 6] + SH_C2[3] * xz * sh[:, 7] +
                   SH_C2[4] * (xx - yy) * sh[:, 8])
            if deg > 2:
                res = (res + SH_C3[0] * y * (3.0 * xx - yy) * sh[:, 9] + SH_C3[1] * xy * z * sh[:, 10] +
                       SH_C3[2] * y * (4.0 * zz - xx - yy) * sh[:, 11] +
                       SH_C3[3] * z * (2.0 * zz - 3.0 * xx - 3.0 * yy) * sh[:, 12] +
                       SH_C3[4] * x * (4.0 * zz - xx - yy) * sh[:, 13] + SH_C3[5] * z * (xx - yy) * sh[:, 14] +
                       SH_C3[6] * x * (xx - 3.0 * yy) * sh[:, 15])
    return res + 0.5


def quat_to_rot(q):
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = torch.stack([
        1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
        2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
        2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], dim=1).reshape(-1, 3, 3)
    return R


def rasterize(means3D, opacities, s, shs=None, colors_precomp=None, scales=None, rotations=None,
              cov3D_precomp=None, means2D_probe=None):
    """Differentiable fp64 forward. Tensors are float64 leaves (or anything autograd-tracked).

    means2D_probe: optional (P,2) zero tensor added to the pixel centres, scaled so that its
    gradient equals the reference's dL_dmeans2D[:, :2].
    Returns color (3,H,W), invdepth (1,H,W), radii (P,) int.
    """
    dt = torch.float64
    W, H = int(s["image_width"]), int(s["image_height"])
    view = s["viewmatrix"].to(dt)      # row-major memory of W2V^T  => p_view = [p,1] @ view
    proj = s["projmatrix"].to(dt)
    campos = s["campos"].to(dt)
    bg = s["bg"].to(dt)
    tanx, tany = float(s["tanfovx"]), float(s["tanfovy"])
    fx, fy = W / (2.0 * tanx), H / (2.0 * tany)
    mod = float(s["scale_modifier"])
    P = means3D.shape[0]
    ones = torch.ones(P, 1, dtype=dt)
    hom = torch.cat([means3D, ones], 1)
    p_view = hom @ view
    p_hom = hom @ proj
    p_w = 1.0 / (p_hom[:, 3] + 1e-7)
    p_proj = p_hom[:, :3] * p_w[:, None]
    tz = p_view[:, 2]
    vis = tz > 0.2

    if cov3D_precomp is None:
        R = quat_to_rot(rotations)                       # standard rotation
        S = mod * scales
        Mx = R * S[:, None, :]                           # R @ diag(S)
        Sigma = Mx @ Mx.transpose(1, 2)
    else:
        c = cov3D_precomp
        Sigma = torch.stack([c[:, 0], c[:, 1], c[:, 2], c[:, 1], c[:, 3], c[:, 4], c[:, 2], c[:, 4], c[:, 5]], 1).reshape(-1, 3, 3)

    limx, limy = 1.3 * tanx, 1.3 * tany
    txtz = p_view[:, 0] / tz
    tytz = p_view[:, 1] / tz
    inx = (txtz >= -limx) & (txtz <= limx)
    iny = (tytz >= -limy) & (tytz <= limy)
    tx = torch.where(inx, p_view[:, 0], (txtz.clamp(-limx, limx) * tz).detach())
    ty = torch.where(iny, p_view[:, 1], (tytz.clamp(-limy, limy) * tz).detach())
    zero = torch.zeros_like(tz)
    J = torch.stack([fx / tz, zero, -(fx * tx) / (tz * tz),
                     zero, fy / tz, -(fy * ty) / (tz * tz),
                     zero, zero, zero], 1).reshape(-1, 3, 3)
    Rw = view[:3, :3].t()                                # W2V rotation
    A = J @ Rw
    cov = A @ Sigma @ A.transpose(1, 2)
    cxx, cxy, cyy = cov[:, 0, 0], cov[:, 0, 1], cov[:, 1, 1]
    det_cov = cxx * cyy - cxy * cxy
    cxx = cxx + 0.3
    cyy = cyy + 0.3
    det = cxx * cyy - cxy * cxy
    if s["antialiasing"]:
        ratio = det_cov / det
        h = torch.sqrt(torch.where(ratio > 0.000025, ratio, torch.full_like(ratio, 0.000025)))
    else:
        h = torch.ones_like(det)
    vis = vis & (det != 0)
    det_inv = 1.0 / det
    ca, cb, cc = cyy * det_inv, -cxy * det_inv, cxx * det_inv
    mid = 0.5 * (cxx + cyy)
    lam = mid + torch.sqrt(torch.clamp(mid * mid - det, min=0.1))
    radius = torch.ceil(3.0 * torch.sqrt(lam)).detach()
    px = ((p_proj[:, 0] + 1.0) * W - 1.0) * 0.5
    py = ((p_proj[:, 1] + 1.0) * H - 1.0) * 0.5
    if means2D_probe is not None:
        px = px + means2D_probe[:, 0] * (0.5 * W)
        py = py + means2D_probe[:, 1] * (0.5 * H)
    gx, gy = (W + 15) // 16, (H + 15) // 16
    pxd, pyd = px.detach(), py.detach()
    rminx = torch.clamp(torch.trunc((pxd - radius) / 16), 0, gx)
    rminy = torch.clamp(torch.trunc((pyd - radius) / 16), 0, gy)
    rmaxx = torch.clamp(torch.trunc((pxd + radius + 15) / 16), 0, gx)
    rmaxy = torch.clamp(torch.trunc((pyd + radius + 15) / 16), 0, gy)
    vis = vis & ((rmaxx - rminx) * (rmaxy - rminy) > 0)
    radii = torch.where(vis, radius, torch.zeros_like(radius)).to(torch.int32)

    if colors_precomp is None:
        d = means3D - campos[None, :]
        d = d / d.norm(dim=1, keepdim=True)
        rgb = torch.clamp_min(eval_sh(int(s["sh_degree"]), shs, d), 0.0)
    else:
        rgb = colors_precomp
    op = opacities.reshape(-1) * h
    invd = 1.0 / tz

    # per-pixel blending in global depth order (ties broken by index: stable sort)
    order = torch.argsort(tz.detach(), stable=True)
    ys, xs = torch.meshgrid(torch.arange(H, dtype=dt), torch.arange(W, dtype=dt), indexing="ij")
    tile_x, tile_y = torch.floor(xs / 16), torch.floor(ys / 16)
    T = torch.ones(H, W, dtype=dt)
    done = torch.zeros(H, W, dtype=torch.bool)
    C = [torch.zeros(H, W, dtype=dt) for _ in range(3)]
    D = torch.zeros(H, W, dtype=dt)
    for i in order.tolist():
        if not bool(vis[i]):
            continue
        in_rect = (tile_x >= rminx[i]) & (tile_x < rmaxx[i]) & (tile_y >= rminy[i]) & (tile_y < rmaxy[i])
        if not bool(in_rect.any()):
            continue
        dx = px[i] - xs
        dy = py[i] - ys
        power = -0.5 * (ca[i] * dx * dx + cc[i] * dy * dy) - cb[i] * dx * dy
        G = torch.exp(torch.clamp(power, max=0.0))
        a_raw = op[i] * G
        # value clamps at 0.99, gradient as if unclamped (reference quirk)
        alpha = a_raw + (torch.clamp(a_raw, max=0.99) - a_raw).detach()
        m = in_rect & (power.detach() <= 0) & (alpha.detach() >= 1.0 / 255.0) & ~done
        test_T = T * (1 - alpha)
        stop = m & (test_T.detach() < 0.0001)
        done = done | stop
        m = m & ~stop
        w = torch.where(m, alpha * T, torch.zeros_like(T))
        for ch in range(3):
            C[ch] = C[ch] + rgb[i, ch] * w
        D = D + invd[i] * w
        T = torch.where(m, test_T, T)
    color = torch.stack([C[ch] + T * bg[ch] for ch in range(3)], 0)
    return color, D[None], radii
