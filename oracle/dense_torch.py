"""Dense PyTorch evaluator of the reference's rasterizer semantics (every Gaussian x every pixel).

TEST INFRASTRUCTURE ONLY (see oracle/vegs_oracle.c).  It serves two purposes:
  * an independent check of the hand-derived backward formulas restated in vegs_oracle.c: here
    the gradients come from torch.autograd over the forward (run in float64), with the reference's
    hard decisions (tile membership, alpha >= 1/255, T < 1e-4 cut) treated as constant masks and the
    0.99 alpha clamp passed straight through, exactly as the reference's backward does
    (cuda_rasterizer/backward.cu:565-691);
  * the "dense PyTorch-CPU compositing evaluator" BASELINE.md §2b names as the second reported
    baseline (O(P * Npix), config 1 only).

Follows cuda_rasterizer/forward.cu:156-257 (preprocess), :262-441 (compositing) and
auxiliary.h:41-77.  The tile-membership mask is applied: a Gaussian contributes
to a pixel only if the pixel's 16x16 tile lies inside the Gaussian's tile rectangle.
"""
from __future__ import annotations

import math

import torch

SH_C0 = 0.28209479177387814
SH_C1 = 0.4886025119029199
SH_C2 = [1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792,
         0.5462742152960396]
SH_C3 = [-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154,
         -0.4570457994644658, 1.445305721320277, -0.5900435899266435]


def eval_sh(deg, sh, dirs):
    """computeColorFromSH (forward.cu:21-72): sh [P,M,3], dirs [P,3] unit -> [P,3] before +0.5."""
    x, y, z = dirs[:, 0:1], dirs[:, 1:2], dirs[:, 2:3]
    res = SH_C0 * sh[:, 0]
    if deg > 0:
        res = res - SH_C1 * y * sh[:, 1] + SH_C1 * z * sh[:, 2] - SH_C1 * x * sh[:, 3]
    if deg > 1:
        xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
        res = (res + SH_C2[0] * xy * sh[:, 4] + SH_C2[1] * yz * sh[:, 5] + SH_C2[2] * (2 * zz - xx - yy) * sh[:, 6]
               + SH_C2[3] * xz * sh[:, 7] + SH_C2[4] * (xx - yy) * sh[:, 8])
    if deg > 2:
        res = (res + SH_C3[0] * y * (3 * xx - yy) * sh[:, 9] + SH_C3[1] * xy * z * sh[:, 10]
               + SH_C3[2] * y * (4 * zz - xx - yy) * sh[:, 11] + SH_C3[3] * z * (2 * zz - 3 * xx - 3 * yy) * sh[:, 12]
               + SH_C3[4] * x * (4 * zz - xx - yy) * sh[:, 13] + SH_C3[5] * z * (xx - yy) * sh[:, 14]
               + SH_C3[6] * x * (xx - 3 * yy) * sh[:, 15])
    return res


def render_dense(means3D, opacities, *, shs=None, colors_precomp=None, scales=None, rotations=None,
                 viewmatrix, projmatrix, campos, bg, W, H, tanfovx, tanfovy, sh_degree=3,
                 scale_modifier=1.0, quat_map=True, pixel_chunk=4096, pixel_range=None):
    """Returns dict(color [3,H,W], depth [1,H,W], alpha [1,H,W], cov_scale [3,H,W],
    cov_quat [4,H,W] (no gradient, as in the reference), radii [P], n_contrib [H,W]).
    With pixel_range=(first, last) only that run of row-major pixel indices is evaluated and the
    image outputs come back flattened ([3,n], [1,n], ...): bench.py's bounded CPU-baseline sample."""
    dt = means3D.dtype
    P = means3D.shape[0]
    view, proj = viewmatrix.to(dt), projmatrix.to(dt)
    campos, bg = campos.to(dt), bg.to(dt)
    focal_x, focal_y = W / (2.0 * tanfovx), H / (2.0 * tanfovy)
    gx, gy = (W + 15) // 16, (H + 15) // 16

    p_view = means3D @ view[:3, :3] + view[3, :3]
    depth = p_view[:, 2]
    p_hom = means3D @ proj[:3, :] + proj[3, :]
    p_w = 1.0 / (p_hom[:, 3] + 1e-7)
    ndc = p_hom[:, :2] * p_w[:, None]

    # 3D covariance: Sigma = R diag(s^2) R^T with the un-normalised quaternion (forward.cu:119-153)
    s = scale_modifier * scales
    r, x, y, z = rotations[:, 0], rotations[:, 1], rotations[:, 2], rotations[:, 3]
    R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                     2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                     2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], -1).view(P, 3, 3)
    Sigma = R @ torch.diag_embed(s * s) @ R.transpose(1, 2)

    # EWA projection (forward.cu:75-114)
    limx, limy = 1.3 * tanfovx, 1.3 * tanfovy
    tz = p_view[:, 2]
    tx = torch.clamp(p_view[:, 0] / tz, -limx, limx) * tz
    ty = torch.clamp(p_view[:, 1] / tz, -limy, limy) * tz
    zero = torch.zeros_like(tz)
    J = torch.stack([focal_x / tz, zero, -(focal_x * tx) / (tz * tz),
                     zero, focal_y / tz, -(focal_y * ty) / (tz * tz)], -1).view(P, 2, 3)
    Wv = view[:3, :3].transpose(0, 1)
    JW = J @ Wv
    cov = JW @ Sigma @ JW.transpose(1, 2)
    a = cov[:, 0, 0] + 0.3
    b = cov[:, 0, 1]
    c = cov[:, 1, 1] + 0.3
    det = a * c - b * b
    det_safe = torch.where(det == 0, torch.ones_like(det), det)
    conic = torch.stack([c / det_safe, -b / det_safe, a / det_safe], -1)
    with torch.no_grad():
        mid = 0.5 * (a + c)
        sq = torch.sqrt(torch.clamp(mid * mid - det, min=0.1))
        radius = torch.ceil(3.0 * torch.sqrt(torch.maximum(mid + sq, mid - sq)))
    px = ((ndc[:, 0] + 1.0) * W - 1.0) * 0.5
    py = ((ndc[:, 1] + 1.0) * H - 1.0) * 0.5
    with torch.no_grad():
        rx0 = torch.clamp(torch.trunc((px - radius) / 16), 0, gx)
        ry0 = torch.clamp(torch.trunc((py - radius) / 16), 0, gy)
        rx1 = torch.clamp(torch.trunc((px + radius + 15) / 16), 0, gx)
        ry1 = torch.clamp(torch.trunc((py + radius + 15) / 16), 0, gy)
        visible = (depth > 0.2) & (det != 0) & ((rx1 - rx0) * (ry1 - ry0) > 0)
        radii = torch.where(visible, radius, torch.zeros_like(radius)).to(torch.int32)

    if colors_precomp is not None:
        rgb = colors_precomp
    else:
        d = means3D - campos
        d = d / d.norm(dim=1, keepdim=True)
        rgb = torch.clamp_min(eval_sh(sh_degree, shs, d) + 0.5, 0.0)

    # depth order (stable: ties keep ascending index) over the visible Gaussians
    with torch.no_grad():
        vis_idx = torch.nonzero(visible).squeeze(1)
        order = vis_idx[torch.sort(depth[vis_idx], stable=True).indices]
    o_px, o_py, o_con = px[order], py[order], conic[order]
    o_op, o_rgb, o_dep = opacities.reshape(-1)[order], rgb[order], depth[order]
    o_scale = scales[order]
    o_rot = rotations[order]
    o_rect = torch.stack([rx0, ry0, rx1, ry1], -1)[order]
    V = order.numel()

    npix = W * H
    out_color = []
    out_depth, out_alpha, out_scale, out_ncontrib = [], [], [], []
    quat = torch.zeros((npix, 4), dtype=dt)
    quat[:, 0] = 1.0
    ys, xs = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")
    pix_x = xs.reshape(-1).to(dt)
    pix_y = ys.reshape(-1).to(dt)
    p_first, p_last = (0, npix) if pixel_range is None else pixel_range
    for s0 in range(p_first, p_last, pixel_chunk):
        s1 = min(s0 + pixel_chunk, p_last)
        fx, fy = pix_x[s0:s1, None], pix_y[s0:s1, None]
        n = fx.shape[0]
        with torch.no_grad():
            tix, tiy = torch.floor(fx / 16), torch.floor(fy / 16)
            member = (tix >= o_rect[None, :, 0]) & (tix < o_rect[None, :, 2]) & \
                     (tiy >= o_rect[None, :, 1]) & (tiy < o_rect[None, :, 3])
        dx = o_px[None, :] - fx
        dy = o_py[None, :] - fy
        power = -0.5 * (o_con[None, :, 0] * dx * dx + o_con[None, :, 2] * dy * dy) - o_con[None, :, 1] * dx * dy
        alpha_raw = o_op[None, :] * torch.exp(torch.clamp(power, max=0.0))
        alpha = alpha_raw + (torch.clamp(alpha_raw, max=0.99) - alpha_raw).detach()   # straight-through clamp
        with torch.no_grad():
            valid = member & (power <= 0) & (alpha >= 1.0 / 255.0)
        A = torch.where(valid, alpha, torch.zeros_like(alpha))
        om = 1.0 - A
        T_incl = torch.cumprod(om, dim=1)
        T_excl = torch.cat([torch.ones((n, 1), dtype=dt), T_incl[:, :-1]], 1)
        with torch.no_grad():
            stop = valid & (T_excl * om < 0.0001)
            stopped = torch.cumsum(stop.to(torch.int32), 1) > 0
            contrib = valid & ~stopped
        w = torch.where(contrib, A * T_excl, torch.zeros_like(A))
        T_final = torch.prod(torch.where(contrib, om, torch.ones_like(om)), dim=1)
        out_color.append(w @ o_rgb + T_final[:, None] * bg[None, :])
        out_depth.append(w @ o_dep)
        out_alpha.append(1.0 - T_final)
        out_scale.append(w @ o_scale)
        with torch.no_grad():
            pos = torch.cumsum(member.to(torch.int64), 1)            # ordinal inside the tile's list
            last = torch.where(contrib, pos, torch.zeros_like(pos)).amax(dim=1) if V > 0 else torch.zeros(n, dtype=torch.int64)
            out_ncontrib.append(last)
            if quat_map and V > 0:
                q = quat[s0:s0 + n]
                wd = w.detach()
                for j in range(V):
                    sel = torch.nonzero(contrib[:, j]).squeeze(1)
                    if sel.numel() == 0:
                        continue
                    aw = wd[sel, j]
                    g = o_rot[j].detach()
                    if float(g[0]) > 1.0 - 1.1920929e-07:
                        rq = torch.stack([1.0 * (1 - aw) + g[0] * aw, g[1] * aw, g[2] * aw, g[3] * aw], -1)
                    else:
                        ang = math.acos(float(g[0]))
                        s0_, s1_ = torch.sin((1 - aw) * ang), torch.sin(aw * ang)
                        rq = torch.stack([s0_ + s1_ * g[0], s1_ * g[1], s1_ * g[2], s1_ * g[3]], -1) / math.sin(ang)
                    p_ = q[sel]
                    q[sel] = torch.stack([
                        p_[:, 0] * rq[:, 0] - p_[:, 1] * rq[:, 1] - p_[:, 2] * rq[:, 2] - p_[:, 3] * rq[:, 3],
                        p_[:, 0] * rq[:, 1] + p_[:, 1] * rq[:, 0] + p_[:, 2] * rq[:, 3] - p_[:, 3] * rq[:, 2],
                        p_[:, 0] * rq[:, 2] + p_[:, 2] * rq[:, 0] + p_[:, 3] * rq[:, 1] - p_[:, 1] * rq[:, 3],
                        p_[:, 0] * rq[:, 3] + p_[:, 3] * rq[:, 0] + p_[:, 1] * rq[:, 2] - p_[:, 2] * rq[:, 1]], -1)
    qn = quat.norm(dim=1, keepdim=True)
    ident = torch.zeros_like(quat); ident[:, 0] = 1
    quat = torch.where(qn > 0, quat / qn.clamp_min(1e-300), ident)
    if pixel_range is not None:
        return dict(color=torch.cat(out_color, 0).t(), depth=torch.cat(out_depth, 0)[None],
                    alpha=torch.cat(out_alpha, 0)[None], cov_scale=torch.cat(out_scale, 0).t(),
                    cov_quat=quat[p_first:p_last].t(), radii=radii, n_contrib=torch.cat(out_ncontrib, 0))
    return dict(color=torch.cat(out_color, 0).t().reshape(3, H, W),
                depth=torch.cat(out_depth, 0).reshape(1, H, W),
                alpha=torch.cat(out_alpha, 0).reshape(1, H, W),
                cov_scale=torch.cat(out_scale, 0).t().reshape(3, H, W),
                cov_quat=quat.t().reshape(4, H, W), radii=radii,
                n_contrib=torch.cat(out_ncontrib, 0).reshape(H, W))
