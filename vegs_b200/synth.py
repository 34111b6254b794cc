"""Synthetic LiDAR-density street scene and KITTI-360-like cameras.

Used by bench.py, the tests and the golden-vector generator.  There is no dataset on the box, so
the workload is generated: a 200 m street corridor (ground, two facades, far wall, clutter boxes)
sampled at LiDAR density, with Gaussians as VEGS leaves them after its normal-based initialisation
(flat discs `scales = (1e-5, 0.1, 0.1)` aligned with the surface normal, reference
utils/norminit_utils.py:217-219).  Cameras follow the reference's conventions exactly
(scene/cameras.py:70-88, utils/graphics_utils.py:266-344).
"""
from __future__ import annotations

import math
from typing import Dict, NamedTuple

import numpy as np
import torch

C0 = 0.28209479177387814

# KITTI-360 perspective camera, nominal intrinsics at 1408x376
KITTI360 = dict(W=1408, H=376, fx=552.554261, fy=552.554261, cx=682.049453, cy=238.769549)


class Cam(NamedTuple):
    W: int
    H: int
    tanfovx: float
    tanfovy: float
    viewmatrix: torch.Tensor   # [4,4] world_view_transform (transposed storage)
    projmatrix: torch.Tensor   # [4,4] full_proj_transform
    campos: torch.Tensor       # [3]


def _projection(znear, zfar, fovX, fovY, fx, fy, cx, cy, w, h) -> torch.Tensor:
    """utils/graphics_utils.py:305-337 (getProjectionMatrixwithPrincipalPointOffset)."""
    top_c = math.tan(fovY / 2) * znear
    right_c = math.tan(fovX / 2) * znear
    dx = (cx - w / 2) / fx * znear
    dy = (cy - h / 2) / fy * znear
    top, bottom = top_c + dy, -top_c + dy
    left, right = -right_c + dx, right_c + dx
    P = torch.zeros(4, 4)
    P[0, 0] = 2.0 * znear / (right - left)
    P[1, 1] = 2.0 * znear / (top - bottom)
    P[0, 2] = (right + left) / (right - left)
    P[1, 2] = (top + bottom) / (top - bottom)
    P[3, 2] = 1.0
    P[2, 2] = (zfar + znear) / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def rot_x(deg: float) -> np.ndarray:
    a = math.radians(deg)
    return np.array([[1, 0, 0], [0, math.cos(a), -math.sin(a)], [0, math.sin(a), math.cos(a)]])


def rot_z(deg: float) -> np.ndarray:
    a = math.radians(deg)
    return np.array([[math.cos(a), -math.sin(a), 0], [math.sin(a), math.cos(a), 0], [0, 0, 1]])


def make_camera(k: int = 0, scale: float = 1.0, add_xrot: float = 0.0, add_zrot: float = 0.0,
                add_tz: float = 0.0, device="cpu") -> Cam:
    """Camera k of the synthetic drive: 1.55 m above the centre line at x = 10 + 5k, looking +x.

    `scale` scales the KITTI-360 intrinsics (0.25 -> 352x94, 2 -> 2816x752).  add_xrot / add_zrot /
    add_tz build the extrapolated views the way scene/cameras.py:178-227 does: pitch applied in the
    camera frame (c2w @ R_x), yaw in the world frame (R_z @ c2w), then a height offset.
    """
    W, H = int(round(KITTI360["W"] * scale)), int(round(KITTI360["H"] * scale))
    fx, fy = KITTI360["fx"] * scale, KITTI360["fy"] * scale
    cx, cy = KITTI360["cx"] * scale, KITTI360["cy"] * scale
    # camera axes in world coordinates: x right = -y_w, y down = -z_w, z forward = +x_w
    R_c2w = np.array([[0.0, 0.0, 1.0], [-1.0, 0.0, 0.0], [0.0, -1.0, 0.0]])
    C = np.array([10.0 + 5.0 * k, 0.0, 1.55])
    if add_xrot:
        R_c2w = R_c2w @ rot_x(add_xrot)
    if add_zrot:
        R_c2w = rot_z(add_zrot) @ R_c2w      # rotation part only; the centre stays
    C = C + np.array([0.0, 0.0, add_tz])
    # getWorld2View2(R, t): Rt[:3,:3] = R^T, Rt[:3,3] = t  with R = c2w rotation, t = -R^T C
    Rt = np.eye(4)
    Rt[:3, :3] = R_c2w.T
    Rt[:3, 3] = -R_c2w.T @ C
    view = torch.tensor(np.float32(Rt)).transpose(0, 1)
    fovx = 2 * math.atan(W / (2 * fx))
    fovy = 2 * math.atan(H / (2 * fy))
    proj = _projection(0.01, 100.0, fovx, fovy, fx, fy, cx, cy, W, H).transpose(0, 1)
    full = (view.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0)
    campos = view.inverse()[3, :3]
    return Cam(W, H, math.tan(fovx * 0.5), math.tan(fovy * 0.5), view.contiguous().to(device),
               full.contiguous().to(device), campos.contiguous().to(device))


def _quat_from_normal(n: torch.Tensor) -> torch.Tensor:
    """Unit quaternion (r,x,y,z), r >= 0, rotating local +x onto the unit normal n."""
    ex = torch.tensor([1.0, 0.0, 0.0], dtype=n.dtype).expand_as(n)
    d = (ex * n).sum(-1, keepdim=True).clamp(-1.0, 1.0)
    axis = torch.linalg.cross(ex, n)
    an = axis.norm(dim=-1, keepdim=True)
    fallback = torch.tensor([0.0, 0.0, 1.0], dtype=n.dtype).expand_as(n)
    axis = torch.where(an > 1e-8, axis / an.clamp_min(1e-20), fallback)
    half = 0.5 * torch.acos(d)
    q = torch.cat([torch.cos(half), axis * torch.sin(half)], dim=-1)
    q = q / q.norm(dim=-1, keepdim=True)
    return torch.where(q[:, :1] < 0, -q, q)


def make_scene(P: int, seed: int = 7, sh_degree: int = 3, device="cpu") -> Dict[str, torch.Tensor]:
    """P Gaussians on the street corridor.  Returns float32 tensors:
    means3D [P,3], scales [P,3], rotations [P,4] (unit, r >= 0), opacities [P,1] (post-sigmoid),
    shs [P,16,3] (DC = RGB2SH(U(0,1)), higher orders N(0,0.05))."""
    g = torch.Generator().manual_seed(seed)
    n_ground = int(0.60 * P)
    n_fac = int(0.35 * P)
    n_clut = P - n_ground - n_fac
    rnd = lambda *s: torch.rand(*s, generator=g, dtype=torch.float64)
    rn = lambda *s: torch.randn(*s, generator=g, dtype=torch.float64)

    pts, nrm = [], []
    # ground z = 0
    p = torch.stack([rnd(n_ground) * 200.0, rnd(n_ground) * 40.0 - 20.0, torch.zeros(n_ground, dtype=torch.float64)], -1)
    pts.append(p); nrm.append(torch.tensor([0.0, 0.0, 1.0], dtype=torch.float64).expand(n_ground, 3))
    # facades y = +-12 (85 % of the facade budget) and the far wall x = 200 (15 %)
    n_wall = int(0.15 * n_fac)
    n_side = n_fac - n_wall
    side = torch.where(rnd(n_side) < 0.5, -1.0, 1.0).to(torch.float64)
    p = torch.stack([rnd(n_side) * 200.0, 12.0 * side, rnd(n_side) * 10.0], -1)
    pts.append(p); nrm.append(torch.stack([torch.zeros_like(side), -side, torch.zeros_like(side)], -1))
    p = torch.stack([torch.full((n_wall,), 200.0, dtype=torch.float64), rnd(n_wall) * 24.0 - 12.0, rnd(n_wall) * 10.0], -1)
    pts.append(p); nrm.append(torch.tensor([-1.0, 0.0, 0.0], dtype=torch.float64).expand(n_wall, 3))
    # clutter: 40 boxes of 2 x 4 x 1.6 m on the road shoulder, points on their faces
    nb = 40
    centers = torch.stack([rnd(nb) * 180.0 + 15.0, (rnd(nb) * 2 - 1) * 8.0, torch.full((nb,), 0.8, dtype=torch.float64)], -1)
    half = torch.tensor([2.0, 1.0, 0.8], dtype=torch.float64)
    bi = torch.randint(0, nb, (n_clut,), generator=g)
    face = torch.randint(0, 5, (n_clut,), generator=g)          # +-x, +-y, +z
    axis = torch.tensor([0, 0, 1, 1, 2])[face]
    sign = torch.tensor([-1.0, 1.0, -1.0, 1.0, 1.0], dtype=torch.float64)[face]
    u = (rnd(n_clut, 3) * 2 - 1) * half
    u[torch.arange(n_clut), axis] = sign * half[axis]
    n = torch.zeros(n_clut, 3, dtype=torch.float64)
    n[torch.arange(n_clut), axis] = sign
    pts.append(centers[bi] + u); nrm.append(n)

    pts = torch.cat(pts, 0)
    nrm = torch.cat(nrm, 0)
    pts = pts + nrm * (rn(P, 1) * 0.02)                          # 2 cm jitter along the normal
    perm = torch.randperm(P, generator=g)                         # LiDAR points arrive unordered
    pts, nrm = pts[perm], nrm[perm]

    scales = torch.tensor([1e-5, 0.1, 0.1], dtype=torch.float64) * torch.exp(rn(P, 3) * 0.3)
    rot = _quat_from_normal(nrm)
    opac = torch.sigmoid(rn(P, 1) * 2.0)
    M = (sh_degree + 1) ** 2
    shs = rn(P, M, 3) * 0.05
    shs[:, 0, :] = (rnd(P, 3) - 0.5) / C0
    out = dict(means3D=pts, scales=scales, rotations=rot, opacities=opac, shs=shs)
    return {k: v.to(torch.float32).contiguous().to(device) for k, v in out.items()}


def cotangents(W: int, H: int, seed: int = 11, device="cpu") -> Dict[str, torch.Tensor]:
    """Fixed random cotangents N(0,1) on the five image outputs (config 1: every backward input
    is exercised)."""
    g = torch.Generator().manual_seed(seed)
    shapes = dict(color=(3, H, W), depth=(1, H, W), cov_quat=(4, H, W), cov_scale=(3, H, W), alpha=(1, H, W))
    return {k: torch.randn(*s, generator=g, dtype=torch.float32).to(device) for k, s in shapes.items()}
