"""The reference's CALLERS of the rasterizer, restated for the tests (test infrastructure).

`/root/reference` does not exist on the GPU box and its `gaussian_renderer` imports half of VEGS
(datasets, plyfile, LoRA ...), so the ~60 lines that matter are restated here against a stub model
and a stub camera.  The point of the tests that use this file is that these callers run UNCHANGED
against either rasterizer module: the reference's own extension (baseline/_ref) or
`vegs_b200.install()`'s drop-in, selected only by what `diff_gaussian_rasterization` resolves to.

Restated from (paths relative to the reference tree):
  render()                   gaussian_renderer/__init__.py:20-119
  GaussianModel getters      scene/gaussian_model.py:100-120 (activations :58-66)
  add_densification_stats    scene/gaussian_model.py:411-413, max_radii2D update train.py:300-301
  loss_normal_guidance       loss/normal_guidance.py:3-23, quaternion_to_matrix
                             utils/graphics_utils.py:204-246, cam_normal_to_world_normal :362-368
"""
from __future__ import annotations

import math
from types import SimpleNamespace

import torch


class StubGaussianModel:
    """The slice of scene/gaussian_model.py's GaussianModel the renderer and the densification
    bookkeeping touch: raw parameters as leaves, activated values through the getters."""

    def __init__(self, scene: dict, device="cuda", sh_degree: int = 3):
        d = torch.device(device)
        opac = scene["opacities"].to(d).clamp(1e-6, 1 - 1e-6)
        self._xyz = scene["means3D"].to(d).clone().requires_grad_(True)
        self._features_dc = scene["shs"][:, :1].to(d).clone().contiguous().requires_grad_(True)
        self._features_rest = scene["shs"][:, 1:].to(d).clone().contiguous().requires_grad_(True)
        self._scaling = torch.log(scene["scales"].to(d)).requires_grad_(True)
        # un-normalised on purpose: the getter normalises (scene/gaussian_model.py:105-106)
        self._rotation = (scene["rotations"].to(d) * 1.7).requires_grad_(True)
        self._opacity = torch.log(opac / (1 - opac)).requires_grad_(True)     # inverse_sigmoid
        self.active_sh_degree = sh_degree
        self.max_sh_degree = sh_degree
        P = self._xyz.shape[0]
        self.max_radii2D = torch.zeros((P,), device=d)
        self.xyz_gradient_accum = torch.zeros((P, 1), device=d)
        self.denom = torch.zeros((P, 1), device=d)

    def leaves(self):
        return dict(xyz=self._xyz, f_dc=self._features_dc, f_rest=self._features_rest,
                    opacity=self._opacity, scaling=self._scaling, rotation=self._rotation)

    @property
    def get_scaling(self):
        return torch.exp(self._scaling)

    @property
    def get_rotation(self):
        return torch.nn.functional.normalize(self._rotation)

    @property
    def get_xyz(self):
        return self._xyz

    @property
    def get_features(self):
        return torch.cat((self._features_dc, self._features_rest), dim=1)

    @property
    def get_opacity(self):
        return torch.sigmoid(self._opacity)

    def add_densification_stats(self, viewspace_point_tensor, update_filter):
        self.xyz_gradient_accum[update_filter] += torch.norm(
            viewspace_point_tensor.grad[update_filter, :2], dim=-1, keepdim=True)
        self.denom[update_filter] += 1


def stub_camera(cam, R_c2w=None, normal=None):
    """A viewpoint camera with the attribute names scene/cameras.py gives it."""
    return SimpleNamespace(
        FoVx=2 * math.atan(cam.tanfovx), FoVy=2 * math.atan(cam.tanfovy), image_height=cam.H,
        image_width=cam.W, world_view_transform=cam.viewmatrix.cuda(), full_proj_transform=cam.projmatrix.cuda(),
        camera_center=cam.campos.cuda(), R=R_c2w, original_normal=normal)


PIPE = SimpleNamespace(debug=False, compute_cov3D_python=False, convert_SHs_python=False)


def render(viewpoint_camera, pc, pipe, bg_color, scaling_modifier=1.0, override_color=None):
    """gaussian_renderer/__init__.py:20-119, with the module looked up the way VEGS does:
    `from diff_gaussian_rasterization import ...` (whatever that name resolves to)."""
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer

    screenspace_points = torch.zeros_like(pc.get_xyz, dtype=pc.get_xyz.dtype, requires_grad=True, device="cuda") + 0
    try:
        screenspace_points.retain_grad()
    except Exception:
        pass
    tanfovx = math.tan(viewpoint_camera.FoVx * 0.5)
    tanfovy = math.tan(viewpoint_camera.FoVy * 0.5)
    raster_settings = GaussianRasterizationSettings(
        image_height=int(viewpoint_camera.image_height), image_width=int(viewpoint_camera.image_width),
        tanfovx=tanfovx, tanfovy=tanfovy, bg=bg_color, scale_modifier=scaling_modifier,
        viewmatrix=viewpoint_camera.world_view_transform, projmatrix=viewpoint_camera.full_proj_transform,
        sh_degree=pc.active_sh_degree, campos=viewpoint_camera.camera_center, prefiltered=False,
        debug=pipe.debug)
    rasterizer = GaussianRasterizer(raster_settings=raster_settings)
    means3D = pc.get_xyz
    means2D = screenspace_points
    opacity = pc.get_opacity
    scales = pc.get_scaling
    rotations = pc.get_rotation
    shs, colors_precomp = (pc.get_features, None) if override_color is None else (None, override_color)
    rendered_image, depth_image, rendered_cov_quat, rendered_cov_scale, alpha, radii = rasterizer(
        means3D=means3D, means2D=means2D, shs=shs, colors_precomp=colors_precomp, opacities=opacity,
        scales=scales, rotations=rotations, cov3D_precomp=None)
    return {"render": rendered_image, "render_depth": depth_image, "render_cov_quat": rendered_cov_quat,
            "render_cov_scale": rendered_cov_scale, "alpha": alpha, "viewspace_points": screenspace_points,
            "visibility_filter": radii > 0, "radii": radii}


def quaternion_to_matrix(quaternions: torch.Tensor) -> torch.Tensor:
    """utils/graphics_utils.py:204-246 (real part first)."""
    r, i, j, k = torch.unbind(quaternions, -1)
    two_s = 2.0 / (quaternions * quaternions).sum(-1)
    o = torch.stack((1 - two_s * (j * j + k * k), two_s * (i * j - k * r), two_s * (i * k + j * r),
                     two_s * (i * j + k * r), 1 - two_s * (i * i + k * k), two_s * (j * k - i * r),
                     two_s * (i * k - j * r), two_s * (j * k + i * r), 1 - two_s * (i * i + j * j)), -1)
    return o.reshape(quaternions.shape[:-1] + (3, 3))


def loss_normal_guidance(viewpoint_cam, cov_quat, cov_scale):
    """loss/normal_guidance.py:3-23."""
    norm_pred = viewpoint_cam.original_normal
    cov_scale = cov_scale.permute(1, 2, 0).reshape(-1, 1, 3).contiguous()
    cov_quat = cov_quat.permute(1, 2, 0).reshape(-1, 4).contiguous()
    cov_rot = quaternion_to_matrix(cov_quat)
    cov_rs = cov_rot.detach() * cov_scale
    R = torch.as_tensor(viewpoint_cam.R, dtype=torch.float32, device=norm_pred.device)
    norm_pred_world = (R[None, None] @ norm_pred.permute(1, 2, 0).unsqueeze(-1)).squeeze(-1).permute(2, 0, 1)
    norm_pred_world = norm_pred_world.permute(1, 2, 0).reshape(-1, 3).contiguous()
    norm_pred_world = norm_pred_world[:, :, None].repeat(1, 1, 3)
    return 0.8 * (cov_rot * norm_pred_world).sum(dim=-2).abs().mean() + \
        0.2 * (cov_rs * norm_pred_world).sum(dim=-2).abs().mean()


def training_iteration(pc, cam_ns, bg, gt_image, lambda_dnormal=0.05):
    """The rasterizer-facing part of one VEGS iteration (train.py:150-168, :292-301): render, L1 +
    normal-guidance loss, backward, densification bookkeeping."""
    pkg = render(cam_ns, pc, PIPE, bg)
    image = pkg["render"]
    loss = (image - gt_image).abs().mean()
    loss = loss + lambda_dnormal * loss_normal_guidance(cam_ns, pkg["render_cov_quat"], pkg["render_cov_scale"])
    loss.backward()
    vis, radii = pkg["visibility_filter"], pkg["radii"]
    pc.max_radii2D[vis] = torch.max(pc.max_radii2D[vis], radii[vis])
    pc.add_densification_stats(pkg["viewspace_points"], vis)
    return pkg, loss
