"""SplatModel: the Gaussian set with its optimiser and its densify / prune bookkeeping, in the shape
the view-parallel driver (vegs_b200/parallel_train.py) needs.

It offers the part of the reference's `GaussianModel` interface that the training loop touches
(scene/gaussian_model.py): the raw parameters `_xyz, _features_dc, _features_rest, _opacity,
_scaling, _rotation`, the activated getters (:100-120), `training_setup` with the reference's six
parameter groups and learning rates (:154-172), `update_learning_rate` (:174-180),
`add_densification_stats` (:411-413), `densify_and_prune` = clone + split + prune (:369-409),
`reset_opacity` (:253-257) and `capture` / `restore` (:68-98).  The reference's own class can be
handed to the driver instead; this one exists because the reference's module imports its dataset
stack and cannot be loaded on a bare box, and because the driver's tests need a model that also
runs on CPU tensors (gloo, stand-in rasterizer).

The mechanics are organised differently from the reference: every per-Gaussian tensor (the six
parameters, their two Adam moments, the three statistics) is reached through one table, and the
two structural operations - keep a subset of rows, append rows - are applied to the whole table at
once.  Results are the reference's: same selection rules, same sampling of split positions, same
reset values.
"""
from __future__ import annotations

import math
from typing import Dict

import torch
import torch.nn as nn

PARAM_NAMES = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")


def inverse_sigmoid(x):
    return torch.log(x / (1 - x))


def expon_lr(lr_init, lr_final, lr_delay_steps=0, lr_delay_mult=1.0, max_steps=1000000):
    """utils/general_utils.py:get_expon_lr_func - log-linear interpolation with an optional warm-up."""
    def f(step):
        if step < 0 or (lr_init == 0.0 and lr_final == 0.0):
            return 0.0
        if lr_delay_steps > 0:
            delay = lr_delay_mult + (1 - lr_delay_mult) * math.sin(0.5 * math.pi * min(max(step / lr_delay_steps, 0), 1))
        else:
            delay = 1.0
        t = min(max(step / max_steps, 0), 1)
        return delay * math.exp(math.log(lr_init) * (1 - t) + math.log(lr_final) * t)
    return f


def quat_to_rotmat(q: torch.Tensor) -> torch.Tensor:
    """utils/general_utils.py:build_rotation (normalises first; (r, x, y, z))."""
    q = q / q.norm(dim=1, keepdim=True)
    r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    return torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                        2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                        2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], -1).view(-1, 3, 3)


class TrainingArgs:
    """The optimisation parameters of arguments/__init__.py:OptimizationParams that this path uses."""
    position_lr_init = 0.00016
    position_lr_final = 0.0000016
    position_lr_delay_mult = 0.01
    position_lr_max_steps = 30_000
    feature_lr = 0.0025
    opacity_lr = 0.05
    scaling_lr = 0.005
    rotation_lr = 0.001
    percent_dense = 0.01
    densification_interval = 100
    opacity_reset_interval = 3000
    densify_from_iter = 500
    densify_until_iter = 15_000
    densify_grad_threshold = 0.0002

    def __init__(self, **kw):
        for k, v in kw.items():
            if not hasattr(type(self), k):
                raise AttributeError(k)
            setattr(self, k, v)


class SplatModel:
    def __init__(self, scene: Dict[str, torch.Tensor], device="cuda", sh_degree: int = 3, spatial_lr_scale: float = 1.0):
        """`scene`: activated values as vegs_b200.synth.make_scene returns them (means3D, scales,
        rotations, opacities, shs [P,M,3]); they are stored through the inverse activations."""
        d = torch.device(device)
        op = scene["opacities"].to(d).clamp(1e-6, 1 - 1e-6)
        self._xyz = nn.Parameter(scene["means3D"].to(d).clone())
        self._features_dc = nn.Parameter(scene["shs"][:, :1].to(d).clone().contiguous())
        self._features_rest = nn.Parameter(scene["shs"][:, 1:].to(d).clone().contiguous())
        self._opacity = nn.Parameter(inverse_sigmoid(op))
        self._scaling = nn.Parameter(torch.log(scene["scales"].to(d)))
        self._rotation = nn.Parameter(scene["rotations"].to(d).clone())
        self.max_sh_degree = sh_degree
        self.active_sh_degree = sh_degree
        self.spatial_lr_scale = spatial_lr_scale
        self.optimizer = None
        self.percent_dense = 0.01
        self._reset_stats()

    # ------------------------------------------------------------------ the reference's getters
    get_xyz = property(lambda self: self._xyz)
    get_scaling = property(lambda self: torch.exp(self._scaling))
    get_rotation = property(lambda self: torch.nn.functional.normalize(self._rotation))
    get_opacity = property(lambda self: torch.sigmoid(self._opacity))
    get_features = property(lambda self: torch.cat((self._features_dc, self._features_rest), dim=1))

    @property
    def device(self):
        return self._xyz.device

    @property
    def num_points(self) -> int:
        return int(self._xyz.shape[0])

    def parameters(self) -> Dict[str, nn.Parameter]:
        return dict(xyz=self._xyz, f_dc=self._features_dc, f_rest=self._features_rest, opacity=self._opacity,
                    scaling=self._scaling, rotation=self._rotation)

    def _bind(self, p: Dict[str, nn.Parameter]) -> None:
        self._xyz, self._features_dc, self._features_rest = p["xyz"], p["f_dc"], p["f_rest"]
        self._opacity, self._scaling, self._rotation = p["opacity"], p["scaling"], p["rotation"]

    def _reset_stats(self) -> None:
        P, d = self.num_points, self.device
        self.xyz_gradient_accum = torch.zeros((P, 1), device=d)
        self.denom = torch.zeros((P, 1), device=d)
        self.max_radii2D = torch.zeros((P,), device=d)

    # ------------------------------------------------------------------ optimiser
    def training_setup(self, args: TrainingArgs, optimizer_cls=None) -> None:
        """Six parameter groups with the reference's learning rates (scene/gaussian_model.py:154-172);
        `optimizer_cls` defaults to torch.optim.Adam, pass vegs_b200.optim.FusedAdam on a GPU."""
        self.percent_dense = args.percent_dense
        self._reset_stats()
        p = self.parameters()
        lrs = dict(xyz=args.position_lr_init * self.spatial_lr_scale, f_dc=args.feature_lr, f_rest=args.feature_lr / 20.0,
                   opacity=args.opacity_lr, scaling=args.scaling_lr, rotation=args.rotation_lr)
        groups = [{"params": [p[n]], "lr": lrs[n], "name": n} for n in PARAM_NAMES]
        self.optimizer = (optimizer_cls or torch.optim.Adam)(groups, lr=0.0, eps=1e-15)
        self.xyz_scheduler_args = expon_lr(args.position_lr_init * self.spatial_lr_scale,
                                           args.position_lr_final * self.spatial_lr_scale,
                                           lr_delay_mult=args.position_lr_delay_mult,
                                           max_steps=args.position_lr_max_steps)

    def update_learning_rate(self, iteration: int) -> float:
        for g in self.optimizer.param_groups:
            if g["name"] == "xyz":
                g["lr"] = self.xyz_scheduler_args(iteration)
                return g["lr"]

    # ------------------------------------------------------------------ densification statistics
    def add_densification_stats(self, viewspace_point_tensor, update_filter) -> None:
        self.xyz_gradient_accum[update_filter] += torch.norm(viewspace_point_tensor.grad[update_filter, :2], dim=-1,
                                                             keepdim=True)
        self.denom[update_filter] += 1

    # ------------------------------------------------------------------ the two structural operations
    def _rebuild(self, transform) -> None:
        """Apply `transform(name, tensor, is_moment) -> tensor` to every parameter and to its Adam
        moments, swap the new parameters into the optimiser's groups (same state keys as torch's
        Adam), rebind the model's attributes."""
        new = {}
        for g in self.optimizer.param_groups:
            old = g["params"][0]
            state = self.optimizer.state.pop(old, None)
            fresh = nn.Parameter(transform(g["name"], old.detach(), False).requires_grad_(True))
            if state is not None and "exp_avg" in state:
                state["exp_avg"] = transform(g["name"], state["exp_avg"], True)
                state["exp_avg_sq"] = transform(g["name"], state["exp_avg_sq"], True)
                self.optimizer.state[fresh] = state
            g["params"][0] = fresh
            new[g["name"]] = fresh
        self._bind(new)

    def keep_rows(self, keep: torch.Tensor) -> None:
        """Drop the Gaussians where `keep` is False (prune_points, :309-325)."""
        self._rebuild(lambda name, t, is_moment: t[keep])
        self.xyz_gradient_accum = self.xyz_gradient_accum[keep]
        self.denom = self.denom[keep]
        self.max_radii2D = self.max_radii2D[keep]

    def append_rows(self, extra: Dict[str, torch.Tensor]) -> None:
        """Append new Gaussians; their moments start at zero and every statistic is cleared
        (densification_postfix, :349-367)."""
        self._rebuild(lambda name, t, is_moment: torch.cat((t, torch.zeros_like(extra[name]) if is_moment else extra[name]), 0))
        self._reset_stats()

    # ------------------------------------------------------------------ densify / prune / reset
    def densify_and_prune(self, max_grad, min_opacity, extent, max_screen_size, prune=True) -> None:
        grads = self.xyz_gradient_accum / self.denom
        grads[grads.isnan()] = 0.0
        big_limit = self.percent_dense * extent
        # clone: small Gaussians with a large screen-space gradient are duplicated in place (:385-398)
        sel = (torch.norm(grads, dim=-1) >= max_grad) & (self.get_scaling.max(dim=1).values <= big_limit)
        self.append_rows({n: t.detach()[sel] for n, t in self.parameters().items()})
        # split: large ones are replaced by N = 2 samples drawn inside them, 1.6 times smaller (:369-383)
        N = 2
        n_before = grads.shape[0]
        padded = torch.zeros((self.num_points,), device=self.device)
        padded[:n_before] = grads.squeeze(-1)
        sel = (padded >= max_grad) & (self.get_scaling.max(dim=1).values > big_limit)
        stds = self.get_scaling.detach()[sel].repeat(N, 1)
        samples = torch.normal(mean=torch.zeros_like(stds), std=stds)
        rots = quat_to_rotmat(self._rotation.detach()[sel]).repeat(N, 1, 1)
        p = self.parameters()
        extra = {n: p[n].detach()[sel].repeat(*([N] + [1] * (p[n].dim() - 1))) for n in PARAM_NAMES}
        extra["xyz"] = torch.bmm(rots, samples.unsqueeze(-1)).squeeze(-1) + self._xyz.detach()[sel].repeat(N, 1)
        extra["scaling"] = torch.log(self.get_scaling.detach()[sel].repeat(N, 1) / (0.8 * N))
        self.append_rows(extra)
        gone = torch.cat((sel, torch.zeros(N * int(sel.sum()), device=self.device, dtype=torch.bool)))
        self.keep_rows(~gone)
        # prune: transparent, and (once sizes are trusted) screen-filling or world-filling Gaussians (:400-407)
        drop = (self.get_opacity < min_opacity).squeeze(-1)
        if prune:
            if max_screen_size:
                drop = drop | (self.max_radii2D > max_screen_size) | (self.get_scaling.max(dim=1).values > 0.1 * extent)
            self.keep_rows(~drop)

    def reset_opacity(self) -> None:
        """Opacities capped at 0.01, their Adam moments cleared (:253-257 + replace_tensor_to_optimizer)."""
        new_raw = inverse_sigmoid(torch.min(self.get_opacity.detach(), torch.ones_like(self._opacity) * 0.01))
        self._rebuild(lambda name, t, is_moment: (torch.zeros_like(t) if is_moment else new_raw) if name == "opacity" else t)

    # ------------------------------------------------------------------ checkpoint
    def capture(self) -> dict:
        return dict(active_sh_degree=self.active_sh_degree, params={n: t.detach().clone() for n, t in self.parameters().items()},
                    max_radii2D=self.max_radii2D.clone(), xyz_gradient_accum=self.xyz_gradient_accum.clone(),
                    denom=self.denom.clone(), optimizer=self.optimizer.state_dict() if self.optimizer else None,
                    spatial_lr_scale=self.spatial_lr_scale)

    def restore(self, ckpt: dict, args: TrainingArgs, optimizer_cls=None) -> None:
        self.active_sh_degree = ckpt["active_sh_degree"]
        self.spatial_lr_scale = ckpt["spatial_lr_scale"]
        self._bind({n: nn.Parameter(t.to(self.device).clone()) for n, t in ckpt["params"].items()})
        self.training_setup(args, optimizer_cls)
        self.max_radii2D = ckpt["max_radii2D"].to(self.device)
        self.xyz_gradient_accum = ckpt["xyz_gradient_accum"].to(self.device)
        self.denom = ckpt["denom"].to(self.device)
        if ckpt.get("optimizer") is not None:
            self.optimizer.load_state_dict(ckpt["optimizer"])
