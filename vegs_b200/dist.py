"""View-parallel training step: training views are sharded across ranks, every rank holds the full
Gaussian set, and one all-reduce per optimisation step combines the parameter gradients and the
densification statistics (the reference is single-GPU and has no counterpart).

What is exchanged, per step:
  * SUM over ranks of one flat fp32 buffer  [grads of xyz(3) sh(3M) opacity(1) scale(3) rot(4),
    xyz_gradient_accum(1), denom(1)]  = (13 + 3M) words per Gaussian (61 at M = 16)
  * MAX over ranks of max_radii2D[P]
The per-view densification statistics follow the reference exactly
(train.py:300-301, scene/gaussian_model.py:411-413): the gradient norm is taken per view, before
the reduction, because it is not linear.

Works with any torch.distributed backend: NCCL over NVLink on the GPU box, gloo in the CPU tests.
"""
from __future__ import annotations

from typing import Dict, List, Sequence

import torch
import torch.distributed as dist

PARAM_ORDER = ("means3D", "shs", "opacities", "scales", "rotations")


def shard_views(n_views: int, rank: int, world: int) -> List[int]:
    """Round-robin: rank r renders views r, r + world, ..."""
    return list(range(rank, n_views, world))


class DensificationStats:
    """xyz_gradient_accum, denom, max_radii2D of scene/gaussian_model.py:154-172."""

    def __init__(self, P: int, device):
        self.xyz_gradient_accum = torch.zeros((P, 1), device=device)
        self.denom = torch.zeros((P, 1), device=device)
        self.max_radii2D = torch.zeros((P,), device=device)

    def attach(self, rasterizer_module) -> bool:
        """Let the backward kernel of `rasterizer_module` update these buffers itself
        (vegs_b200's set_densification_accumulators); False if the module cannot (the reference)."""
        reg = getattr(rasterizer_module, "set_densification_accumulators", None)
        self.fused = reg is not None
        if self.fused:
            reg(self.xyz_gradient_accum, self.denom, self.max_radii2D)
        return self.fused

    def detach(self, rasterizer_module) -> None:
        if getattr(self, "fused", False):
            rasterizer_module.set_densification_accumulators()
            self.fused = False

    def add_view(self, viewspace_grad: torch.Tensor, radii: torch.Tensor) -> None:
        """train.py:300-301 + add_densification_stats for one rendered view (no-op when the
        backward kernel already did it, see attach)."""
        if getattr(self, "fused", False):
            return
        vis = radii > 0
        self.max_radii2D = torch.where(vis, torch.maximum(self.max_radii2D, radii.to(self.max_radii2D.dtype)),
                                       self.max_radii2D)
        norm = torch.norm(viewspace_grad[:, :2], dim=-1, keepdim=True)
        self.xyz_gradient_accum += torch.where(vis.unsqueeze(1), norm, torch.zeros_like(norm))
        self.denom += vis.unsqueeze(1).to(self.denom.dtype)


def flat_size(params: Dict[str, torch.Tensor]) -> int:
    P = params["means3D"].shape[0]
    return sum(params[k].numel() for k in PARAM_ORDER) + 2 * P


def pack(params: Dict[str, torch.Tensor], stats: DensificationStats, out: torch.Tensor = None) -> torch.Tensor:
    """Gradients of the five parameter tensors + the two additive statistics as one flat buffer."""
    parts = []
    for k in PARAM_ORDER:
        g = params[k].grad
        parts.append((g if g is not None else torch.zeros_like(params[k])).reshape(-1))
    parts += [stats.xyz_gradient_accum.reshape(-1), stats.denom.reshape(-1)]
    if out is None:
        return torch.cat(parts)
    torch.cat(parts, out=out)
    return out


def unpack(flat: torch.Tensor, params: Dict[str, torch.Tensor], stats: DensificationStats) -> None:
    """Hand the reduced values back as views of the flat buffer (no copy)."""
    off = 0
    for k in PARAM_ORDER:
        n = params[k].numel()
        params[k].grad = flat[off:off + n].view_as(params[k])
        off += n
    P = params["means3D"].shape[0]
    stats.xyz_gradient_accum = flat[off:off + P].view(P, 1); off += P
    stats.denom = flat[off:off + P].view(P, 1)


def allreduce_step(params: Dict[str, torch.Tensor], stats: DensificationStats, group=None,
                   flat: torch.Tensor = None) -> torch.Tensor:
    """One SUM all-reduce of the flat buffer and one MAX all-reduce of max_radii2D.  Returns the
    flat buffer (so the caller can reuse it next step)."""
    flat = pack(params, stats, flat)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
        dist.all_reduce(stats.max_radii2D, op=dist.ReduceOp.MAX, group=group)
    unpack(flat, params, stats)
    return flat
