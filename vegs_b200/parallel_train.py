"""View-parallel training driver (the reference trains on one GPU, train.py:112-284).

Every rank holds the full Gaussian set.  An optimisation step takes a batch of `views_per_step`
training views, shards them round-robin over the ranks (vegs_b200.dist.shard_views), renders and
back-propagates the local ones, all-reduces the gradients and the step's densification statistics
through one flat buffer (vegs_b200.dist.FlatGradBuffer), and then runs the SAME optimiser step,
densification, pruning and opacity reset on every rank: those are deterministic functions of
identical inputs (the random split positions are drawn from a generator re-seeded identically on
every rank), so the replicas stay identical without ever exchanging parameters.  That is checked,
not assumed: after each densification the ranks compare Gaussian count and a checksum of the
parameters.  Rank 0 writes the checkpoints.

The driver is rasterizer- and model-agnostic:
  model       the reference's GaussianModel interface (scene/gaussian_model.py) - the reference's
              own class or vegs_b200.splat_model.SplatModel;
  render_view callable(model, camera, rasterizer_kwargs) -> (loss, viewspace_points, radii): one
              view's forward and loss (what train.py:150-168 does with gaussian_renderer.render);
              the driver calls loss.backward().  `rasterizer_kwargs` are the per-instance extension
              arguments for `GaussianRasterizer(settings, **kwargs)`: {} for the reference's module.
Works with any torch.distributed backend (NCCL on the GPU box, gloo in the CPU tests) and without
one (single process).
"""
from __future__ import annotations

import os
import random
from typing import Callable, List, Optional, Sequence

import torch
import torch.distributed as dist

from . import dist as vdist

_LEAF_TO_FLAT = dict(xyz="means3D", f_dc="shs_dc", f_rest="shs_rest", opacity="opacities", scaling="scales",
                     rotation="rotations")


def _world(group=None):
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


class ViewParallelTrainer:
    def __init__(self, model, cameras: Sequence, render_view: Callable, args, *, views_per_step: int = 8,
                 rasterizer_module=None, cameras_extent: float = 1.0, group=None, seed: int = 0,
                 checkpoint_dir: Optional[str] = None, checkpoint_iterations: Sequence[int] = (),
                 white_background: bool = False, camera_center: Optional[Callable] = None):
        """`camera_center(camera) -> [3] tensor` (optional, with a rasterizer module that has the
        extensions): switches the step to the deferred SH gradient - every view leaves dL/dRGB [P,3],
        the ranks all-gather those instead of all-reducing the SH gradient (FlatGradBuffer)."""
        self.model, self.cameras, self.render_view, self.args = model, list(cameras), render_view, args
        self.views_per_step = views_per_step
        self.module = rasterizer_module
        self.extent = cameras_extent
        self.group = group
        self.rank, self.world = _world(group)
        self.seed = seed
        self.ckpt_dir, self.ckpt_iters = checkpoint_dir, set(checkpoint_iterations)
        self.white_background = white_background
        self.camera_center = camera_center
        self._order_rng = random.Random(seed)        # the same camera order on every rank
        self._stack: List[int] = []
        self._buf = None
        self._buf_ids = None
        self.iteration = 0
        self.log: List[dict] = []

    # ------------------------------------------------------------------ batches
    def _next_batch(self) -> List[int]:
        """train.py:131-134 pops random cameras off a stack that is refilled when empty; here the
        same, `views_per_step` at a time, from a generator every rank seeds identically."""
        out = []
        for _ in range(self.views_per_step):
            if not self._stack:
                self._stack = list(range(len(self.cameras)))
            out.append(self._stack.pop(self._order_rng.randint(0, len(self._stack) - 1)))
        return out

    def _flat_params(self):
        m = self.model
        p = dict(xyz=m._xyz, f_dc=m._features_dc, f_rest=m._features_rest, opacity=m._opacity, scaling=m._scaling,
                 rotation=m._rotation)
        return {_LEAF_TO_FLAT[k]: v for k, v in p.items()}

    # ------------------------------------------------------------------ one optimisation step
    def step(self) -> dict:
        self.iteration += 1
        it, m, a = self.iteration, self.model, self.args
        m.update_learning_rate(it)
        params = self._flat_params()
        ids = tuple(id(v) for v in params.values())
        if self._buf is None or self._buf_ids != ids:       # first step, or densification replaced the tensors
            defer = ("shs_dc", "shs_rest") if (self.camera_center is not None and self.module is not None) else ()
            self._buf = vdist.FlatGradBuffer(params, order=tuple(params.keys()), defer_sh=defer,
                                             views_per_step=self.views_per_step)
            self._buf_ids = ids
        buf = self._buf
        batch = self._next_batch()
        buf.sh_degree = int(getattr(m, "active_sh_degree", 3))
        buf.begin_step([self.camera_center(self.cameras[v]) for v in batch] if buf.defer_names else None)
        mine = [batch[i] for i in vdist.shard_views(len(batch), self.rank, self.world)]
        module = self.module if self.module is not None else object
        loss_sum = 0.0
        for i, v in enumerate(mine):
            kw = buf.rasterizer_kwargs(module, first=(i == 0), slot=i)
            loss, viewspace, radii = self.render_view(m, self.cameras[v], kw)
            loss.backward()
            buf.step_stats.add_view(viewspace.grad, radii)
            loss_sum += float(loss.detach())
        if not mine:
            buf.zero_grads()                                 # a rank without a view contributes zeros
        running = vdist.DensificationStats.wrap(m.xyz_gradient_accum, m.denom, m.max_radii2D)
        densifying = it < a.densify_until_iter
        buf.allreduce(group=self.group, accumulate_into=running if densifying else None)
        info = dict(iteration=it, views=batch, local_views=mine, loss=loss_sum, P=int(params["means3D"].shape[0]))
        # ---- densification (train.py:292-316), identically on every rank
        if densifying:
            if it > a.densify_from_iter and it % a.densification_interval == 0:
                size_threshold = 20 if it > a.opacity_reset_interval else None
                self._seed_all(it)
                m.densify_and_prune(a.densify_grad_threshold, 0.005, self.extent, size_threshold)
                info["densified_to"] = self._check_replicas()
            if it % a.opacity_reset_interval == 0 or (self.white_background and it == a.densify_from_iter):
                m.reset_opacity()
        # ---- optimiser step (train.py:318-320); tensors replaced by densification have no .grad yet
        m.optimizer.step()
        m.optimizer.zero_grad(set_to_none=True)
        if it in self.ckpt_iters and self.rank == 0 and self.ckpt_dir:
            os.makedirs(self.ckpt_dir, exist_ok=True)
            torch.save((m.capture(), it), os.path.join(self.ckpt_dir, f"chkpnt{it}.pth"))
            info["checkpoint"] = True
        self.log.append(info)
        return info

    def _seed_all(self, it: int) -> None:
        torch.manual_seed(self.seed * 1_000_003 + it)
        if torch.cuda.is_available():
            torch.cuda.manual_seed_all(self.seed * 1_000_003 + it)

    def _check_replicas(self) -> int:
        """Every rank must hold the same Gaussians after a densification: compare the count and a
        checksum of the parameters across ranks (MIN and MAX must agree)."""
        params = self._flat_params()
        P = params["means3D"].shape[0]
        sig = torch.stack([torch.tensor(float(P), device=params["means3D"].device, dtype=torch.float64)] +
                          [v.detach().double().sum() for v in params.values()])
        if self.world > 1:
            lo, hi = sig.clone(), sig.clone()
            dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=self.group)
            dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=self.group)
            if not torch.allclose(lo, hi, rtol=1e-9, atol=0.0) or lo[0] != hi[0]:
                raise RuntimeError(f"replicas diverged after densification at iteration {self.iteration}: "
                                   f"min {lo.tolist()} max {hi.tolist()}")
        return int(P)

    def run(self, iterations: int) -> List[dict]:
        for _ in range(iterations):
            self.step()
        return self.log
