"""FusedAdam: torch.optim.Adam's update for VEGS's parameter groups as ONE kernel launch
(vegs_adam_step in include/vegs_b200.h), optionally folding the view-parallel step's reduced
densification statistics into the running ones and clearing the gradients in the same pass.

It is a `torch.optim.Optimizer` with the same per-parameter state keys as torch's Adam (`step`,
`exp_avg`, `exp_avg_sq`) and the same `param_groups` (each with its own `lr` and `name`), so the
reference's optimiser surgery during densification - `replace_tensor_to_optimizer`,
`_prune_optimizer`, `cat_tensors_to_optimizer` (scene/gaussian_model.py:278-367), which edit
`optimizer.state[...]["exp_avg"]` and swap `group["params"][0]` - and its learning-rate schedule
(`update_learning_rate`, :174-180) work on it unchanged.  Replaces `gaussians.optimizer.step()` +
`optimizer.zero_grad()` of train.py:319-320 (torch.optim.Adam(l, lr=0.0, eps=1e-15),
scene/gaussian_model.py:168).  CUDA float32 parameters only; there is no CPU path.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence, Tuple

import torch

from . import _lib


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr: float = 1e-3, betas: Tuple[float, float] = (0.9, 0.999), eps: float = 1e-8):
        defaults = dict(lr=lr, betas=betas, eps=eps)
        super().__init__(params, defaults)
        _lib.load()

    def _segments(self):
        segs = []
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is None:
                    continue
                if not (p.is_cuda and p.dtype == torch.float32 and p.is_contiguous()):
                    raise RuntimeError("FusedAdam: parameters must be contiguous float32 CUDA tensors (no CPU path)")
                g = p.grad
                if not (g.is_contiguous() and g.dtype == torch.float32 and g.device == p.device):
                    raise RuntimeError("FusedAdam: gradients must be contiguous float32 tensors on the parameter's device")
                state = self.state[p]
                if len(state) == 0:
                    state["step"] = torch.tensor(0.0)
                    state["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    state["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                state["step"] += 1
                t = int(state["step"].item()) if torch.is_tensor(state["step"]) else int(state["step"])
                segs.append((p, g, state["exp_avg"], state["exp_avg_sq"], float(group["lr"]), group["betas"],
                             float(group["eps"]), t))
        return segs

    @torch.no_grad()
    def step(self, closure=None, *, zero_grad: bool = False,
             stat_updates: Optional[Sequence[Tuple[torch.Tensor, torch.Tensor, str]]] = None):
        """One Adam step for every parameter that has a gradient.

        zero_grad=True clears each `.grad` in the same pass (the tensors stay allocated: what the
        fused accumulation of the next step adds into).  stat_updates = [(running, step_value, "add" |
        "max"), ...] folds the step's reduced densification statistics into the running ones (and,
        with zero_grad, clears the step values)."""
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        lib = _lib.load()
        segs = self._segments()
        # launches: parameters that share (betas, eps, step, device) go together, up to the segment
        # limit; VEGS's six groups (+ the three statistic pairs) always fit one launch
        jobs = []      # [key, [(param, grad, exp_avg, exp_avg_sq, count, lr, kind), ...]]
        for (p, g, m, v, lr, betas, eps, t) in segs:
            key = (tuple(betas), eps, t, p.device)
            if not jobs or jobs[-1][0] != key or len(jobs[-1][1]) >= _lib.ADAM_MAX_SEGMENTS:
                jobs.append([key, []])
            jobs[-1][1].append((p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), p.numel(), lr, 0))
        for run, val, how in (stat_updates or []):
            if not (run.is_cuda and val.is_cuda and run.dtype == val.dtype == torch.float32
                    and run.is_contiguous() and val.is_contiguous() and run.numel() == val.numel()):
                raise RuntimeError("FusedAdam: statistic pairs must be matching contiguous float32 CUDA tensors")
            if how not in ("add", "max"):
                raise ValueError("FusedAdam: statistic update must be 'add' or 'max'")
            if not jobs or len(jobs[-1][1]) >= _lib.ADAM_MAX_SEGMENTS or jobs[-1][0][3] != run.device:
                jobs.append([(tuple(self.defaults["betas"]), self.defaults["eps"], 1, run.device), []])
            jobs[-1][1].append((run.data_ptr(), val.data_ptr(), None, None, run.numel(), 0.0, 1 if how == "add" else 2))
        for (betas, eps, t, dev), items in jobs:
            a = _lib.VegsAdamArgs()
            for n, (pp, gp, mp, vp, count, lr, kind) in enumerate(items):
                sg = a.segments[n]
                sg.param, sg.grad, sg.exp_avg, sg.exp_avg_sq = pp, gp, mp, vp
                sg.count, sg.lr, sg.kind = count, lr, kind
            a.num_segments = len(items)
            a.step, a.beta1, a.beta2, a.eps, a.zero_grad = t, float(betas[0]), float(betas[1]), float(eps), int(bool(zero_grad))
            with torch.cuda.device(dev):
                stream = torch.cuda.current_stream(dev)
                _lib.check(lib.vegs_adam_step(C.byref(a), C.c_void_p(stream.cuda_stream)), "vegs_adam_step")
        return loss
