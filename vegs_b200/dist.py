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
    """xyz_gradient_accum, denom, max_radii2D of scene/gaussian_model.py:154-172.

    The three tensors keep their identity for the lifetime of the object (reductions copy into them
    in place), so a rasterizer that was handed them through `rasterizer_kwargs` keeps updating the
    tensors this object reads."""

    def __init__(self, P: int, device):
        self.xyz_gradient_accum = torch.zeros((P, 1), device=device)
        self.denom = torch.zeros((P, 1), device=device)
        self.max_radii2D = torch.zeros((P,), device=device)
        self.fused = False

    @classmethod
    def wrap(cls, xyz_gradient_accum, denom, max_radii2D) -> "DensificationStats":
        """A view of three existing tensors (e.g. a GaussianModel's running statistics)."""
        s = cls.__new__(cls)
        s.xyz_gradient_accum, s.denom, s.max_radii2D, s.fused = xyz_gradient_accum, denom, max_radii2D, False
        return s

    def zero_(self) -> None:
        self.xyz_gradient_accum.zero_(); self.denom.zero_(); self.max_radii2D.zero_()

    def rasterizer_kwargs(self, rasterizer_module) -> dict:
        """Keyword arguments for `rasterizer_module.GaussianRasterizer(settings, **kwargs)` that let
        the backward kernel of that instance update these buffers itself (vegs_b200's per-instance
        `densification_accumulators`); {} for a module without the extension (the reference), in
        which case `add_view` does the work.  Sets `self.fused` accordingly."""
        self.fused = hasattr(rasterizer_module, "RasterizerOptions")
        if self.fused:
            return dict(densification_accumulators=(self.xyz_gradient_accum, self.denom, self.max_radii2D))
        return {}

    def add_view(self, viewspace_grad: torch.Tensor, radii: torch.Tensor) -> None:
        """train.py:300-301 + add_densification_stats for one rendered view (no-op when the
        backward kernel already did it, see rasterizer_kwargs).  In place."""
        if self.fused:
            return
        vis = radii > 0
        self.max_radii2D.copy_(torch.where(vis, torch.maximum(self.max_radii2D, radii.to(self.max_radii2D.dtype)),
                                           self.max_radii2D))
        norm = torch.norm(viewspace_grad[:, :2], dim=-1, keepdim=True)
        self.xyz_gradient_accum += torch.where(vis.unsqueeze(1), norm, torch.zeros_like(norm))
        self.denom += vis.unsqueeze(1).to(self.denom.dtype)


def flat_size(params: Dict[str, torch.Tensor]) -> int:
    return _layout(params)[1]


def _order(params: Dict[str, torch.Tensor], order=None):
    if order is not None:
        return tuple(order)
    return PARAM_ORDER if all(k in params for k in PARAM_ORDER) and len(params) == len(PARAM_ORDER) else tuple(params.keys())


def _layout(params: Dict[str, torch.Tensor], order=None, tail=()):
    """[(name, offset, numel)] of the flat buffer: the five gradients, then the two additive
    statistics.  Every segment starts on a multiple of four floats, so that each `.grad` view is
    16-byte aligned (what the backward kernel's bulk stores / reductions need) whatever P is."""
    P = params["means3D"].shape[0]
    out, off = [], 0
    al = lambda n: (n + 3) // 4 * 4
    for k in _order(params, order):
        if k in tail:
            continue
        out.append((k, off, params[k].numel())); off = al(off + params[k].numel())
    out.append(("xyz_gradient_accum", off, P)); off = al(off + P)
    out.append(("denom", off, P)); off = al(off + P)
    for k in tail:                       # segments placed after the statistics (see FlatGradBuffer, deferred SH)
        out.append((k, off, params[k].numel())); off = al(off + params[k].numel())
    return out, off


def pack(params: Dict[str, torch.Tensor], stats: DensificationStats, out: torch.Tensor = None) -> torch.Tensor:
    """Gradients of the five parameter tensors + the two additive statistics as one flat buffer.
    Parts that already ARE the matching slice of `out` (a `.grad` bound by `unpack` /
    `FlatGradBuffer.bind` and accumulated into in place) are left where they are."""
    layout, total = _layout(params)
    if out is None:
        out = torch.empty(total, dtype=torch.float32, device=params["means3D"].device)
    assert out.numel() == total and out.is_contiguous()
    if total != sum(n for _, _, n in layout):
        # alignment padding (at most three words after a segment) must reduce as zeros
        ends = [off for _, off, _ in layout[1:]] + [total]
        for (_, off, n), nxt in zip(layout, ends):
            if off + n < nxt:
                out[off + n:nxt].zero_()
    for name, off, n in layout:
        if name in params:
            g = params[name].grad
        else:
            g = getattr(stats, name)
        dst = out[off:off + n]
        if g is None:
            dst.zero_()
        elif g.is_contiguous() and g.data_ptr() == dst.data_ptr():
            continue
        else:
            dst.copy_(g.reshape(-1))
    return out


def unpack(flat: torch.Tensor, params: Dict[str, torch.Tensor], stats: DensificationStats) -> None:
    """Hand the reduced gradients back as views of the flat buffer (no copy) and copy the reduced
    statistics into the stats tensors IN PLACE (they keep their identity: a rasterizer instance that
    was handed them keeps updating the tensors this object reads)."""
    layout, _ = _layout(params)
    for name, off, n in layout:
        if name in params:
            params[name].grad = flat[off:off + n].view_as(params[name])
        else:
            t = getattr(stats, name)
            src = flat[off:off + n].view_as(t)
            if t.data_ptr() != src.data_ptr():
                t.copy_(src)


SH_C0 = 0.28209479177387814      # the constant SH basis function (reference cuda_rasterizer/auxiliary.h:23)


class ShFactorCapture:
    """Keeps, for every backward pass of a step, the factor of that view's SH gradient.

    One view's dL/dSH is an outer product: dL/dsh[k][c] = basis_k(view direction) * dL/dRGB[c] (the
    colour is linear in the coefficients; clamped channels are already zero), and basis_0 is the
    constant SH_C0, so the DC row of the gradient a backward pass hands to the SH leaf IS that
    view's dL/dRGB up to that factor.  A tensor hook on the leaf (the `shs` parameter, or
    `features_dc` when the coefficients are split) copies `grad[:, 0, :] / SH_C0` of each pass into
    the next of `slots` rows; autograd's accumulation into `.grad` goes on untouched, so the caller
    and the rasterizer (the plain drop-in module) see nothing.  `allreduce_step(..., sh_factor=...)`
    then sends these rows instead of the SH gradient.

        cap = ShFactorCapture(params["shs"], slots=views_per_rank)
        per step:  cap.begin_step();  <views: forward, loss.backward()>;
                   allreduce_step(params, stats, sh_factor=cap, view_campos=<[V, 3], batch order>)
    """

    def __init__(self, sh_dc_param: torch.Tensor, slots: int):
        if sh_dc_param.dim() != 3 or sh_dc_param.shape[2] != 3:
            raise ValueError("the SH parameter must be [P, M, 3]")
        self.slots = int(slots)
        self.buf = torch.zeros(max(self.slots, 1), sh_dc_param.shape[0], 3, device=sh_dc_param.device,
                               dtype=torch.float32)
        self.n = 0
        self.enabled = False
        self._handle = sh_dc_param.register_hook(self._hook)

    def _hook(self, grad):
        if self.enabled:
            if self.n >= self.slots:
                raise RuntimeError(f"ShFactorCapture: more than {self.slots} backward passes in one step")
            torch.mul(grad[:, 0, :], 1.0 / SH_C0, out=self.buf[self.n])
            self.n += 1
        return None

    def begin_step(self) -> None:
        self.buf.zero_()              # a slot without a view reduces as "not rendered"
        self.n = 0
        self.enabled = True

    def end_step(self) -> None:
        self.enabled = False

    def remove(self) -> None:
        self._handle.remove()


def _exchange_sh_factor(params, sh_names, drgb_local, campos_all, sh_degree, group):
    """All-gather of every rank's dL/dRGB rows ([slots, P, 3] each) and the rebuild of the step's SH
    gradient from them (vegs_sh_grad_from_views, csrc/sh_views.cu) in place of this rank's own
    `.grad`: 3 words per view and Gaussian between the GPUs instead of 48 per Gaussian.
    `campos_all`: [world * slots, 3], row r * slots + i = camera centre of rank r's i-th view (rows of
    unused slots are never read: their dL/dRGB is zero).
    Returns the rebuild as a callable (the caller runs it after its other collectives)."""
    import ctypes as C
    from . import _lib
    lib = _lib.load()
    world = dist.get_world_size(group)
    dc = params[sh_names[0]].grad
    rest = params[sh_names[1]].grad if len(sh_names) > 1 else None
    means = params["means3D"].detach()
    P, dev = int(means.shape[0]), means.device
    M = int(dc.shape[1]) + (int(rest.shape[1]) if rest is not None else 0)
    slots = int(drgb_local.shape[0])
    campos_all = campos_all.to(device=dev, dtype=torch.float32).contiguous()
    if tuple(campos_all.shape) != (world * slots, 3) or tuple(drgb_local.shape) != (slots, P, 3):
        raise ValueError(f"SH factor exchange: camera centres must be [{world * slots}, 3] and the local factors "
                         f"[{slots}, {P}, 3], got {tuple(campos_all.shape)} and {tuple(drgb_local.shape)}")
    drgb_all = torch.empty(world * slots, P, 3, device=dev, dtype=torch.float32)
    dist.all_gather_into_tensor(drgb_all.view(-1), drgb_local.contiguous().view(-1), group=group)

    def rebuild():
        with torch.cuda.device(dev):
            rc = lib.vegs_sh_grad_from_views(P, int(sh_degree), M, world * slots, means.data_ptr(),
                                             campos_all.data_ptr(), drgb_all.data_ptr(), dc.data_ptr(),
                                             rest.data_ptr() if (rest is not None and rest.numel()) else None, 0,
                                             C.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
            _lib.check(rc, "vegs_sh_grad_from_views")
    return rebuild


def _slot_campos(view_campos, world: int, slots: int, device) -> torch.Tensor:
    """Camera centres of a step's views (batch order) in all-gather order: view j is rendered by
    rank j % world as its (j // world)-th view (shard_views), i.e. row (j % world) * slots + j // world."""
    vc = torch.as_tensor(view_campos) if not isinstance(view_campos, (list, tuple)) else torch.stack(
        [torch.as_tensor(c).reshape(3) for c in view_campos])
    vc = vc.to(device=device, dtype=torch.float32).reshape(-1, 3)
    V = int(vc.shape[0])
    if V > world * slots:
        raise ValueError(f"{V} views in the step but only {world} ranks x {slots} slots")
    j = torch.arange(V, device=device)
    out = torch.zeros(world * slots, 3, device=device, dtype=torch.float32)
    out[(j % world) * slots + j // world] = vc
    return out


def allreduce_step(params: Dict[str, torch.Tensor], stats: DensificationStats, group=None,
                   flat: torch.Tensor = None, accumulate_into: DensificationStats = None,
                   per_tensor: bool = None, single_view_campos: torch.Tensor = None,
                   sh_degree: int = 3, sh_factor: "ShFactorCapture" = None, view_campos=None) -> torch.Tensor:
    """The reduction of a step whose gradients live in autograd's own `.grad` tensors (the plain
    drop-in module, either arm of bench.py): SUM over ranks of the five gradients and the two
    additive statistics, MAX of max_radii2D.

    Two equivalent routes.  per_tensor=False packs everything into one flat buffer (one copy pass
    over all gradients), reduces it with one call and hands the slices back as the `.grad` tensors;
    it returns the buffer so the caller can reuse it next step.  per_tensor=True reduces the large
    tensors (the SH gradient) where they are and the small ones through one small staging buffer:
    no pack copy of the 1.15 GB SH gradient (0.35 ms at the 6 M scene), three collectives in all.
    Default: per tensor on NCCL, packed elsewhere (gloo pays per call).

    SH gradient as its factor (per-tensor route on CUDA; `sh_degree` = the active degree the views
    were rendered with; the result equals the plain all-reduce to rounding, tests/test_gpu_multi.py):
    * `sh_factor` (a ShFactorCapture that watched this step's backward passes) + `view_campos`
      (the camera centres of ALL the step's views, [V, 3] or a list, in batch order): the ranks
      all-gather each view's dL/dRGB and rebuild the SH gradient (_exchange_sh_factor): 13 + 2 + 3 V
      words per Gaussian between the GPUs instead of 61 + 2 for a V-view step;
    * `single_view_campos` ([world, 3], row r = camera centre of rank r's view): the same without a
      hook, on the promise that every rank rendered exactly ONE view this step - its own `.grad`
      then still is that one outer product and the factor is read off its DC row.

    `stats` must hold THIS STEP's contribution only (fresh or zeroed at the start of the step): the
    SUM over ranks of a running accumulator would count the earlier steps world-size times.  Pass
    the running accumulator as `accumulate_into` to have the reduced step added to it."""
    active = dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
    if per_tensor is None:
        per_tensor = active and dist.get_backend(group) == "nccl" and flat is None
    if (single_view_campos is not None or sh_factor is not None) and active and not per_tensor:
        raise ValueError("the SH factor exchange (single_view_campos / sh_factor) is a per-tensor-route option "
                         "(per_tensor=True, no `flat` buffer)")
    if per_tensor:
        if active:
            # the large tensors (the SH gradient is 48 of the 61 words per Gaussian) are reduced where
            # they lie; the small ones share one staging buffer, because every collective costs a
            # fixed ~70 us on top of its bytes (measured at 8 GPUs: 7 separate calls 4.23 ms, one
            # flat call 3.61 ms for the same 1.49 GB)
            names = _order(params)
            for k in names:
                if params[k].grad is None:
                    params[k].grad = torch.zeros_like(params[k])
            sh_names = []
            if single_view_campos is not None or sh_factor is not None:
                sh_names = ["shs"] if "shs" in params else [k for k in ("shs_dc", "shs_rest") if k in params]
                if not sh_names or not params["means3D"].is_cuda:
                    raise ValueError("the SH factor exchange needs an SH parameter ('shs' or 'shs_dc' + 'shs_rest') on a CUDA device")
                world = dist.get_world_size(group)
                if sh_factor is not None:
                    if view_campos is None:
                        raise ValueError("sh_factor needs view_campos (the camera centres of the step's views, batch order)")
                    sh_factor.end_step()
                    rebuild_sh = _exchange_sh_factor(params, sh_names, sh_factor.buf,
                                                     _slot_campos(view_campos, world, sh_factor.buf.shape[0],
                                                                  params["means3D"].device), sh_degree, group)
                else:
                    dc = params[sh_names[0]].grad
                    rebuild_sh = _exchange_sh_factor(params, sh_names, (dc[:, 0, :] * (1.0 / SH_C0)).unsqueeze(0),
                                                     single_view_campos, sh_degree, group)
            big = [k for k in names if params[k].numel() >= 8 * params["means3D"].numel() and k not in sh_names]
            small = [params[k].grad for k in names if k not in big and k not in sh_names]
            small += [stats.xyz_gradient_accum, stats.denom]
            for k in big:
                dist.all_reduce(params[k].grad, op=dist.ReduceOp.SUM, group=group)
            stage = torch.cat([t.reshape(-1) for t in small])
            dist.all_reduce(stage, op=dist.ReduceOp.SUM, group=group)
            off = 0
            for t in small:
                t.copy_(stage[off:off + t.numel()].view_as(t)); off += t.numel()
            dist.all_reduce(stats.max_radii2D, op=dist.ReduceOp.MAX, group=group)
            if sh_names:
                # (running this kernel underneath the collectives above - async collectives, kernel on
                # the current stream - was measured at 8 GPUs: 1010 against 1021 views/s, no gain)
                rebuild_sh()
    else:
        flat = pack(params, stats, flat)
        if active:
            dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
            dist.all_reduce(stats.max_radii2D, op=dist.ReduceOp.MAX, group=group)
        unpack(flat, params, stats)
    if accumulate_into is not None:
        accumulate_into.xyz_gradient_accum += stats.xyz_gradient_accum
        accumulate_into.denom += stats.denom
        torch.maximum(accumulate_into.max_radii2D, stats.max_radii2D, out=accumulate_into.max_radii2D)
    return flat


class FlatGradBuffer:
    """The view-parallel step without a pack copy: ONE persistent fp32 buffer whose slices ARE the
    parameters' `.grad` tensors and the step's two additive statistics, so the backward kernel's
    fused accumulation (GaussianRasterizer(..., fused_grad_accumulation=..., densification_accumulators=...))
    writes straight into what `all_reduce` sends.  Layout = `pack`'s.

        buf = FlatGradBuffer(params)
        per step:  buf.begin_step()                      # binds .grad, clears the statistics
                   for i, view in enumerate(my_views):
                       rast = GaussianRasterizer(settings, **buf.rasterizer_kwargs(module, first=(i == 0)))
                       ... forward, loss.backward() ...  # view 0 assigns, later views add
                       buf.step_stats.add_view(means2D.grad, radii)   # no-op when fused
                   buf.allreduce(accumulate_into=running_stats)
    """

    def __init__(self, params: Dict[str, torch.Tensor], order=None, defer_sh=(), sh_degree: int = 3,
                 views_per_step: int = 0):
        """`params`: name -> leaf tensor; must contain "means3D" (its first dimension is the Gaussian
        count).  `order` fixes the segment order (default: PARAM_ORDER, else the dict's order).

        Deferred SH gradient (fewer bytes between the GPUs): `defer_sh` names the SH parameter(s) -
        ("shs",) or ("shs_dc", "shs_rest") - and `views_per_step` the batch size.  Each view's
        backward then leaves only its dL/dRGB [P,3] (RasterizerOptions.deferred_sh_grad); the step
        all-gathers those (3 words per view and Gaussian), all-reduces everything except the SH
        segments, and rebuilds the SH gradient from the views' camera centres
        (vegs_sh_grad_from_views): 13 + 3 V instead of 13 + 3 M words per Gaussian on the wire."""
        self.defer_names = tuple(defer_sh)
        layout, total = _layout(params, order, tail=self.defer_names)
        dev = params["means3D"].device
        P = params["means3D"].shape[0]
        self.params = params
        self.flat = torch.zeros(total, dtype=torch.float32, device=dev)
        self.grad_views = {}
        self.step_stats = DensificationStats.__new__(DensificationStats)
        for name, off, n in layout:
            v = self.flat[off:off + n]
            if name in params:
                self.grad_views[name] = v.view_as(params[name])
            else:
                setattr(self.step_stats, name, v.view(P, 1))
        self.step_stats.max_radii2D = torch.zeros((P,), device=dev)
        self.step_stats.fused = False
        self._stat_off = [off for name, off, _ in layout if name == "xyz_gradient_accum"][0]
        self._dirty = True        # gradient slices hold stale values until the first view assigns
        # deferred SH: the SH segments lie after the statistics, everything before them is reduced as usual
        self._sh_off = layout[-len(self.defer_names)][1] if self.defer_names else total
        self._defer_active = False
        self.sh_degree = sh_degree
        if self.defer_names:
            rank, world = (dist.get_rank(), dist.get_world_size()) if (dist.is_available() and dist.is_initialized()) else (0, 1)
            self._rank, self._world = rank, world
            self.slots_per_rank = max(1, -(-int(views_per_step) // world))
            self.drgb_all = torch.zeros((world * self.slots_per_rank, P, 3), dtype=torch.float32, device=dev)
            self.drgb_local = self.drgb_all[rank * self.slots_per_rank:(rank + 1) * self.slots_per_rank]
            self.campos_all = torch.zeros((world * self.slots_per_rank, 3), dtype=torch.float32, device=dev)

    def begin_step(self, view_campos=None) -> None:
        """Bind every parameter's `.grad` to its slice and clear the step's statistics.  The gradient
        slices are NOT cleared here when the first view is rendered with `first=True` (the kernel
        overwrites them); call `zero_grads()` for a path that only ever adds.

        With a deferred SH gradient, `view_campos` = the camera centres ([3] tensors) of ALL views of
        the step in batch order (view j is rendered by rank j % world as its (j // world)-th view)."""
        for k, v in self.grad_views.items():
            self.params[k].grad = v
        self.flat[self._stat_off:self._sh_off].zero_()
        self.step_stats.max_radii2D.zero_()
        self._dirty = True
        self._defer_active = False
        if self.defer_names and view_campos is not None:
            if len(view_campos) > self._world * self.slots_per_rank:
                raise ValueError("more views in the step than views_per_step")
            self.campos_all.zero_()
            self.drgb_local.zero_()                 # unused slots must reduce as "not rendered"
            for j, c in enumerate(view_campos):
                self.campos_all[(j % self._world) * self.slots_per_rank + j // self._world].copy_(c.reshape(3))
            self._defer_active = True

    def zero_grads(self) -> None:
        self.flat[:self._stat_off].zero_()
        self.flat[self._sh_off:].zero_()
        self._dirty = False

    def rasterizer_kwargs(self, rasterizer_module, first: bool, slot: int = None) -> dict:
        """Constructor keyword arguments for the rasterizer of one view of the step ({} for a module
        without the extensions, e.g. the reference: then autograd adds into the bound `.grad` views,
        which needs `zero_grads()` first).  `slot` = the index of the view among this rank's views of
        the step (needed for the deferred SH gradient)."""
        kw = self.step_stats.rasterizer_kwargs(rasterizer_module)
        if self.step_stats.fused:
            kw["fused_grad_accumulation"] = "assign" if (first and self._dirty) else "add"
            self._dirty = False
            if self._defer_active:
                if slot is None or not (0 <= slot < self.slots_per_rank):
                    raise ValueError("deferred SH gradient: pass the view's slot (0 .. views per rank - 1)")
                kw["deferred_sh_grad"] = self.drgb_local[slot]
        else:
            self._defer_active = False           # a module without the extension: plain accumulation
            if self._dirty:
                self.zero_grads()
        return kw

    def allreduce(self, group=None, accumulate_into: DensificationStats = None) -> None:
        active = dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
        if self._defer_active:
            if active:
                dist.all_gather_into_tensor(self.drgb_all.view(-1), self.drgb_local.reshape(-1).clone(), group=group)
                dist.all_reduce(self.flat[:self._sh_off], op=dist.ReduceOp.SUM, group=group)
                dist.all_reduce(self.step_stats.max_radii2D, op=dist.ReduceOp.MAX, group=group)
            self._rebuild_sh_grad()
        elif active:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=group)
            dist.all_reduce(self.step_stats.max_radii2D, op=dist.ReduceOp.MAX, group=group)
        if accumulate_into is not None:
            s = self.step_stats
            accumulate_into.xyz_gradient_accum += s.xyz_gradient_accum
            accumulate_into.denom += s.denom
            torch.maximum(accumulate_into.max_radii2D, s.max_radii2D, out=accumulate_into.max_radii2D)

    def _rebuild_sh_grad(self) -> None:
        """dL/dSH of the whole step from every view's dL/dRGB (vegs_sh_grad_from_views), written
        straight into the SH gradient segment(s) of the flat buffer."""
        import ctypes as C
        from . import _lib
        lib = _lib.load()
        names = self.defer_names
        dc = self.grad_views[names[0]]
        rest = self.grad_views[names[1]] if len(names) > 1 else None
        M = int(dc.shape[1]) + (int(rest.shape[1]) if rest is not None else 0)
        means = self.params["means3D"].detach()
        P = int(means.shape[0])
        dev = means.device
        with torch.cuda.device(dev):
            stream = torch.cuda.current_stream(dev)
            rc = lib.vegs_sh_grad_from_views(P, int(self.sh_degree), M, int(self.drgb_all.shape[0]), means.data_ptr(),
                                             self.campos_all.data_ptr(), self.drgb_all.data_ptr(), dc.data_ptr(),
                                             rest.data_ptr() if (rest is not None and rest.numel()) else None, 0,
                                             C.c_void_p(stream.cuda_stream))
            _lib.check(rc, "vegs_sh_grad_from_views")
