"""Drop-in replacement for VEGS's `diff_gaussian_rasterization` Python module.

Same names, argument order, return order and error behaviour as the reference module
(submodules/diff_gaussian_rasterization/diff_gaussian_rasterization/__init__.py):

    GaussianRasterizationSettings   (:181-193)
    GaussianRasterizer.forward      (:211-244)  -> color, depth, cov_quat, cov_scale, alpha, radii
    GaussianRasterizer.markVisible  (:200-209)
    rasterize_gaussians / _RasterizeGaussians (:21-179)

Underneath, torch is used only for device memory, streams and autograd bookkeeping; the work is
done by the hand-written sm_100a kernels in libvegs_b200.so through its C ABI
(include/vegs_b200.h).  There is no CPU path.
"""
from __future__ import annotations

import ctypes as C
from typing import NamedTuple

import torch
import torch.nn as nn

from .. import _lib

_lib.load()  # fail at import time if the CUDA library is missing

_pinned = {}


def _pinned_word(device: torch.device, stream) -> torch.Tensor:
    key = (device.index, stream.cuda_stream)
    t = _pinned.get(key)
    if t is None:
        t = torch.zeros(2, dtype=torch.int32).pin_memory()
        _pinned[key] = t
    return t


def _ptr(t):
    """Device pointer of a tensor, or NULL for None / empty tensors (the reference's convention:
    empty `torch.Tensor([])` stands for "not provided", __init__.py:221-231)."""
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()


def _dev_f32(t: torch.Tensor, device: torch.device, name: str) -> torch.Tensor:
    if t is None or t.numel() == 0:
        return t
    if t.dtype != torch.float32:
        raise TypeError(f"{name} must be float32, got {t.dtype}")
    if t.device != device:
        t = t.to(device)
    return t.contiguous()


def cpu_deep_copy_tuple(input_tuple):
    return tuple(item.cpu().clone() if isinstance(item, torch.Tensor) else item
                 for item in input_tuple)


class _State:
    """What the forward keeps for the backward (the counterpart of the reference's geomBuffer /
    binningBuffer / imgBuffer) plus, when requested for tests, the forward-only scratch."""
    __slots__ = ("P", "D", "M", "W", "H", "R", "geom", "img", "binning", "scratch", "bin_scratch",
                 "args")


def _forward_impl(bg, means3D, colors_precomp, opacities, scales, rotations, scale_modifier,
                  cov3Ds_precomp, viewmatrix, projmatrix, tanfovx, tanfovy, image_height,
                  image_width, sh, degree, campos, prefiltered, debug, keep_scratch=False,
                  activations=0, sh_rest=None):
    lib = _lib.load()
    if means3D.dim() != 2 or means3D.shape[1] != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    if not means3D.is_cuda:
        raise RuntimeError("means3D must be a CUDA tensor: this rasterizer has no CPU path")
    device = means3D.device
    P = int(means3D.shape[0])
    H, W = int(image_height), int(image_width)
    f32 = dict(dtype=torch.float32, device=device)

    st = _State()
    st.P, st.D, st.W, st.H, st.R = P, int(degree), W, H, 0
    st.M = int(sh.shape[1]) if (sh is not None and sh.numel() != 0) else 0
    if sh_rest is not None:
        st.M += int(sh_rest.shape[1])      # sh is f_dc [P,1,3], sh_rest is f_rest [P,M-1,3]
    st.geom = st.img = st.binning = st.scratch = st.bin_scratch = None
    st.args = None

    if P == 0:
        # reference: outputs stay at their zero fill, nothing is launched (rasterize_points.cu:88)
        z = lambda c: torch.zeros((c, H, W), **f32)
        radii = torch.zeros((0,), dtype=torch.int32, device=device)
        return z(3), z(1), z(4), z(3), z(1), radii, st

    with torch.cuda.device(device):
        stream = torch.cuda.current_stream(device)
        means3D = _dev_f32(means3D, device, "means3D")
        opacities = _dev_f32(opacities, device, "opacities")
        sh = _dev_f32(sh, device, "shs")
        sh_rest = _dev_f32(sh_rest, device, "features_rest")
        colors_precomp = _dev_f32(colors_precomp, device, "colors_precomp")
        scales = _dev_f32(scales, device, "scales")
        rotations = _dev_f32(rotations, device, "rotations")
        cov3Ds_precomp = _dev_f32(cov3Ds_precomp, device, "cov3D_precomp")
        bg = _dev_f32(bg, device, "bg")
        viewmatrix = _dev_f32(viewmatrix, device, "viewmatrix")
        projmatrix = _dev_f32(projmatrix, device, "projmatrix")
        campos = _dev_f32(campos, device, "campos")

        out_color = torch.empty((3, H, W), **f32)
        out_depth = torch.empty((1, H, W), **f32)
        out_quat = torch.empty((4, H, W), **f32)
        out_scale = torch.empty((3, H, W), **f32)
        out_alpha = torch.empty((1, H, W), **f32)
        radii = torch.empty((P,), dtype=torch.int32, device=device)

        u8 = dict(dtype=torch.uint8, device=device)
        st.geom = torch.empty(lib.vegs_geom_bytes(P), **u8)
        st.img = torch.empty(lib.vegs_img_bytes(W, H), **u8)
        st.scratch = torch.empty(lib.vegs_scratch_bytes(P), **u8)
        word = _pinned_word(device, stream)

        a = _lib.VegsForwardArgs()
        a.P, a.D, a.M, a.W, a.H = P, st.D, st.M, W, H
        a.background = _ptr(bg)
        a.means3D = _ptr(means3D)
        a.shs = _ptr(sh)
        a.colors_precomp = _ptr(colors_precomp)
        a.opacities = _ptr(opacities)
        a.scales = _ptr(scales)
        a.rotations = _ptr(rotations)
        a.cov3D_precomp = _ptr(cov3Ds_precomp)
        a.viewmatrix = _ptr(viewmatrix)
        a.projmatrix = _ptr(projmatrix)
        a.campos = _ptr(campos)
        a.scale_modifier = float(scale_modifier)
        a.tan_fovx, a.tan_fovy = float(tanfovx), float(tanfovy)
        a.prefiltered, a.debug = int(bool(prefiltered)), int(bool(debug))
        a.out_color, a.out_depth = out_color.data_ptr(), out_depth.data_ptr()
        a.out_cov_quat, a.out_cov_scale = out_quat.data_ptr(), out_scale.data_ptr()
        a.out_alpha, a.radii = out_alpha.data_ptr(), radii.data_ptr()
        a.geom, a.img, a.scratch = st.geom.data_ptr(), st.img.data_ptr(), st.scratch.data_ptr()
        a.num_rendered_host = word.data_ptr()
        a.activations = int(activations)
        a.shs_rest = _ptr(sh_rest)

        sp = C.c_void_p(stream.cuda_stream)
        _lib.check(lib.vegs_forward_prepare(C.byref(a), sp), "vegs_forward_prepare")
        stream.synchronize()  # the one host sync of the forward (reference: rasterizer_impl.cu:286)
        R = int(word[0].item())
        st.R = R
        if R > 0:
            st.binning = torch.empty(lib.vegs_binning_bytes(R), **u8)
            st.bin_scratch = torch.empty(lib.vegs_bin_scratch_bytes(R, W, H), **u8)
            a.binning, a.bin_scratch = st.binning.data_ptr(), st.bin_scratch.data_ptr()
        _lib.check(lib.vegs_forward_render(C.byref(a), R, sp), "vegs_forward_render")
        if keep_scratch:
            st.args = a
            # keep the inputs alive for as long as the args struct may be used
            st.args._keep = (bg, means3D, sh, sh_rest, colors_precomp, opacities, scales, rotations,
                             cov3Ds_precomp, viewmatrix, projmatrix, campos, out_color, out_depth,
                             out_quat, out_scale, out_alpha, radii)
        else:
            st.scratch = None
            st.bin_scratch = None
    return out_color, out_depth, out_quat, out_scale, out_alpha, radii, st


def export_state(st: _State, radii: torch.Tensor):
    """Test support: the binning / geometry state in the reference's formats (see
    vegs_export_state in include/vegs_b200.h).  Needs a state made with keep_scratch=True."""
    lib = _lib.load()
    if st.args is None:
        raise RuntimeError("export_state needs a forward run with keep_scratch=True")
    device = radii.device
    P, W, H, R = st.P, st.W, st.H, st.R
    T = ((W + 15) // 16) * ((H + 15) // 16)
    out = {
        "point_list_keys": torch.zeros(max(R, 1), dtype=torch.int64, device=device),
        "point_list": torch.zeros(max(R, 1), dtype=torch.int32, device=device),
        "ranges": torch.zeros((T, 2), dtype=torch.int32, device=device),
        "tiles_touched": torch.zeros(P, dtype=torch.int32, device=device),
        "depths": torch.zeros(P, dtype=torch.float32, device=device),
        "means2D": torch.zeros((P, 2), dtype=torch.float32, device=device),
        "conic_opacity": torch.zeros((P, 4), dtype=torch.float32, device=device),
        "rgb": torch.zeros((P, 3), dtype=torch.float32, device=device),
        "n_contrib": torch.zeros((H, W), dtype=torch.int32, device=device),
    }
    with torch.cuda.device(device):
        stream = torch.cuda.current_stream(device)
        rc = lib.vegs_export_state(
            C.byref(st.args), R, out["point_list_keys"].data_ptr(), out["point_list"].data_ptr(),
            out["ranges"].data_ptr(), out["tiles_touched"].data_ptr(), out["depths"].data_ptr(),
            out["means2D"].data_ptr(), out["conic_opacity"].data_ptr(), out["rgb"].data_ptr(),
            out["n_contrib"].data_ptr(), C.c_void_p(stream.cuda_stream))
        _lib.check(rc, "vegs_export_state")
        stream.synchronize()
    out["point_list_keys"] = out["point_list_keys"][:R]
    out["point_list"] = out["point_list"][:R]
    out["num_rendered"] = R
    return out


_ACC_BITS = dict(means3D=1, sh=2, sh_rest=2, colors=4, opacity=8, scales=16, rotations=32, cov3D=64)   # VEGS_ACC_*


class RasterizerOptions:
    """Opt-in extensions of ONE GaussianRasterizer instance (no reference counterpart; a plain
    `GaussianRasterizer(raster_settings)` has none of them and behaves exactly like the reference's
    module).  They are a property of the rasterizer object, not of the module: a second render in
    the same iteration through another instance (VEGS's augmented-camera render, train.py:172-193, or
    a box model with its own Gaussian count, scene/gaussian_model.py:517) is not affected.

    fused_grad_accumulation
        For loops that accumulate several views into `.grad` before an optimiser step.  The backward
        pass adds its result straight into the existing `.grad` of every rasterizer input that is a
        leaf tensor (float32, contiguous, same shape) instead of returning a fresh tensor for
        autograd to add in a separate pass over the parameters; inputs without a `.grad` yet, and
        non-leaf inputs, take the normal route.  With the value "assign" the existing `.grad` is
        overwritten instead of added to (first view of a step whose `.grad` tensors are views of a
        persistent flat buffer that was not cleared; only valid when this rasterizer call is the
        leaf's only consumer in that backward pass, because a gradient autograd adds from another
        path may land before the overwrite).  Hooks registered on those leaves' gradient
        accumulation do not fire for the fused inputs.
    densification_accumulators
        `(xyz_gradient_accum [P,1], denom [P,1], max_radii2D [P])`, VEGS's three densification
        statistics (scene/gaussian_model.py:154-172; float32, CUDA, contiguous).  The backward pass
        of this instance updates them itself for the Gaussians its view rendered (radii > 0):
        `xyz_gradient_accum += ||means2D.grad[:, :2]||`, `denom += 1`,
        `max_radii2D = max(., radii)` - what train.py:300-301 + add_densification_stats do in
        separate passes over all Gaussians.
    deferred_sh_grad
        A `[P,3]` tensor.  The backward pass of this instance writes the view's clamp-masked dL/dRGB
        into it INSTEAD of producing dL/dSH (the SH inputs get no gradient from this call); the sum
        over a step's views is rebuilt later from those 12 bytes per view and Gaussian by
        `vegs_sh_grad_from_views` (vegs_b200.dist.FlatGradBuffer does both ends).  The gradient of one
        view is the outer product basis(view direction) x dL/dRGB, so nothing is lost; a
        view-parallel step then exchanges 3 instead of 3 M words per view and Gaussian.
    live_map_gradients
        CHANGES RESULTS, never default.  The reference computes a gradient of the covariance maps
        (cov_quat, cov_scale) towards the Gaussians' rotations and scales and then overwrites it
        (SURVEY.md Q1), so `loss_normal_guidance` trains nothing there; by default this package
        reproduces that.  With this flag the cotangents of the two maps do reach `scales` and
        `rotations`, with the blending weights held constant (see VegsBackwardArgs in
        include/vegs_b200.h for the exact definition).
    """
    __slots__ = ("fused_grad_accumulation", "densification_accumulators", "live_map_gradients", "deferred_sh_grad")

    def __init__(self, fused_grad_accumulation=False, densification_accumulators=None, live_map_gradients=False,
                 deferred_sh_grad=None):
        self.live_map_gradients = bool(live_map_gradients)
        if deferred_sh_grad is not None:
            t = deferred_sh_grad
            if not (isinstance(t, torch.Tensor) and t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()
                    and t.dim() == 2 and t.shape[1] == 3):
                raise ValueError("deferred_sh_grad must be a contiguous float32 CUDA tensor [P,3]")
        self.deferred_sh_grad = deferred_sh_grad
        if fused_grad_accumulation not in (False, True, "add", "assign"):
            raise ValueError("fused_grad_accumulation must be False, True / 'add', or 'assign'")
        self.fused_grad_accumulation = "add" if fused_grad_accumulation is True else fused_grad_accumulation
        if densification_accumulators is not None:
            acc = tuple(densification_accumulators)
            if len(acc) != 3:
                raise ValueError("densification accumulators: (xyz_gradient_accum, denom, max_radii2D)")
            for t in acc:
                if not (isinstance(t, torch.Tensor) and t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()):
                    raise ValueError("densification accumulators must be contiguous float32 CUDA tensors")
            if not (acc[0].numel() == acc[1].numel() == acc[2].numel()):
                raise ValueError("densification accumulators must have one element per Gaussian")
            densification_accumulators = acc
        self.densification_accumulators = densification_accumulators

    @property
    def any(self) -> bool:
        return bool(self.fused_grad_accumulation) or self.densification_accumulators is not None \
            or self.live_map_gradients or self.deferred_sh_grad is not None


def _accum_target(t, shape, mode):
    """`.grad` of leaf `t` if the backward kernel may write into it, else None."""
    if not mode or t is None or not isinstance(t, torch.Tensor):
        return None
    if not (t.is_leaf and t.requires_grad) or t.grad is None:
        return None
    g = t.grad
    ok = (g.dtype == torch.float32 and g.is_cuda and g.device == t.device and g.is_contiguous()
          and tuple(g.shape) == tuple(shape) and not g.requires_grad)
    return g if ok else None


def _backward_impl(st: _State, bg, means3D, radii, colors_precomp, scales, rotations,
                   scale_modifier, cov3Ds_precomp, viewmatrix, projmatrix, tanfovx, tanfovy,
                   grad_color, grad_depth, grad_cov_quat, grad_cov_scale, grad_alpha, sh, campos,
                   out_alpha, opacities_shape, debug, leaves=None, options=None, activations=0,
                   opacities_raw=None, sh_rest=None, out_cov_quat=None):
    lib = _lib.load()
    device = means3D.device
    P, M = st.P, st.M
    f32 = dict(dtype=torch.float32, device=device)
    have_sh = sh is not None and sh.numel() != 0
    have_sr = scales is not None and scales.numel() != 0
    have_cov = cov3Ds_precomp is not None and cov3Ds_precomp.numel() != 0

    if P == 0:
        e = lambda *s: torch.zeros(s, **f32)
        return (e(0, 3), e(0, 3), e(0, M, 3) if have_sh else None, None if have_sh else e(0, 3),
                torch.zeros(opacities_shape, **f32), e(0, 3) if have_sr else None,
                e(0, 4) if have_sr else None, e(0, 6) if have_cov else None)

    with torch.cuda.device(device):
        stream = torch.cuda.current_stream(device)
        g = lambda t, n: _dev_f32(t, device, n) if t is not None else None
        grad_color, grad_depth = g(grad_color, "grad_color"), g(grad_depth, "grad_depth")
        grad_alpha = g(grad_alpha, "grad_alpha")
        grad_cov_quat, grad_cov_scale = g(grad_cov_quat, "grad_cov_quat"), g(grad_cov_scale, "grad_cov_scale")
        bg = _dev_f32(bg, device, "bg")
        viewmatrix = _dev_f32(viewmatrix, device, "viewmatrix")
        projmatrix = _dev_f32(projmatrix, device, "projmatrix")
        campos = _dev_f32(campos, device, "campos")
        means3D = _dev_f32(means3D, device, "means3D")
        sh = _dev_f32(sh, device, "shs")
        sh_rest = _dev_f32(sh_rest, device, "features_rest")
        opacities_raw = _dev_f32(opacities_raw, device, "opacities")
        colors_precomp = _dev_f32(colors_precomp, device, "colors_precomp")
        scales = _dev_f32(scales, device, "scales")
        rotations = _dev_f32(rotations, device, "rotations")
        cov3Ds_precomp = _dev_f32(cov3Ds_precomp, device, "cov3D_precomp")
        split = sh_rest is not None
        M_dc = 1 if split else M          # rows of the tensor `sh` points at

        # outputs: a fresh tensor, or (fused accumulation) the leaf's existing .grad, added to in place
        leaves = leaves or {}
        acc_mode = options.fused_grad_accumulation if options is not None else False
        acc_mask = 0
        fused = {}

        def out(name, shape, wanted=True):
            nonlocal acc_mask
            if not wanted:
                return None
            tgt = _accum_target(leaves.get(name), shape, acc_mode) if M <= 16 or name not in ("sh", "sh_rest") else None
            if tgt is not None:
                if acc_mode == "add":
                    acc_mask |= _ACC_BITS[name]     # "assign": the kernel's plain overwrite, into .grad
                fused[name] = True
                return tgt
            return torch.empty(shape, **f32)

        dmeans3D = out("means3D", (P, 3))
        dmeans2D = torch.empty((P, 3), **f32)
        dopac = out("opacity", tuple(opacities_shape))
        defer = options.deferred_sh_grad if (options is not None and have_sh and M <= 16) else None
        if defer is not None and (defer.shape[0] != P or defer.device != device):
            raise RuntimeError("deferred_sh_grad does not match this call (P or device)")
        if split and (_accum_target(leaves.get("sh"), (P, 1, 3), acc_mode) is None
                      or _accum_target(leaves.get("sh_rest"), (P, M - 1, 3), acc_mode) is None):
            leaves = {k: v for k, v in leaves.items() if k not in ("sh", "sh_rest")}   # VEGS_ACC_SH covers both
        dsh = out("sh", (P, M_dc, 3), have_sh and defer is None)
        dsh_rest = out("sh_rest", (P, M - 1, 3), split and defer is None)
        dcolors = out("colors", (P, 3), not have_sh)
        dscales = out("scales", (P, 3), have_sr)
        drots = out("rotations", (P, 4), have_sr)
        dcov = out("cov3D", (P, 6), have_cov)
        bwd_scratch = torch.empty(lib.vegs_bwd_scratch_bytes(P), dtype=torch.uint8, device=device)

        b = _lib.VegsBackwardArgs()
        b.P, b.D, b.M, b.W, b.H, b.R = P, st.D, M, st.W, st.H, st.R
        b.background = _ptr(bg)
        b.means3D = _ptr(means3D)
        b.shs = _ptr(sh)
        b.colors_precomp = _ptr(colors_precomp)
        b.scales = _ptr(scales)
        b.rotations = _ptr(rotations)
        b.cov3D_precomp = _ptr(cov3Ds_precomp)
        b.viewmatrix, b.projmatrix, b.campos = _ptr(viewmatrix), _ptr(projmatrix), _ptr(campos)
        b.scale_modifier = float(scale_modifier)
        b.tan_fovx, b.tan_fovy = float(tanfovx), float(tanfovy)
        b.debug = int(bool(debug))
        b.radii, b.out_alpha = radii.data_ptr(), out_alpha.data_ptr()
        b.dL_dout_color, b.dL_dout_depth = _ptr(grad_color), _ptr(grad_depth)
        b.dL_dout_cov_quat, b.dL_dout_cov_scale = _ptr(grad_cov_quat), _ptr(grad_cov_scale)
        b.dL_dout_alpha = _ptr(grad_alpha)
        b.geom, b.img = st.geom.data_ptr(), st.img.data_ptr()
        b.binning = st.binning.data_ptr() if st.binning is not None else None
        b.bwd_scratch = bwd_scratch.data_ptr()
        b.dL_dmeans3D, b.dL_dmeans2D = dmeans3D.data_ptr(), dmeans2D.data_ptr()
        b.dL_dsh = _ptr(dsh)
        b.dL_dcolors = _ptr(dcolors)
        b.dL_dopacity = dopac.data_ptr()
        b.dL_dscales, b.dL_drotations = _ptr(dscales), _ptr(drots)
        b.dL_dcov3D = _ptr(dcov)
        b.accumulate_mask = acc_mask
        b.activations = int(activations)
        b.opacities = _ptr(opacities_raw)
        b.shs_rest = _ptr(sh_rest)
        b.dL_dsh_rest = _ptr(dsh_rest)
        b.dL_drgb_deferred = defer.data_ptr() if defer is not None else None
        if options is not None and options.live_map_gradients:
            if out_cov_quat is None or not have_sr:
                raise RuntimeError("live_map_gradients needs the scale / rotation inputs")
            b.live_map_gradients = 1
            b.out_cov_quat = out_cov_quat.data_ptr()
        if options is not None and options.densification_accumulators is not None:
            sg, sd, sr = options.densification_accumulators
            if sg.numel() != P or sg.device != device:
                raise RuntimeError("densification accumulators do not match this call (P or device)")
            b.stat_grad_accum, b.stat_denom, b.stat_max_radii = sg.data_ptr(), sd.data_ptr(), sr.data_ptr()
        _lib.check(lib.vegs_backward(C.byref(b), C.c_void_p(stream.cuda_stream)), "vegs_backward")
    # inputs whose .grad already received the result get no gradient from autograd
    nf = lambda name, t: None if name in fused else t
    res = (nf("means3D", dmeans3D), dmeans2D, nf("sh", dsh), nf("colors", dcolors), nf("opacity", dopac),
           nf("scales", dscales), nf("rotations", drots), nf("cov3D", dcov))
    return res + (nf("sh_rest", dsh_rest),) if split else res


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                        cov3Ds_precomp, raster_settings, options=None):
    """Same positional arguments as the reference's function (:21-42); `options` (a
    RasterizerOptions, default none) carries the opt-in extensions of the calling rasterizer."""
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales,
                                     rotations, cov3Ds_precomp, raster_settings, options)


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                cov3Ds_precomp, raster_settings, options=None):
        rs = raster_settings
        args = (rs.bg, means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier,
                cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy,
                rs.image_height, rs.image_width, sh, rs.sh_degree, rs.campos, rs.prefiltered,
                rs.debug)
        if rs.debug:
            cpu_args = cpu_deep_copy_tuple(args)  # copy them before they can be corrupted
            try:
                color, depth, cov_quat, cov_scale, alpha, radii, st = _forward_impl(*args)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_fw.dump")
                print("\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.")
                raise ex
        else:
            color, depth, cov_quat, cov_scale, alpha, radii, st = _forward_impl(*args)

        ctx.raster_settings = rs
        ctx.state = st
        ctx.num_rendered = st.R
        ctx.opacities_shape = tuple(opacities.shape)
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh,
                              alpha)
        # the input objects themselves, to find leaves whose .grad the backward may add into
        ctx.options = options if (options is not None and options.any) else None
        ctx.cov_quat = cov_quat.detach() if (ctx.options is not None and ctx.options.live_map_gradients) else None
        ctx.leaves = dict(means3D=means3D, sh=sh, colors=colors_precomp, opacity=opacities, scales=scales,
                          rotations=rotations, cov3D=cov3Ds_precomp) \
            if (ctx.options is not None and ctx.options.fused_grad_accumulation) else None
        ctx.mark_non_differentiable(radii)
        return color, depth, cov_quat, cov_scale, alpha, radii

    @staticmethod
    def backward(ctx, grad_out_color, grad_out_depth, grad_out_cov_quat, grad_out_cov_scale,
                 grad_out_alpha, _):
        rs = ctx.raster_settings
        colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, alpha = ctx.saved_tensors
        args = (ctx.state, rs.bg, means3D, radii, colors_precomp, scales, rotations,
                rs.scale_modifier, cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, rs.tanfovx,
                rs.tanfovy, grad_out_color, grad_out_depth, grad_out_cov_quat, grad_out_cov_scale,
                grad_out_alpha, sh, rs.campos, alpha, ctx.opacities_shape, rs.debug)
        if rs.debug:
            cpu_args = cpu_deep_copy_tuple(args[1:])
            try:
                grads = _backward_impl(*args, leaves=ctx.leaves, options=ctx.options, out_cov_quat=ctx.cov_quat)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_bw.dump")
                print("\nAn error occured in backward. Writing snapshot_bw.dump for debugging.\n")
                raise ex
        else:
            grads = _backward_impl(*args, leaves=ctx.leaves, options=ctx.options, out_cov_quat=ctx.cov_quat)
        (grad_means3D, grad_means2D, grad_sh, grad_colors_precomp, grad_opacities, grad_scales,
         grad_rotations, grad_cov3Ds_precomp) = grads
        return (grad_means3D, grad_means2D, grad_sh, grad_colors_precomp, grad_opacities,
                grad_scales, grad_rotations, grad_cov3Ds_precomp, None, None)


ACT_OPACITY_SIGMOID, ACT_SCALE_EXP, ACT_ROT_NORMALIZE = 1, 2, 4       # VEGS_ACT_* (include/vegs_b200.h)
_ACT_ALL = ACT_OPACITY_SIGMOID | ACT_SCALE_EXP | ACT_ROT_NORMALIZE


class _RasterizeGaussiansRaw(torch.autograd.Function):
    """GaussianRasterizer.forward_raw: the rasterizer on VEGS's RAW parameters.  No reference
    counterpart: it stands for `render()`'s getter calls + the rasterizer call
    (gaussian_renderer/__init__.py:55-82 with scene/gaussian_model.py:100-120), with
    sigmoid / exp / normalize / cat and their backward folded into the two per-Gaussian kernels."""

    @staticmethod
    def forward(ctx, xyz, means2D, features_dc, features_rest, opacity_raw, scaling_raw, rotation_raw,
                raster_settings, options=None):
        rs = raster_settings
        empty = torch.Tensor([])
        color, depth, cov_quat, cov_scale, alpha, radii, st = _forward_impl(
            rs.bg, xyz, empty, opacity_raw, scaling_raw, rotation_raw, rs.scale_modifier, empty, rs.viewmatrix,
            rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, features_dc, rs.sh_degree,
            rs.campos, rs.prefiltered, rs.debug, activations=_ACT_ALL, sh_rest=features_rest)
        ctx.raster_settings = rs
        ctx.state = st
        ctx.opacities_shape = tuple(opacity_raw.shape)
        ctx.save_for_backward(xyz, features_dc, features_rest, opacity_raw, scaling_raw, rotation_raw, radii, alpha)
        ctx.options = options if (options is not None and options.any) else None
        ctx.cov_quat = cov_quat.detach() if (ctx.options is not None and ctx.options.live_map_gradients) else None
        ctx.leaves = dict(means3D=xyz, sh=features_dc, sh_rest=features_rest, opacity=opacity_raw,
                          scales=scaling_raw, rotations=rotation_raw) \
            if (ctx.options is not None and ctx.options.fused_grad_accumulation) else None
        ctx.mark_non_differentiable(radii)
        return color, depth, cov_quat, cov_scale, alpha, radii

    @staticmethod
    def backward(ctx, grad_out_color, grad_out_depth, grad_out_cov_quat, grad_out_cov_scale, grad_out_alpha, _):
        rs = ctx.raster_settings
        xyz, f_dc, f_rest, opacity_raw, scaling_raw, rotation_raw, radii, alpha = ctx.saved_tensors
        empty = torch.Tensor([])
        g = _backward_impl(ctx.state, rs.bg, xyz, radii, empty, scaling_raw, rotation_raw, rs.scale_modifier, empty,
                           rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_out_color, grad_out_depth,
                           grad_out_cov_quat, grad_out_cov_scale, grad_out_alpha, f_dc, rs.campos, alpha,
                           ctx.opacities_shape, rs.debug, leaves=ctx.leaves, options=ctx.options,
                           activations=_ACT_ALL, opacities_raw=opacity_raw, sh_rest=f_rest, out_cov_quat=ctx.cov_quat)
        if len(g) == 8:                      # no higher-order coefficients: features_rest was empty
            g = g + (torch.zeros_like(f_rest),)
        g_xyz, g_m2d, g_dc, _gc, g_op, g_sc, g_rot, _gcov, g_rest = g
        return g_xyz, g_m2d, g_dc, g_rest, g_op, g_sc, g_rot, None, None


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings, *, fused_grad_accumulation=False, densification_accumulators=None,
                 live_map_gradients=False, deferred_sh_grad=None):
        """`GaussianRasterizer(raster_settings)` is the reference's constructor (:195-198).  The two
        keyword-only arguments are this package's opt-in extensions for THIS instance; see
        RasterizerOptions."""
        super().__init__()
        self.raster_settings = raster_settings
        self.options = RasterizerOptions(fused_grad_accumulation, densification_accumulators, live_map_gradients,
                                         deferred_sh_grad)

    def markVisible(self, positions):
        """Boolean mask of the points in front of the near plane (view z > 0.2)."""
        lib = _lib.load()
        with torch.no_grad():
            rs = self.raster_settings
            if not positions.is_cuda:
                raise RuntimeError("positions must be a CUDA tensor: there is no CPU path")
            device = positions.device
            P = int(positions.shape[0])
            visible = torch.zeros((P,), dtype=torch.bool, device=device)
            if P != 0:
                pos = _dev_f32(positions, device, "positions")
                view = _dev_f32(rs.viewmatrix, device, "viewmatrix")
                proj = _dev_f32(rs.projmatrix, device, "projmatrix")
                with torch.cuda.device(device):
                    stream = torch.cuda.current_stream(device)
                    rc = lib.vegs_mark_visible(P, pos.data_ptr(), view.data_ptr(), proj.data_ptr(),
                                               visible.data_ptr(), C.c_void_p(stream.cuda_stream))
                    _lib.check(rc, "vegs_mark_visible")
        return visible

    def forward_raw(self, xyz, means2D, opacity_raw, features_dc, features_rest, scaling_raw, rotation_raw):
        """Opt-in extension (no reference counterpart): rasterise from VEGS's RAW parameters
        `_xyz [P,3], _opacity [P,1], _features_dc [P,1,3], _features_rest [P,M-1,3], _scaling [P,3],
        _rotation [P,4]`.  Equivalent to
            forward(xyz, means2D, sigmoid(opacity_raw), shs=cat((features_dc, features_rest), 1),
                    scales=exp(scaling_raw), rotations=normalize(rotation_raw))
        (scene/gaussian_model.py:100-120) with the four activation passes and their backward passes
        fused into the per-Gaussian kernels.  Same six return values; gradients arrive at the raw
        parameters."""
        if not (features_dc.dim() == 3 and features_dc.shape[1] == 1 and features_rest.dim() == 3):
            raise Exception('features_dc must be [P,1,3] and features_rest [P,M-1,3]')
        return _RasterizeGaussiansRaw.apply(xyz, means2D, features_dc, features_rest, opacity_raw, scaling_raw,
                                            rotation_raw, self.raster_settings,
                                            self.options if self.options.any else None)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None,
                rotations=None, cov3D_precomp=None):
        raster_settings = self.raster_settings

        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')

        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')

        if shs is None:
            shs = torch.Tensor([])
        if colors_precomp is None:
            colors_precomp = torch.Tensor([])
        if scales is None:
            scales = torch.Tensor([])
        if rotations is None:
            rotations = torch.Tensor([])
        if cov3D_precomp is None:
            cov3D_precomp = torch.Tensor([])

        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales,
                                   rotations, cov3D_precomp, raster_settings,
                                   self.options if self.options.any else None)
