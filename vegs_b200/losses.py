"""Drop-in for VEGS's `loss.normal_guidance.loss_normal_guidance(viewpoint_cam, cov_quat, cov_scale)`
(reference loss/normal_guidance.py:3-23): same arguments, same value, same gradients with respect to
the two covariance maps, computed by ONE kernel (vegs_normal_guidance in include/vegs_b200.h) that
produces the loss and both gradient maps in a single pass over the pixels, instead of a dozen
element-wise / permute / repeat torch kernels over [Npix,3,3] temporaries.  CUDA float32 only."""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib


class _NormalGuidance(torch.autograd.Function):
    @staticmethod
    def forward(ctx, cov_quat, cov_scale, normal_cam, R_c2w, w_rot, w_scale):
        lib = _lib.load()
        if not (cov_quat.is_cuda and cov_scale.is_cuda and normal_cam.is_cuda):
            raise RuntimeError("loss_normal_guidance expects CUDA tensors: there is no CPU path")
        dev = cov_quat.device
        H, W = int(cov_quat.shape[1]), int(cov_quat.shape[2])
        q = cov_quat.detach().contiguous().float()
        s = cov_scale.detach().contiguous().float()
        n = normal_cam.detach().to(dev).contiguous().float()
        R = np.ascontiguousarray(np.asarray(R_c2w, dtype=np.float32).reshape(3, 3))
        loss = torch.empty((), dtype=torch.float32, device=dev)
        dq = torch.empty_like(q) if ctx.needs_input_grad[0] else None
        ds = torch.empty_like(s) if ctx.needs_input_grad[1] else None
        with torch.cuda.device(dev):
            stream = torch.cuda.current_stream(dev)
            rc = lib.vegs_normal_guidance(W, H, q.data_ptr(), s.data_ptr(), n.data_ptr(), R.ctypes.data_as(C.c_void_p),
                                          float(w_rot), float(w_scale), loss.data_ptr(),
                                          dq.data_ptr() if dq is not None else None,
                                          ds.data_ptr() if ds is not None else None, C.c_void_p(stream.cuda_stream))
            _lib.check(rc, "vegs_normal_guidance")
        ctx.dq, ctx.ds = dq, ds
        return loss

    @staticmethod
    def backward(ctx, g):
        dq = ctx.dq * g if ctx.dq is not None else None
        ds = ctx.ds * g if ctx.ds is not None else None
        return dq, ds, None, None, None, None


def loss_normal_guidance(viewpoint_cam, cov_quat, cov_scale, w_rot: float = 0.8, w_scale: float = 0.2):
    """`viewpoint_cam.original_normal` [3,H,W] (camera-space normals) and `viewpoint_cam.R` (3x3
    camera-to-world rotation, numpy or tensor), as scene/cameras.py provides them."""
    R = viewpoint_cam.R
    if torch.is_tensor(R):
        R = R.detach().cpu().numpy()
    return _NormalGuidance.apply(cov_quat, cov_scale, viewpoint_cam.original_normal, R, w_rot, w_scale)
