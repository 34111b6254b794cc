"""Drop-in replacement for `simple_knn._C` (reference submodules/simple-knn/ext.cpp:15-16,
spatial.cu:15-26): `distCUDA2(points[P,3] float32 cuda) -> float32[P]`, the mean of the three
smallest squared distances from each point to the other points.  Runs the sm_100a kernels of
libvegs_b200.so (vegs_dist2 in include/vegs_b200.h); there is no CPU path.
"""
from __future__ import annotations

import ctypes as C

import torch

from .. import _lib

_lib.load()


def distCUDA2(points: torch.Tensor) -> torch.Tensor:
    lib = _lib.load()
    if not points.is_cuda:
        raise RuntimeError("distCUDA2 expects a CUDA tensor: there is no CPU path")
    if points.dtype != torch.float32:
        raise TypeError(f"points must be float32, got {points.dtype}")
    if points.dim() != 2 or points.shape[1] != 3:
        raise RuntimeError("points must have dimensions (num_points, 3)")
    device = points.device
    P = int(points.shape[0])
    means = torch.zeros((P,), dtype=torch.float32, device=device)
    if P == 0:
        return means
    pts = points.contiguous()
    with torch.cuda.device(device):
        stream = torch.cuda.current_stream(device)
        nbytes = lib.vegs_dist2_workspace_bytes(P)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=device)
        rc = lib.vegs_dist2(P, pts.data_ptr(), means.data_ptr(), ws.data_ptr(), nbytes,
                            C.c_void_p(stream.cuda_stream))
        _lib.check(rc, "vegs_dist2")
    return means
