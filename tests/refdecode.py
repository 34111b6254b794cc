"""Decoding of the REFERENCE extension's opaque byte buffers, used by the
golden-vector generator and by the live reference-vs-ours GPU tests.

geomBuffer / binningBuffer / imgBuffer are carved by `obtain(chunk, ptr, count, 128)`
(reference cuda_rasterizer/rasterizer_impl.h:21-27, rasterizer_impl.cu:155-193): each array starts
at the next 128-byte boundary.
"""
import numpy as np
import torch


def _carve(buf: torch.Tensor, off: int, dtype, count: int):
    off = (off + 127) & ~127
    nbytes = count * np.dtype(dtype).itemsize
    arr = buf[off:off + nbytes].cpu().numpy().view(dtype).copy()
    return arr, off + nbytes


def decode_geom(buf: torch.Tensor, P: int):
    o = {}
    off = 0
    o["depths"], off = _carve(buf, off, np.float32, P)
    o["clamped"], off = _carve(buf, off, np.uint8, 3 * P)
    o["internal_radii"], off = _carve(buf, off, np.int32, P)
    o["means2D"], off = _carve(buf, off, np.float32, 2 * P)
    o["cov3D"], off = _carve(buf, off, np.float32, 6 * P)
    o["conic_opacity"], off = _carve(buf, off, np.float32, 4 * P)
    o["rgb"], off = _carve(buf, off, np.float32, 3 * P)
    o["tiles_touched"], off = _carve(buf, off, np.uint32, P)
    o["means2D"] = o["means2D"].reshape(P, 2)
    o["cov3D"] = o["cov3D"].reshape(P, 6)
    o["conic_opacity"] = o["conic_opacity"].reshape(P, 4)
    o["rgb"] = o["rgb"].reshape(P, 3)
    o["clamped"] = o["clamped"].reshape(P, 3)
    return o


def decode_binning(buf: torch.Tensor, R: int):
    o = {}
    off = 0
    o["point_list"], off = _carve(buf, off, np.uint32, R)
    _, off = _carve(buf, off, np.uint32, R)
    o["point_list_keys"], off = _carve(buf, off, np.uint64, R)
    return o


def decode_img(buf: torch.Tensor, W: int, H: int):
    o = {}
    off = 0
    N = W * H
    T = ((W + 15) // 16) * ((H + 15) // 16)
    o["n_contrib"], off = _carve(buf, off, np.uint32, N)
    ranges, off = _carve(buf, off, np.uint32, 2 * N)
    o["n_contrib"] = o["n_contrib"].reshape(H, W)
    o["ranges"] = ranges.reshape(N, 2)[:T].copy()
    return o
