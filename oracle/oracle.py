"""numpy/ctypes front end of the CPU oracle (oracle/vegs_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline
legs, never by the vegs_b200 package.  See the header of vegs_oracle.c for what it restates and
how it is pinned.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libvegs_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "vegs_oracle.c")
    if force or not os.path.exists(LIB_PATH) or os.path.getmtime(src) > os.path.getmtime(LIB_PATH):
        subprocess.run(["make", "-C", _HERE, "-B", "libvegs_oracle.so"], check=True,
                       capture_output=True)
    return LIB_PATH


def load() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        _lib = C.CDLL(LIB_PATH)
        _lib.oracle_num_threads.restype = C.c_int
    return _lib


def num_threads() -> int:
    return int(load().oracle_num_threads())


def set_num_threads(n: int) -> None:
    load().oracle_set_num_threads(C.c_int(n))


def _f(a, shape=None):
    if a is None:
        return None
    a = np.ascontiguousarray(a, dtype=np.float32)
    if shape is not None:
        a = a.reshape(shape)
    return a


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _np(t):
    """torch tensor / array-like / None -> numpy (host)."""
    if t is None:
        return None
    if hasattr(t, "detach"):
        t = t.detach().cpu().numpy()
    a = np.asarray(t)
    return None if a.size == 0 else a


def forward(*, means3D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
            cov3D_precomp=None, viewmatrix, projmatrix, campos, bg, W, H, tanfovx, tanfovy,
            sh_degree=3, scale_modifier=1.0, render=True):
    """Full forward of the reference algorithm on the CPU.  Returns a dict with every intermediate
    (the reference's geomBuffer / binningBuffer / imgBuffer contents) and the six outputs."""
    lib = load()
    means3D = _f(_np(means3D)); P = means3D.shape[0]
    opac = _f(_np(opacities)).reshape(-1)
    shs = _f(_np(shs)); colors = _f(_np(colors_precomp)); scales = _f(_np(scales))
    rots = _f(_np(rotations)); cov3 = _f(_np(cov3D_precomp))
    view = _f(_np(viewmatrix)).reshape(16); proj = _f(_np(projmatrix)).reshape(16)
    cam = _f(_np(campos)).reshape(3); bg = _f(_np(bg)).reshape(3)
    M = 0 if shs is None else shs.shape[1]
    gx, gy = (W + 15) // 16, (H + 15) // 16
    T = gx * gy
    o = dict(P=P, W=W, H=H, M=M, D=sh_degree)
    o["radii"] = np.zeros(P, np.int32)
    o["depths"] = np.zeros(P, np.float32)
    o["means2D"] = np.zeros((P, 2), np.float32)
    o["conic_opacity"] = np.zeros((P, 4), np.float32)
    o["rgb"] = np.zeros((P, 3), np.float32)
    o["clamped"] = np.zeros((P, 3), np.uint8)
    o["tiles_touched"] = np.zeros(P, np.uint32)
    o["rect"] = np.zeros((P, 4), np.int32)
    o["cov3D"] = np.zeros((P, 6), np.float32)
    lib.oracle_preprocess(
        C.c_int(P), C.c_int(sh_degree), C.c_int(M), _p(means3D), _p(scales),
        C.c_float(scale_modifier), _p(rots), _p(opac), _p(shs), _p(cov3), _p(colors), _p(view),
        _p(proj), _p(cam), C.c_int(W), C.c_int(H), C.c_float(tanfovx), C.c_float(tanfovy),
        _p(o["radii"]), _p(o["depths"]), _p(o["means2D"]), _p(o["conic_opacity"]), _p(o["rgb"]),
        _p(o["clamped"]), _p(o["tiles_touched"]), _p(o["rect"]), _p(o["cov3D"]))
    R = int(o["tiles_touched"].astype(np.int64).sum())
    o["num_rendered"] = R
    o["point_list_keys"] = np.zeros(max(R, 1), np.uint64)
    o["point_list"] = np.zeros(max(R, 1), np.uint32)
    o["ranges"] = np.zeros((T, 2), np.uint32)
    lib.oracle_bin(C.c_int(P), _p(o["radii"]), _p(o["rect"]), _p(o["depths"]), C.c_int(W),
                   C.c_int(H), C.c_uint64(R), _p(o["point_list_keys"]), _p(o["point_list"]),
                   _p(o["ranges"]))
    o["point_list_keys"] = o["point_list_keys"][:R]
    o["point_list"] = o["point_list"][:R]
    if render:
        o["color"] = np.zeros((3, H, W), np.float32)
        o["depth"] = np.zeros((1, H, W), np.float32)
        o["cov_quat"] = np.zeros((4, H, W), np.float32)
        o["cov_scale"] = np.zeros((3, H, W), np.float32)
        o["alpha"] = np.zeros((1, H, W), np.float32)
        o["n_contrib"] = np.zeros((H, W), np.uint32)
        feat = colors if colors is not None else o["rgb"]
        plist = o["point_list"] if R > 0 else np.zeros(1, np.uint32)
        lib.oracle_render(_p(o["ranges"]), _p(plist), C.c_int(W), C.c_int(H), _p(o["means2D"]),
                          _p(feat), _p(o["depths"]), _p(o["conic_opacity"]), _p(rots), _p(scales),
                          _p(bg), _p(o["color"]), _p(o["depth"]), _p(o["cov_quat"]),
                          _p(o["cov_scale"]), _p(o["alpha"]), _p(o["n_contrib"]))
    o["_inputs"] = dict(means3D=means3D, opac=opac, shs=shs, colors=colors, scales=scales,
                        rots=rots, cov3=cov3, view=view, proj=proj, cam=cam, bg=bg,
                        tanfovx=tanfovx, tanfovy=tanfovy, scale_modifier=scale_modifier)
    return o


def backward(fwd, *, grad_color=None, grad_depth=None, grad_alpha=None):
    """Backward of the reference algorithm given a forward() result and image cotangents.
    (Cotangents of cov_quat / cov_scale do not reach any returned gradient in the reference —
    SURVEY.md Q1 — so they are not inputs here.)"""
    lib = load()
    i = fwd["_inputs"]
    P, W, H, M, D = fwd["P"], fwd["W"], fwd["H"], fwd["M"], fwd["D"]
    gc, gd, ga = _f(_np(grad_color)), _f(_np(grad_depth)), _f(_np(grad_alpha))
    dmean2D = np.zeros((P, 3), np.float32); dconic = np.zeros((P, 4), np.float32)
    dopac = np.zeros(P, np.float32); dcolors = np.zeros((P, 3), np.float32)
    ddepths = np.zeros(P, np.float32)
    feat = i["colors"] if i["colors"] is not None else fwd["rgb"]
    R = fwd["num_rendered"]
    plist = fwd["point_list"] if R > 0 else np.zeros(1, np.uint32)
    lib.oracle_render_backward(
        C.c_int(P), _p(fwd["ranges"]), _p(plist), C.c_int(W), C.c_int(H), _p(i["bg"]),
        _p(fwd["means2D"]), _p(fwd["conic_opacity"]), _p(feat), _p(fwd["depths"]),
        _p(fwd["alpha"]), _p(fwd["n_contrib"]), _p(gc), _p(gd), _p(ga), _p(dmean2D), _p(dconic),
        _p(dopac), _p(dcolors), _p(ddepths))
    dmeans3D = np.zeros((P, 3), np.float32)
    dsh = np.zeros((P, M, 3), np.float32) if M else None
    dcov3D = np.zeros((P, 6), np.float32)
    dscales = np.zeros((P, 3), np.float32) if i["scales"] is not None else None
    drots = np.zeros((P, 4), np.float32) if i["rots"] is not None else None
    lib.oracle_preprocess_backward(
        C.c_int(P), C.c_int(D), C.c_int(M), _p(i["means3D"]), _p(fwd["radii"]), _p(i["shs"]),
        _p(fwd["clamped"]), _p(i["scales"]), _p(i["rots"]), C.c_float(i["scale_modifier"]),
        _p(i["cov3"]), _p(i["view"]), _p(i["proj"]), _p(i["cam"]), C.c_int(W), C.c_int(H),
        C.c_float(i["tanfovx"]), C.c_float(i["tanfovy"]), _p(dmean2D), _p(dconic), _p(dcolors),
        _p(ddepths), _p(dmeans3D), _p(dsh), _p(dcov3D), _p(dscales), _p(drots))
    return dict(means3D=dmeans3D, means2D=dmean2D, sh=dsh, colors_precomp=dcolors,
                opacities=dopac, scales=dscales, rotations=drots, cov3D=dcov3D)


def mark_visible(means3D, viewmatrix):
    lib = load()
    m = _f(_np(means3D)); v = _f(_np(viewmatrix)).reshape(16)
    out = np.zeros(m.shape[0], np.uint8)
    lib.oracle_mark_visible(C.c_int(m.shape[0]), _p(m), _p(v), _p(out))
    return out.astype(bool)


def dist2(points, method="auto"):
    """distCUDA2 on the CPU: brute force (exact definition) or the reference's box algorithm."""
    lib = load()
    p = _f(_np(points)); P = p.shape[0]
    out = np.zeros(P, np.float32)
    if method == "auto":
        method = "bruteforce" if P <= 30000 else "boxes"
    fn = lib.oracle_dist2_bruteforce if method == "bruteforce" else lib.oracle_dist2_boxes
    fn(C.c_int(P), _p(p), _p(out))
    return out
