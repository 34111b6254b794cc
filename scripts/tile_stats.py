"""Diagnostics: per-tile list length, deepest contributor and CTA cycles of the compositing kernels
for one view of the bench scene.  Run on the GPU box; prints a summary, saves an .npz."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import util  # noqa: E402
from vegs_b200 import synth, _lib  # noqa: E402
import vegs_b200.diff_gaussian_rasterization as dgr  # noqa: E402

P = int(sys.argv[1]) if len(sys.argv) > 1 else 6_000_000
out = sys.argv[2] if len(sys.argv) > 2 else "gpurun_out/tile_stats.npz"
dev = torch.device("cuda")
sc = synth.make_scene(P)
if len(sys.argv) > 3 and sys.argv[3] == "sorted":      # experiment: Gaussians stored in depth order
    perm = torch.argsort(sc["means3D"][:, 0])
    sc = {k: v[perm].contiguous() for k, v in sc.items()}
sc = {k: v.to(dev) for k, v in sc.items()}
cam = synth.make_camera(0, device=dev)
lib = _lib.load()
e = torch.Tensor([])
res = dgr._forward_impl(torch.zeros(3, device=dev), sc["means3D"], e, sc["opacities"], sc["scales"], sc["rotations"], 1.0, e,
                        cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, cam.H, cam.W, sc["shs"], 3, cam.campos,
                        False, False, keep_scratch=True)
st = res[6]
state = dgr.export_state(st, res[5])
g = torch.randn(3, cam.H, cam.W, device=dev)
grads = dgr._backward_impl(st, torch.zeros(3, device=dev), sc["means3D"], res[5], e, sc["scales"], sc["rotations"], 1.0, e,
                           cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, g, None, None, None, None,
                           sc["shs"], cam.campos, res[4], (P, 1), False)
T = state["ranges"].shape[0]
fwd = torch.zeros(T, dtype=torch.int32, device=dev); bwd = torch.zeros(T, dtype=torch.int32, device=dev)
dbg = torch.zeros(T * 16, dtype=torch.int32, device=dev)
_lib.check(lib.vegs_export_tile_clocks(st.img.data_ptr(), cam.W, cam.H, fwd.data_ptr(), bwd.data_ptr(), dbg.data_ptr(),
                                       C.c_void_p(torch.cuda.current_stream().cuda_stream)), "clocks")
torch.cuda.synchronize()
rl = (state["ranges"][:, 1] - state["ranges"][:, 0]).cpu().numpy().astype(np.int64)
gx, gy = (cam.W + 15) // 16, (cam.H + 15) // 16
nc = state["n_contrib"].cpu().numpy().astype(np.int64)
pad = np.zeros((gy * 16, gx * 16), np.int64); pad[:cam.H, :cam.W] = nc
tmax = pad.reshape(gy, 16, gx, 16).max(axis=(1, 3)).reshape(-1)
tsum = pad.reshape(gy, 16, gx, 16).sum(axis=(1, 3)).reshape(-1)
f = fwd.cpu().numpy().astype(np.int64) * 1024; b = bwd.cpu().numpy().astype(np.int64) * 1024
np.savez_compressed(out, len=rl, tmax=tmax, tsum=tsum, fwd_cycles=f, bwd_cycles=b, gx=gx, gy=gy)
order = np.argsort(-f)
print("tiles", T, "R", rl.sum(), "sum fwd Mcycles", f.sum() / 1e6, "sum bwd Mcycles", b.sum() / 1e6)
print("top fwd tiles: (tile, row, len, maxcontrib, fwd max-unit Mcyc, bwd max-unit Mcyc, fwd cyc/entry)")
for t in order[:12]:
    print(t, t // gx, rl[t], tmax[t], round(f[t] / 1e6, 2), round(b[t] / 1e6, 2), round(f[t] / max(rl[t], 1), 1))
draw = dbg.cpu().numpy().astype(np.int64).reshape(T, 16)
d = draw[:, :8] * 1024
if d.any():
    print("phase Mcycles of unit 0 of the top tiles: fwd scan(+S4) S0 S2 S3 | bwd S0 pass1 pass2 S3")
    for t in order[:12]:
        print(t, rl[t], " ".join("%.2f" % (x / 1e6) for x in d[t, :4]), "|", " ".join("%.2f" % (x / 1e6) for x in d[t, 4:]),
              "| sets %d cand %d hit-entries %d seg0 pixel-hits %d" % tuple(draw[t, 8:12]))
    print("all tiles:", " ".join("%.1f" % (x / 1e6) for x in d.sum(axis=0)))
print("len percentiles", np.percentile(rl, [50, 90, 99, 100]))
print("fwd cycles percentiles (M)", np.percentile(f, [50, 90, 99, 100]) / 1e6)
print("bwd cycles percentiles (M)", np.percentile(b, [50, 90, 99, 100]) / 1e6)
for row in range(gy):
    sel = slice(row * gx, (row + 1) * gx)
    print("row", row, "len mean", int(rl[sel].mean()), "max", rl[sel].max(), "fwd Mcyc mean", round(f[sel].mean() / 1e6, 2), "max", round(f[sel].max() / 1e6, 2),
          "bwd mean", round(b[sel].mean() / 1e6, 2), "max", round(b[sel].max() / 1e6, 2), "contrib/px", round(tsum[sel].sum() / (gx * 256), 1))
