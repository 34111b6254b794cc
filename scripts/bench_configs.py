#!/usr/bin/env python
"""Times BASELINE.md configs 1, 2, 4, 5 for this implementation and, when baseline/_ref is
importable, for the reference's CUDA extension on the same GPU (CUDA events, median of N).
Config 3 is bench.py.  Output: one JSON object (and a markdown table on stderr)."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import util  # noqa: E402
from vegs_b200 import synth  # noqa: E402
import vegs_b200.diff_gaussian_rasterization as ours_mod  # noqa: E402
from vegs_b200.simple_knn._C import distCUDA2 as ours_knn  # noqa: E402

dev = torch.device("cuda")


def median_ms(fn, warm=3, n=10):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return float(np.median(ts)), float(np.min(ts))


def make_step(mod, params, cam, cot, backward=True):
    rs = mod.GaussianRasterizationSettings(cam.H, cam.W, cam.tanfovx, cam.tanfovy, torch.zeros(3, device=dev), 1.0,
                                           cam.viewmatrix, cam.projmatrix, 3, cam.campos, False, False)
    rast = mod.GaussianRasterizer(rs)

    def step():
        for t in params.values():
            t.grad = None
        m2 = torch.zeros_like(params["means3D"], requires_grad=backward)
        with torch.set_grad_enabled(backward):
            out = rast(means3D=params["means3D"], means2D=m2, opacities=params["opacities"], shs=params["shs"],
                       scales=params["scales"], rotations=params["rotations"])
            if backward:
                loss = sum((o * c).sum() for o, c in zip(out[:5], cot))
                loss.backward()
    return step


def main():
    impls = {"ours": ours_mod}
    ref_knn = None
    if util.reference_available():
        ref = util.import_reference()
        impls["reference"] = ref["diff_gaussian_rasterization"]
        ref_knn = ref.get("simple_knn_C")
    res = {}
    cfgs = [("config1_100k_352x94_fwdbwd", 100_000, 0.25, "all"),
            ("config2_2M_1408x376_fwdbwd", 2_000_000, 1.0, "train")]
    for name, P, scale, cotsel in cfgs:
        sc = synth.make_scene(P)
        cam = synth.make_camera(0, scale=scale, device=dev)
        c = util.np_cotangents(cam.W, cam.H)
        keys = ["color", "depth", "cov_quat", "cov_scale", "alpha"]
        cot = [torch.from_numpy(c[k]).to(dev) for k in keys]
        if cotsel == "train":          # colour + cov_quat + cov_scale, as train.py produces
            cot[1].zero_(); cot[4].zero_()
        for iname, mod in impls.items():
            params = {k: v.to(dev).requires_grad_(True) for k, v in sc.items()}
            med, mn = median_ms(make_step(mod, params, cam, cot))
            res.setdefault(name, {})[iname] = dict(ms_median=med, ms_min=mn, views_per_s=1000.0 / med)
    # config 5: render-only sweep at 6M
    sc = synth.make_scene(6_000_000)
    poses = [(0, 0, 0), (0, 30, 0), (0, -30, 0), (0, 60, 0), (0, -60, 0), (-10, 0, 1.0)]
    for scale, tag in ((1.0, "1408x376"), (2.0, "2816x752")):
        for iname, mod in impls.items():
            params = {k: v.to(dev) for k, v in sc.items()}
            tot = 0.0
            for k in (0, 3):
                for (rx, rz, tz) in poses:
                    cam = synth.make_camera(k, scale=scale, add_xrot=rx, add_zrot=rz, add_tz=tz, device=dev)
                    med, _ = median_ms(make_step(mod, params, cam, None, backward=False), warm=1, n=3)
                    tot += med
            n = 2 * len(poses)
            res.setdefault(f"config5_6M_{tag}_render_only", {})[iname] = dict(ms_median=tot / n, views_per_s=1000.0 * n / tot,
                                                                              note=f"mean over {n} poses (cameras 0 and 3)")
    # config 4: distCUDA2 on the 6M positions
    pts = sc["means3D"].to(dev)
    med, mn = median_ms(lambda: ours_knn(pts), warm=1, n=5)
    res["config4_distCUDA2_6M"] = {"ours": dict(ms_median=med, ms_min=mn, mpoints_per_s=6.0 / (med * 1e-3))}
    if ref_knn is not None:
        med, mn = median_ms(lambda: ref_knn.distCUDA2(pts), warm=1, n=3)
        res["config4_distCUDA2_6M"]["reference"] = dict(ms_median=med, ms_min=mn, mpoints_per_s=6.0 / (med * 1e-3))
        same = bool(torch.equal(ours_knn(pts), ref_knn.distCUDA2(pts)))
        res["config4_distCUDA2_6M"]["bitwise_equal_to_reference"] = same
    print(json.dumps(res))
    for name, r in res.items():
        sys.stderr.write(f"| {name} | " + " | ".join(f"{k}: {v.get('ms_median', v) if isinstance(v, dict) else v}" for k, v in r.items()) + " |\n")


if __name__ == "__main__":
    main()
