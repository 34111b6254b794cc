"""Generates tests/golden/*.npz by running the REFERENCE's own CUDA extensions (installed,
unmodified, in baseline/_ref — see DESIGN.md) on the GPU of the box this script runs on:

    gpurun -- 'python tests/golden/make_golden.py gpurun_out/golden'   # then copy into tests/golden/

Nothing in here uses vegs_b200's kernels: the files pin the oracle and the CUDA path to the
reference's behaviour."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import golden_io  # noqa: E402
import util  # noqa: E402
from vegs_b200 import synth  # noqa: E402


def main(out_dir):
    os.makedirs(out_dir, exist_ok=True)
    assert util.reference_available(), "baseline/_ref is missing"
    big = util.scene_with_specials(2500, seed=7)
    small = util.scene_with_specials(1200, seed=9)
    rs = np.random.RandomState(5)
    small_colors = dict(small)
    small_colors["colors"] = torch.from_numpy(rs.uniform(0, 1, (small["means3D"].shape[0], 3)).astype(np.float32))
    del small_colors["shs"]
    cases = [
        ("raster_sh3", big, synth.make_camera(0, scale=0.125), dict(sh_degree=3, bg=(0.0, 0.0, 0.0), scale_modifier=1.0, mode="sh")),
        ("raster_sh1_bg", small, synth.make_camera(1, scale=0.125), dict(sh_degree=1, bg=(0.2, 0.4, 0.6), scale_modifier=0.8, mode="sh")),
        ("raster_sh0_odd", small, synth.make_camera(0, scale=0.142), dict(sh_degree=0, bg=(0.0, 0.0, 0.0), scale_modifier=1.0, mode="sh")),
        ("raster_sh2_yaw", small, synth.make_camera(2, scale=0.125, add_zrot=30.0, add_xrot=-10.0, add_tz=1.0), dict(sh_degree=2, bg=(1.0, 1.0, 1.0), scale_modifier=1.0, mode="sh")),
        ("raster_colors", small_colors, synth.make_camera(0, scale=0.125), dict(sh_degree=0, bg=(0.0, 0.0, 0.0), scale_modifier=1.0, mode="colors")),
    ]
    for name, scene, cam, kw in cases:
        cot = util.np_cotangents(cam.W, cam.H, seed=11)
        ref = util.run_reference(scene, cam, cot=cot, **kw)
        golden_io.save_case(os.path.join(out_dir, name + ".npz"), scene, cam, kw, 11, ref)
        print(name, "P", scene["means3D"].shape[0], "R", ref["num_rendered"], "V", int((ref["radii"] > 0).sum()))
    knn = util.import_reference()["simple_knn_C"]
    pts = synth.make_scene(5000, seed=3)["means3D"]
    pts[100] = pts[101]
    d = knn.distCUDA2(pts.cuda()).cpu().numpy()
    np.savez_compressed(os.path.join(out_dir, "knn_5000.npz"), points=pts.numpy(), dist2=d)
    print("knn", d[:4])


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else HERE)
