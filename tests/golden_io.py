"""Reading / writing of the golden-vector files in tests/golden/ (one .npz per case).

A case stores the inputs (so nothing depends on a random generator's version), the camera, the
call options, and what the REFERENCE's CUDA extension returned for them on a B200: the six
outputs, its decoded geometry / binning / image buffers, and the eight gradients for the seeded
cotangents of util.np_cotangents.  Written by tests/golden/make_golden.py."""
import json

import numpy as np
import torch

from vegs_b200 import synth
import util

REF_KEYS = ("num_rendered", "color", "depth", "cov_quat", "cov_scale", "alpha", "radii", "depths",
            "means2D", "conic_opacity", "rgb", "tiles_touched", "point_list", "point_list_keys",
            "n_contrib", "ranges", "clamped")


def save_case(path, scene, cam, kw, cot_seed, ref):
    d = {}
    for k, v in scene.items():
        d["in_" + k] = v.cpu().numpy()
    d["cam_view"] = cam.viewmatrix.cpu().numpy(); d["cam_proj"] = cam.projmatrix.cpu().numpy()
    d["cam_pos"] = cam.campos.cpu().numpy()
    d["cam_misc"] = np.array([cam.W, cam.H, cam.tanfovx, cam.tanfovy], np.float64)
    d["kw"] = np.array(json.dumps(kw))
    d["cot_seed"] = np.array(cot_seed)
    for k in REF_KEYS:
        d["ref_" + k] = np.asarray(ref[k])
    for k, v in ref["grads"].items():
        d["refg_" + k] = v
    np.savez_compressed(path, **d)


def load_case(path):
    z = np.load(path)
    scene = {k[3:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("in_")}
    W, H, tx, ty = z["cam_misc"]
    cam = synth.Cam(int(W), int(H), float(tx), float(ty), torch.from_numpy(z["cam_view"]),
                    torch.from_numpy(z["cam_proj"]), torch.from_numpy(z["cam_pos"]))
    kw = json.loads(str(z["kw"]))
    kw["bg"] = tuple(kw["bg"])
    ref = {k[4:]: z[k] for k in z.files if k.startswith("ref_")}
    ref["grads"] = {k[5:]: z[k] for k in z.files if k.startswith("refg_")}
    cot = util.np_cotangents(cam.W, cam.H, seed=int(z["cot_seed"]))
    return dict(scene=scene, cam=cam, kw=kw, cot=cot, ref=ref)
