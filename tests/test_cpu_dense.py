"""CPU test: the oracle's C restatement against the dense PyTorch evaluator (oracle/dense_torch.py),
forward in float32/float64 and backward against torch.autograd in float64.  This checks the
hand-derived gradient formulas of vegs_oracle.c (the same ones the CUDA kernels implement)
independently of the reference's code."""
import numpy as np
import pytest
import torch

import util
from oracle import dense_torch
from vegs_b200 import synth


def _scene(n=250, seed=4):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.rand(*s, generator=g)
    means = torch.stack([13 + 10 * r(n), (r(n) - 0.5) * 8, 1.55 + (r(n) - 0.5) * 2.5], 1)
    return dict(means3D=means, scales=0.05 + 0.25 * r(n, 3),
                rotations=torch.nn.functional.normalize(torch.randn(n, 4, generator=g), dim=1),
                opacities=0.1 + 0.6 * r(n, 1), shs=torch.randn(n, 16, 3, generator=g) * 0.15)


def _dense(scene, cam, dtype, bg=(0.1, 0.3, 0.5), requires_grad=False, deg=3):
    t = {k: v.to(dtype).clone().requires_grad_(requires_grad) for k, v in scene.items()}
    out = dense_torch.render_dense(
        t["means3D"], t["opacities"], shs=t["shs"], scales=t["scales"], rotations=t["rotations"],
        viewmatrix=cam.viewmatrix, projmatrix=cam.projmatrix, campos=cam.campos,
        bg=torch.tensor(bg), W=cam.W, H=cam.H, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, sh_degree=deg)
    return t, out


def test_forward_oracle_matches_dense_evaluator():
    scene = _scene()
    cam = synth.make_camera(0, scale=0.125)
    f = util.run_oracle(scene, cam, bg=(0.1, 0.3, 0.5))
    _, d = _dense(scene, cam, torch.float64)
    assert np.array_equal(f["radii"], d["radii"].numpy())
    nc_bad = (f["n_contrib"] != d["n_contrib"].numpy()).mean()
    assert nc_bad < 2e-3
    for k in ("color", "depth", "alpha", "cov_scale", "cov_quat"):
        err = np.abs(f[k] - d[k].numpy())
        assert (err > 2e-5).mean() < 2e-3, (k, float(err.max()))
        assert np.median(err) < 2e-6, k


@pytest.mark.parametrize("deg", [3, 1])
def test_backward_oracle_matches_autograd(deg):
    scene = _scene()
    cam = synth.make_camera(0, scale=0.125)
    cot = util.np_cotangents(cam.W, cam.H, seed=5)
    f = util.run_oracle(scene, cam, bg=(0.1, 0.3, 0.5), cot=cot, sh_degree=deg)
    t, d = _dense(scene, cam, torch.float64, requires_grad=True, deg=deg)
    loss = ((d["color"] * torch.from_numpy(cot["color"]).double()).sum()
            + (d["depth"] * torch.from_numpy(cot["depth"]).double()).sum()
            + (d["alpha"] * torch.from_numpy(cot["alpha"]).double()).sum())
    loss.backward()
    pairs = {"means3D": "means3D", "sh": "shs", "opacities": "opacities", "scales": "scales",
             "rotations": "rotations"}
    for gk, tk in pairs.items():
        ref = t[tk].grad.numpy().reshape(f["grads"][gk].shape)
        util.grad_close(f["grads"][gk], ref, f"oracle grad {gk} vs autograd", rtol=2e-3)
