"""GPU parity tests AT THE SIZES THE NUMBERS ARE QUOTED ON (BASELINE.json configs 2-5), against the
reference's own CUDA extension run live on the same box (baseline/_ref), plus the reference's
callers run unchanged against both modules.

  config 2   2 M Gaussians, 1408x376, camera 0, fwd+bwd
  config 3   6 M Gaussians, 1408x376, cameras 0 and 7 (the horizon-tile case: lists of 1e5 entries,
             hundreds of working sets per unit, ring wrap-around, 8192-instance emission parts)
  config 4   distCUDA2 on the 6 M positions, bitwise
  config 5   6 M forward-only at 2816x752 (T = 8272: 14 tile bits, the second tile-sort pass is live)
             and yawed / pitched extrapolated poses (scene/cameras.py:178-227) at both resolutions

Contract: radii, depth bits, means2D, conics, tiles touched, sort keys, point list,
tile ranges and n_contrib bit-exact; images within 1e-5 max-abs; gradients within 1e-4 of the
tensor's max norm AND element-wise above a magnitude floor, with the reference's own run-to-run
spread (float atomics) measured by the same statistics and recorded beside the error
(gpurun_out/r02_parity.json -> profiles/r02_parity.json).
"""
import functools

import numpy as np
import pytest
import torch

import util
from vegs_b200 import synth
from test_gpu_parity import _check_state_exact, _check_images

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not util.reference_available(), reason="baseline/_ref not present")]

GRAD_NAMES = ("means3D", "means2D", "sh", "opacities", "scales", "rotations")


@functools.lru_cache(maxsize=2)
def _scene(P):
    return synth.make_scene(P, seed=7)


def _train_cotangents(W, H):
    """What train.py:162-168 produces: cotangents on colour, cov_quat and cov_scale only."""
    cot = util.np_cotangents(W, H)
    cot["depth"][:] = 0
    cot["alpha"][:] = 0
    return cot


def _fwd_bwd_case(tag, P, k, full_cotangents):
    scene = _scene(P)
    cam = synth.make_camera(k)
    cot = util.np_cotangents(cam.W, cam.H) if full_cotangents else _train_cotangents(cam.W, cam.H)
    ours = util.run_ours(scene, cam, cot=cot)
    ref = util.run_reference(scene, cam, cot=cot)
    rec = util.parity_record(tag)
    vis = ref["radii"] > 0
    _check_state_exact(ours, ref, vis)
    _check_images(ours, ref, strict=True)
    rec["sizes"] = dict(P=P, V=int(vis.sum()), R=int(ref["num_rendered"]), W=cam.W, H=cam.H, camera=k,
                        longest_tile_list=int((ref["ranges"][:, 1].astype(np.int64) - ref["ranges"][:, 0]).max()))
    rec["images_max_abs_err"] = {n: float(np.abs(ours[n].astype(np.float64) - ref[n]).max())
                                 for n in ("color", "depth", "alpha", "cov_quat", "cov_scale")}
    rec["n_contrib_mismatches"] = int((ours["state"]["n_contrib"].astype(np.uint32) != ref["n_contrib"]).sum())
    ref2 = util.run_reference(scene, cam, cot=cot)
    grads = rec.setdefault("grads", {})
    for name in GRAD_NAMES:
        g = ours["grads"][name]
        util.grad_check_full(g, np.asarray(ref["grads"][name]).reshape(g.shape),
                             np.asarray(ref2["grads"][name]).reshape(g.shape), name, rtol=1e-4, record=grads)


def test_config2_2M_camera0_fwd_bwd():
    _fwd_bwd_case("config2_2M_1408x376_cam0", 2_000_000, 0, full_cotangents=False)


@pytest.mark.parametrize("k", [0, 7])
def test_config3_6M_fwd_bwd(k):
    # camera 0 with cotangents on all five outputs (every backward input live), camera 7 as train.py feeds it
    _fwd_bwd_case(f"config3_6M_1408x376_cam{k}", 6_000_000, k, full_cotangents=(k == 0))


@pytest.mark.parametrize("scale,pose", [(2.0, (0.0, 0.0, 0.0)), (2.0, (-10.0, 60.0, 1.0)), (1.0, (-10.0, 60.0, 1.0)),
                                        (1.0, (0.0, -30.0, 0.0))])
def test_config5_6M_render_only(scale, pose):
    rx, rz, tz = pose
    scene = _scene(6_000_000)
    cam = synth.make_camera(3, scale=scale, add_xrot=rx, add_zrot=rz, add_tz=tz)
    ours = util.run_ours(scene, cam)
    ref = util.run_reference(scene, cam)
    _check_state_exact(ours, ref, ref["radii"] > 0)
    _check_images(ours, ref, strict=True)
    rec = util.parity_record(f"config5_6M_{cam.W}x{cam.H}_rx{rx:g}_rz{rz:g}_tz{tz:g}")
    rec["sizes"] = dict(P=6_000_000, V=int((ref["radii"] > 0).sum()), R=int(ref["num_rendered"]), W=cam.W, H=cam.H,
                        tiles=int(ref["ranges"].shape[0]))
    rec["images_max_abs_err"] = {n: float(np.abs(ours[n].astype(np.float64) - ref[n]).max())
                                 for n in ("color", "depth", "alpha", "cov_quat", "cov_scale")}


def test_config4_dist2_6M_bitwise():
    from vegs_b200.simple_knn._C import distCUDA2
    knn = util.import_reference()["simple_knn_C"]
    pts = _scene(6_000_000)["means3D"].cuda()
    ours = distCUDA2(pts).cpu().numpy()
    ref = knn.distCUDA2(pts).cpu().numpy()
    util.assert_bit_equal(ours, ref, "distCUDA2 6M")
    util.parity_record("config4_distCUDA2_6M")["bitwise_equal"] = True


# ------------------------------------------------------------------ the reference's callers, unchanged
def _run_caller(module, scene, cams, gts, normals, n_iter=2):
    """Two iterations of the rasterizer-facing part of VEGS's loop (tests/vegs_caller.py) with
    `diff_gaussian_rasterization` resolving to `module`."""
    import vegs_caller as vc
    pc = vc.StubGaussianModel(scene)
    bg = torch.zeros(3, device="cuda")
    R_c2w = np.array([[0.0, 0.0, 1.0], [-1.0, 0.0, 0.0], [0.0, -1.0, 0.0]])
    pkgs = []
    with util.use_rasterizer_module(module):
        for i in range(n_iter):
            cam_ns = vc.stub_camera(cams[i], R_c2w=R_c2w, normal=normals[i])
            pkg, loss = vc.training_iteration(pc, cam_ns, bg, gts[i])
            pkgs.append({k: v.detach().clone() for k, v in pkg.items() if k != "viewspace_points"}
                        | {"viewspace_grad": pkg["viewspace_points"].grad.clone(), "loss": loss.detach().clone()})
    return pc, pkgs


def test_reference_callers_run_unchanged_on_both_modules():
    """gaussian_renderer.render() + loss_normal_guidance + add_densification_stats (restated in
    tests/vegs_caller.py, non-leaf means2D with retain_grad, non-leaf opacity / scales / rotations /
    SH from the model's getters), once through the reference's module and once through
    vegs_b200.install(): same dict, same viewspace gradient, same raw-parameter gradients, same
    densification statistics."""
    import sys
    import vegs_b200
    ref_mod = util.import_reference()["diff_gaussian_rasterization"]
    prev = {k: sys.modules.get(k) for k in ("diff_gaussian_rasterization", "simple_knn", "simple_knn._C")}
    vegs_b200.install()
    ours_mod = sys.modules["diff_gaussian_rasterization"]
    for k, v in prev.items():
        if v is None:
            sys.modules.pop(k, None)
        else:
            sys.modules[k] = v
    assert ours_mod is not ref_mod

    scene = util.scene_with_specials(200_000)
    cams = [synth.make_camera(k, scale=0.5) for k in range(2)]
    g = torch.Generator().manual_seed(5)
    gts = [torch.rand(3, c.H, c.W, generator=g).cuda() for c in cams]
    normals = [torch.nn.functional.normalize(torch.randn(3, c.H, c.W, generator=g), dim=0).cuda() for c in cams]

    pc_r, pk_r = _run_caller(ref_mod, scene, cams, gts, normals)
    pc_r2, _ = _run_caller(ref_mod, scene, cams, gts, normals)
    pc_o, pk_o = _run_caller(ours_mod, scene, cams, gts, normals)

    rec = util.parity_record("unchanged_callers_200k_704x188")
    for i, (a, b) in enumerate(zip(pk_o, pk_r)):
        assert set(a) == set(b)
        assert torch.equal(a["radii"], b["radii"]) and torch.equal(a["visibility_filter"], b["visibility_filter"])
        for k in ("render", "render_depth", "render_cov_quat", "render_cov_scale", "alpha"):
            assert float((a[k] - b[k]).abs().max()) <= 1e-5, (i, k)
        util.grad_close(a["viewspace_grad"].cpu().numpy(), b["viewspace_grad"].cpu().numpy(),
                        f"iteration {i} viewspace_points.grad", rtol=1e-4)
        assert abs(float(a["loss"]) - float(b["loss"])) <= 1e-6
    grads = rec.setdefault("raw_parameter_grads_after_2_iterations", {})
    for name, leaf in pc_o.leaves().items():
        util.grad_check_full(leaf.grad.cpu().numpy(), pc_r.leaves()[name].grad.cpu().numpy(),
                             pc_r2.leaves()[name].grad.cpu().numpy(), name, rtol=1e-4, record=grads)
    assert torch.equal(pc_o.denom, pc_r.denom) and torch.equal(pc_o.max_radii2D, pc_r.max_radii2D)
    util.grad_close(pc_o.xyz_gradient_accum.cpu().numpy(), pc_r.xyz_gradient_accum.cpu().numpy(),
                    "xyz_gradient_accum", rtol=1e-4)


def test_giant_anisotropic_offscreen_splats():
    """Large, elongated Gaussians whose centres lie hundreds of pixels outside the tile they reach
    (the row-interval cull's fixed margins vs float rounding of terms of 1e5-1e6)."""
    base = synth.make_scene(20_000, seed=9)
    rs = np.random.RandomState(4)
    n = 400
    means = np.stack([rs.uniform(14, 60, n), rs.uniform(-40, 40, n), rs.uniform(-3, 8, n)], 1).astype(np.float32)
    scales = np.stack([rs.uniform(0.002, 0.02, n), rs.uniform(2.0, 12.0, n), rs.uniform(0.01, 0.3, n)], 1).astype(np.float32)
    q = rs.standard_normal((n, 4)).astype(np.float32)
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    extra = dict(means3D=torch.from_numpy(means), scales=torch.from_numpy(scales), rotations=torch.from_numpy(q),
                 opacities=torch.from_numpy(rs.uniform(0.3, 0.99, (n, 1)).astype(np.float32)),
                 shs=torch.from_numpy((rs.standard_normal((n, 16, 3)) * 0.2).astype(np.float32)))
    scene = {k: torch.cat([base[k], extra[k]], 0).contiguous() for k in base}
    for k, scale in ((0, 1.0), (1, 2.0)):
        cam = synth.make_camera(k, scale=scale)
        ours = util.run_ours(scene, cam)
        ref = util.run_reference(scene, cam)
        _check_state_exact(ours, ref, ref["radii"] > 0)
        _check_images(ours, ref, strict=True)
