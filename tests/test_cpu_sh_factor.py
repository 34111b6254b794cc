"""The identity behind the SH factor exchange (vegs_b200/dist.py: ShFactorCapture,
FlatGradBuffer(defer_sh=...), csrc/sh_views.cu), checked on the CPU against the oracle's restatement
of the reference's SH backward (cuda_rasterizer/backward.cu:21-141):

    dL/dsh[k][c] of one view  =  basis_k(normalize(mean - campos)) * dL/dRGB[c],   dL/dRGB = dL/dsh[0] / SH_C0

so the SH gradient of a step is rebuilt from the DC rows of the per-view gradients and the views'
camera centres.  Clamped channels (colour < 0) and Gaussians a view does not render have a zero
DC row and contribute nothing, as in the reference."""
import numpy as np

import util
from vegs_b200 import synth
from vegs_b200.dist import SH_C0

SH_C1 = 0.4886025119029199
SH_C2 = (1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396)
SH_C3 = (-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
         1.445305721320277, -0.5900435899266435)


def sh_basis(dirs, degree):
    """[P, (degree+1)^2] real SH basis in the reference's sign convention (forward.cu:21-72)."""
    x, y, z = dirs[:, 0], dirs[:, 1], dirs[:, 2]
    cols = [np.full_like(x, SH_C0)]
    if degree > 0:
        cols += [-SH_C1 * y, SH_C1 * z, -SH_C1 * x]
    if degree > 1:
        xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
        cols += [SH_C2[0] * xy, SH_C2[1] * yz, SH_C2[2] * (2 * zz - xx - yy), SH_C2[3] * xz, SH_C2[4] * (xx - yy)]
    if degree > 2:
        cols += [SH_C3[0] * y * (3 * xx - yy), SH_C3[1] * xy * z, SH_C3[2] * y * (4 * zz - xx - yy),
                 SH_C3[3] * z * (2 * zz - 3 * xx - 3 * yy), SH_C3[4] * x * (4 * zz - xx - yy),
                 SH_C3[5] * z * (xx - yy), SH_C3[6] * x * (xx - 3 * yy)]
    return np.stack(cols, axis=1)


def test_step_sh_gradient_is_rebuilt_from_the_views_dc_rows():
    scene = util.scene_with_specials(4000)
    views = (0, 3, 5)
    total = np.zeros_like(np.asarray(scene["shs"], np.float64))
    rebuilt = np.zeros_like(total)
    rendered_somewhere = np.zeros(total.shape[0], bool)
    for deg in (3, 1):
        total[:] = 0.0; rebuilt[:] = 0.0
        for v in views:
            cam = synth.make_camera(v, scale=0.125)
            cot = util.np_cotangents(cam.W, cam.H)
            g = util.run_oracle(scene, cam, sh_degree=deg, cot=cot)["grads"]["sh"].astype(np.float64)
            total += g
            drgb = g[:, 0, :] / SH_C0                                            # what the hook keeps
            d = np.asarray(scene["means3D"], np.float64) - np.asarray(cam.campos, np.float64)[None, :]
            d /= np.linalg.norm(d, axis=1, keepdims=True)
            basis = sh_basis(d, deg)                                             # [P, used]
            rebuilt[:, :basis.shape[1], :] += basis[:, :, None] * drgb[:, None, :]
            rendered_somewhere |= np.abs(drgb).sum(1) > 0
        assert rendered_somewhere.sum() > 500
        scale = np.abs(total).max()
        assert scale > 0
        assert np.abs(rebuilt - total).max() <= 2e-6 * scale, (deg, np.abs(rebuilt - total).max() / scale)
        # coefficients above the active degree get no gradient, in either form
        used = (deg + 1) ** 2
        assert np.abs(total[:, used:, :]).max(initial=0.0) == 0.0 and np.abs(rebuilt[:, used:, :]).max(initial=0.0) == 0.0
