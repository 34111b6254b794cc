"""CPU tests: the C-ABI library loads and exports every declared symbol, the drop-in modules keep
the reference's interface and error behaviour, and the synthetic cameras follow the reference's
matrix conventions.  No compute calls (no GPU here)."""
import ctypes
import inspect
import os
import re

import numpy as np
import pytest
import torch

import util

ROOT = util.ROOT


def _declared_functions():
    hdr = open(os.path.join(ROOT, "include", "vegs_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(vegs_[a-z0-9_]+)\s*\(", hdr)))


def test_library_exports_every_declared_symbol():
    from vegs_b200 import _lib
    lib = _lib.load()
    names = _declared_functions()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/vegs_b200.h but not exported"
    assert set(names) == set(_lib.SIGNATURES), "ctypes signature table out of sync with the header"
    import re
    hdr = open(os.path.join(ROOT, "include", "vegs_b200.h")).read()
    assert lib.vegs_abi_version() == int(re.search(r"#define VEGS_ABI_VERSION (\d+)", hdr).group(1))


def test_struct_layout_matches_header_order():
    from vegs_b200 import _lib
    hdr = open(os.path.join(ROOT, "include", "vegs_b200.h")).read()
    for sname, cls in (("VegsForwardArgs", _lib.VegsForwardArgs), ("VegsBackwardArgs", _lib.VegsBackwardArgs)):
        body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (sname, sname), hdr, re.S).group(1)
        body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
        fields = []
        for decl in body.split(";"):
            decl = decl.strip()
            if not decl:
                continue
            names = decl.split(None, 1)[1] if " " in decl else decl
            for part in decl.replace("*", " ").split(",") if False else [decl]:
                pass
            # "const float* a" / "int32_t W, H" / "float tan_fovx, tan_fovy"
            tail = re.sub(r"^(const\s+)?(float|int32_t|uint32_t|void|uint8_t)\s*\**", "", decl).strip()
            fields += [x.strip().lstrip("*").strip() for x in tail.split(",")]
        assert fields == [f[0] for f in cls._fields_], sname


def test_size_queries_are_monotone_and_aligned():
    from vegs_b200 import _lib
    lib = _lib.load()
    prev = 0
    for P in (0, 1, 1000, 100000, 6000000):
        b = lib.vegs_geom_bytes(P)
        assert b >= prev and b >= 49 * P
        prev = b
    assert lib.vegs_img_bytes(1408, 376) >= 1408 * 376 * 4 + 2112 * 8
    assert lib.vegs_binning_bytes(10) >= 40
    assert lib.vegs_bwd_scratch_bytes(1000) >= 48000
    assert lib.vegs_dist2_workspace_bytes(1000) > 0


def test_invalid_arguments_return_einval_not_abort():
    from vegs_b200 import _lib
    lib = _lib.load()
    a = _lib.VegsForwardArgs()
    a.P, a.W, a.H, a.D = 10, 0, 10, 3
    word = (ctypes.c_int32 * 2)(-1, -1)
    a.num_rendered_host = ctypes.addressof(word)
    assert lib.vegs_forward_prepare(ctypes.byref(a), None) == 1
    assert b"bad sizes" in lib.vegs_last_error()
    a.W = 16
    assert lib.vegs_forward_prepare(ctypes.byref(a), None) == 1   # null required pointers
    a.P = 0
    assert lib.vegs_forward_prepare(ctypes.byref(a), None) == 0 and word[0] == 0
    assert lib.vegs_dist2(5, None, None, None, 0, None) == 1
    assert lib.vegs_mark_visible(5, None, None, None, None, None) == 1
    with pytest.raises(ValueError):
        _lib.check(1, "x")
    with pytest.raises(RuntimeError):
        _lib.check(2, "x")


def test_dropin_interface_matches_reference():
    import vegs_b200
    vegs_b200.install()
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    from simple_knn._C import distCUDA2
    assert GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix",
        "projmatrix", "sh_degree", "campos", "prefiltered", "debug")
    sig = inspect.signature(GaussianRasterizer.forward)
    assert list(sig.parameters) == ["self", "means3D", "means2D", "opacities", "shs", "colors_precomp",
                                    "scales", "rotations", "cov3D_precomp"]
    assert callable(distCUDA2) and hasattr(GaussianRasterizer, "markVisible")


def _settings():
    from vegs_b200.diff_gaussian_rasterization import GaussianRasterizationSettings
    z = torch.zeros
    return GaussianRasterizationSettings(16, 16, 1.0, 1.0, z(3), 1.0, torch.eye(4), torch.eye(4), 3, z(3),
                                         False, False)


def test_argument_validation_raises_like_reference():
    from vegs_b200.diff_gaussian_rasterization import GaussianRasterizer
    r = GaussianRasterizer(_settings())
    m = torch.zeros(4, 3)
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        r(m, m, torch.zeros(4, 1), scales=m, rotations=torch.zeros(4, 4))
    with pytest.raises(Exception, match="SHs or precomputed colors"):
        r(m, m, torch.zeros(4, 1), shs=torch.zeros(4, 16, 3), colors_precomp=m, scales=m, rotations=torch.zeros(4, 4))
    with pytest.raises(Exception, match="scale/rotation pair"):
        r(m, m, torch.zeros(4, 1), shs=torch.zeros(4, 16, 3), scales=m)
    with pytest.raises(Exception, match="scale/rotation pair"):
        r(m, m, torch.zeros(4, 1), shs=torch.zeros(4, 16, 3), scales=m, rotations=torch.zeros(4, 4),
          cov3D_precomp=torch.zeros(4, 6))


def test_no_cpu_fallback():
    """CPU tensors are refused loudly: the product has no CPU path."""
    from vegs_b200.diff_gaussian_rasterization import GaussianRasterizer
    from vegs_b200.simple_knn._C import distCUDA2
    r = GaussianRasterizer(_settings())
    m = torch.zeros(4, 3)
    with pytest.raises(RuntimeError, match="no CPU path"):
        r(m, m, torch.zeros(4, 1), shs=torch.zeros(4, 16, 3), scales=m, rotations=torch.zeros(4, 4))
    with pytest.raises(RuntimeError, match="no CPU path"):
        r.markVisible(m)
    with pytest.raises(RuntimeError, match="no CPU path"):
        distCUDA2(m)


def test_product_does_not_import_oracle():
    """The oracle is test infrastructure; nothing under vegs_b200/ may reference it."""
    pkg = os.path.join(ROOT, "vegs_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, f)).read()
                assert "import oracle" not in src and "from oracle" not in src and "libvegs_oracle" not in src, f


def test_camera_conventions():
    """scene/cameras.py:70-88: view is stored transposed, full = view @ proj, campos = inverse(view)[3,:3]."""
    from vegs_b200 import synth
    cam = synth.make_camera(2)
    assert (cam.W, cam.H) == (1408, 376)
    assert np.allclose(cam.campos.numpy(), [20.0, 0.0, 1.55], atol=1e-5)
    # a point 10 m ahead on the optical axis has view z = 10 and projects near the principal point
    p = torch.tensor([30.0, 0.0, 1.55, 1.0])
    pv = p @ cam.viewmatrix
    assert abs(pv[2].item() - 10.0) < 1e-4 and abs(pv[0].item()) < 1e-5
    ph = p @ cam.projmatrix
    ndc = ph[:2] / ph[3]
    px = ((ndc[0] + 1) * cam.W - 1) * 0.5
    py = ((ndc[1] + 1) * cam.H - 1) * 0.5
    assert abs(px.item() - (synth.KITTI360["cx"] - 0.5)) < 0.05
    assert abs(py.item() - (synth.KITTI360["cy"] - 0.5)) < 0.05
    assert abs(cam.tanfovx - 704.0 / synth.KITTI360["fx"]) < 1e-6
    small = synth.make_camera(0, scale=0.25)
    assert (small.W, small.H) == (352, 94)
    # extrapolated view: yaw keeps the centre, tz raises it (scene/cameras.py:178-227)
    ex = synth.make_camera(0, add_zrot=30.0, add_xrot=-10.0, add_tz=1.0)
    assert np.allclose(ex.campos.numpy(), [10.0, 0.0, 2.55], atol=1e-5)


def test_scene_generator_is_seeded_and_shaped():
    from vegs_b200 import synth
    a = synth.make_scene(5000, seed=7)
    b = synth.make_scene(5000, seed=7)
    for k in a:
        assert torch.equal(a[k], b[k])
    assert a["means3D"].shape == (5000, 3) and a["shs"].shape == (5000, 16, 3)
    assert torch.allclose(a["rotations"].norm(dim=1), torch.ones(5000), atol=1e-5)
    assert (a["rotations"][:, 0] >= 0).all() and (a["opacities"] > 0).all() and (a["opacities"] < 1).all()
    assert a["scales"][:, 0].max() < 1e-4 and a["scales"][:, 1:].min() > 0.01


def test_entry_scripts_compile():
    """bench.py / __graft_entry__.py / the scripts are only run on the GPU box: catch syntax errors here."""
    import py_compile
    paths = [os.path.join(ROOT, "bench.py"), os.path.join(ROOT, "__graft_entry__.py")]
    paths += [os.path.join(ROOT, "scripts", f) for f in os.listdir(os.path.join(ROOT, "scripts")) if f.endswith(".py")]
    for path in paths:
        py_compile.compile(path, doraise=True)
