"""ctypes binding of libvegs_b200.so (the C ABI declared in include/vegs_b200.h).

There is no CPU or PyTorch fallback: if the library has not been built, importing the operator
modules raises.  Build it with `python -m vegs_b200.build` (needs nvcc, no GPU).
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libvegs_b200.so")

c_f32p = C.c_void_p  # device pointers travel as integers


class VegsForwardArgs(C.Structure):
    _fields_ = [
        ("P", C.c_int32), ("D", C.c_int32), ("M", C.c_int32), ("W", C.c_int32), ("H", C.c_int32),
        ("background", C.c_void_p), ("means3D", C.c_void_p), ("shs", C.c_void_p),
        ("colors_precomp", C.c_void_p), ("opacities", C.c_void_p), ("scales", C.c_void_p),
        ("rotations", C.c_void_p), ("cov3D_precomp", C.c_void_p), ("viewmatrix", C.c_void_p),
        ("projmatrix", C.c_void_p), ("campos", C.c_void_p),
        ("scale_modifier", C.c_float), ("tan_fovx", C.c_float), ("tan_fovy", C.c_float),
        ("prefiltered", C.c_int32), ("debug", C.c_int32),
        ("out_color", C.c_void_p), ("out_depth", C.c_void_p), ("out_cov_quat", C.c_void_p),
        ("out_cov_scale", C.c_void_p), ("out_alpha", C.c_void_p), ("radii", C.c_void_p),
        ("geom", C.c_void_p), ("img", C.c_void_p), ("binning", C.c_void_p),
        ("scratch", C.c_void_p), ("bin_scratch", C.c_void_p),
        ("num_rendered_host", C.c_void_p),
        ("activations", C.c_uint32), ("shs_rest", C.c_void_p),
    ]


class VegsBackwardArgs(C.Structure):
    _fields_ = [
        ("P", C.c_int32), ("D", C.c_int32), ("M", C.c_int32), ("W", C.c_int32), ("H", C.c_int32),
        ("R", C.c_int32),
        ("background", C.c_void_p), ("means3D", C.c_void_p), ("shs", C.c_void_p),
        ("colors_precomp", C.c_void_p), ("scales", C.c_void_p), ("rotations", C.c_void_p),
        ("cov3D_precomp", C.c_void_p), ("viewmatrix", C.c_void_p), ("projmatrix", C.c_void_p),
        ("campos", C.c_void_p),
        ("scale_modifier", C.c_float), ("tan_fovx", C.c_float), ("tan_fovy", C.c_float),
        ("debug", C.c_int32),
        ("radii", C.c_void_p), ("out_alpha", C.c_void_p),
        ("dL_dout_color", C.c_void_p), ("dL_dout_depth", C.c_void_p),
        ("dL_dout_cov_quat", C.c_void_p), ("dL_dout_cov_scale", C.c_void_p),
        ("dL_dout_alpha", C.c_void_p),
        ("geom", C.c_void_p), ("img", C.c_void_p), ("binning", C.c_void_p),
        ("bwd_scratch", C.c_void_p),
        ("dL_dmeans3D", C.c_void_p), ("dL_dmeans2D", C.c_void_p), ("dL_dsh", C.c_void_p),
        ("dL_dcolors", C.c_void_p), ("dL_dopacity", C.c_void_p), ("dL_dscales", C.c_void_p),
        ("dL_drotations", C.c_void_p), ("dL_dcov3D", C.c_void_p),
        ("accumulate_mask", C.c_uint32),
        ("stat_grad_accum", C.c_void_p), ("stat_denom", C.c_void_p), ("stat_max_radii", C.c_void_p),
        ("activations", C.c_uint32), ("opacities", C.c_void_p), ("shs_rest", C.c_void_p),
        ("dL_dsh_rest", C.c_void_p),
        ("live_map_gradients", C.c_int32), ("out_cov_quat", C.c_void_p),
        ("dL_drgb_deferred", C.c_void_p),
    ]


ADAM_MAX_SEGMENTS = 12


class VegsAdamSegment(C.Structure):
    _fields_ = [("param", C.c_void_p), ("grad", C.c_void_p), ("exp_avg", C.c_void_p), ("exp_avg_sq", C.c_void_p),
                ("count", C.c_int64), ("lr", C.c_float), ("kind", C.c_int32)]


class VegsAdamArgs(C.Structure):
    _fields_ = [("segments", VegsAdamSegment * ADAM_MAX_SEGMENTS), ("num_segments", C.c_int32), ("step", C.c_int32),
                ("zero_grad", C.c_int32), ("beta1", C.c_double), ("beta2", C.c_double), ("eps", C.c_double)]


# name -> (restype, argtypes); the list doubles as the symbol inventory checked by the tests
SIGNATURES = {
    "vegs_abi_version": (C.c_int, []),
    "vegs_last_error": (C.c_char_p, []),
    "vegs_geom_bytes": (C.c_size_t, [C.c_int32]),
    "vegs_scratch_bytes": (C.c_size_t, [C.c_int32]),
    "vegs_img_bytes": (C.c_size_t, [C.c_int32, C.c_int32]),
    "vegs_binning_bytes": (C.c_size_t, [C.c_int64]),
    "vegs_bin_scratch_bytes": (C.c_size_t, [C.c_int64, C.c_int32, C.c_int32]),
    "vegs_bwd_scratch_bytes": (C.c_size_t, [C.c_int32]),
    "vegs_forward_prepare": (C.c_int, [C.POINTER(VegsForwardArgs), C.c_void_p]),
    "vegs_forward_render": (C.c_int, [C.POINTER(VegsForwardArgs), C.c_int32, C.c_void_p]),
    "vegs_backward": (C.c_int, [C.POINTER(VegsBackwardArgs), C.c_void_p]),
    "vegs_mark_visible": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_void_p]),
    "vegs_dist2_workspace_bytes": (C.c_size_t, [C.c_int32]),
    "vegs_dist2": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t,
                             C.c_void_p]),
    "vegs_dist2_noalloc": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "vegs_adam_step": (C.c_int, [C.POINTER(VegsAdamArgs), C.c_void_p]),
    "vegs_sh_grad_from_views": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                          C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p]),
    "vegs_normal_guidance": (C.c_int, [C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_float,
                                       C.c_float, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "vegs_export_state": (C.c_int, [C.POINTER(VegsForwardArgs), C.c_int32] + [C.c_void_p] * 10),
    "vegs_export_tile_clocks": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "vegs_kernel_launch_count": (C.c_uint64, []),
    "vegs_profile_enable": (None, [C.c_int32]),
    "vegs_debug_unit_masks": (None, [C.c_int32]),
    "vegs_profile_collect": (C.c_int, [C.c_char_p, C.c_size_t]),
}

_lib = None


def load() -> C.CDLL:
    """Load the shared library (once) and type its entry points.  Raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: the CUDA library has not been built. "
            "Run `python -m vegs_b200.build` (requires nvcc). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if a declared symbol is not exported
        fn.restype = res
        fn.argtypes = args
    if lib.vegs_abi_version() != 4:
        raise ImportError("libvegs_b200.so ABI version mismatch; rebuild it")
    _lib = lib
    return lib


def check(rc: int, what: str) -> None:
    """Turn a non-zero return code into a Python exception (never abort)."""
    if rc != 0:
        msg = load().vegs_last_error().decode("utf-8", "replace")
        if rc == 1:
            raise ValueError(f"{what}: {msg}")
        raise RuntimeError(f"{what}: {msg}")


def launch_count() -> int:
    return int(load().vegs_kernel_launch_count())


def profile_enable(on: bool) -> None:
    load().vegs_profile_enable(1 if on else 0)


def profile_collect() -> dict:
    """{kernel name: (total_ms, launches)} since profiling was enabled / last collected."""
    buf = C.create_string_buffer(1 << 16)
    check(load().vegs_profile_collect(buf, len(buf)), "vegs_profile_collect")
    out = {}
    for line in buf.value.decode().splitlines():
        name, ms, n = line.split()
        out[name] = (float(ms), int(n))
    return out
