"""GPU test of vegs_dist2_noalloc, the distCUDA2 entry point for callers that may not allocate
device memory (a host that owns every allocation and gives the library a fixed budget, or a loop
that cannot afford the device-wide stall of cudaMalloc / cudaFree).  Its scratch is an array built
into the library instead of a caller-provided workspace.

The file name sorts last on purpose: if this test fails in a run that stops at the first failure,
every other test has already run.  It must also pass on its own in a fresh process, where its call
is the first thing that touches the library on the device.

Checked through the C ABI: the result equals vegs_dist2's bit for bit (the tolerance the distCUDA2
tests use) on two calls in a row — the second call runs on the scratch the first one left behind —
and with a different, smaller input afterwards; free device memory does not change across a call;
sizes beyond the built-in maximum are refused with a clear error."""
import ctypes as C

import numpy as np
import pytest
import torch

import util
from vegs_b200 import _lib, synth

pytestmark = pytest.mark.gpu


def _dist2_workspace(lib, pts):
    P = pts.shape[0]
    out = torch.full((P,), -1.0, device=pts.device)
    n = lib.vegs_dist2_workspace_bytes(P)
    ws = torch.empty(n, dtype=torch.uint8, device=pts.device)
    stream = torch.cuda.current_stream()
    _lib.check(lib.vegs_dist2(P, pts.data_ptr(), out.data_ptr(), ws.data_ptr(), n,
                              C.c_void_p(stream.cuda_stream)), "vegs_dist2")
    torch.cuda.synchronize()
    return out.cpu().numpy()


def _dist2_noalloc(lib, pts):
    P = pts.shape[0]
    out = torch.full((P,), -1.0, device=pts.device)
    stream = torch.cuda.current_stream()
    _lib.check(lib.vegs_dist2_noalloc(P, pts.data_ptr(), out.data_ptr(), C.c_void_p(stream.cuda_stream)),
               "vegs_dist2_noalloc")
    torch.cuda.synchronize()
    return out.cpu().numpy()


def test_dist2_noalloc_matches_dist2_on_consecutive_calls():
    lib = _lib.load()
    torch.cuda.set_device(0)
    pts_a = synth.make_scene(100000, seed=21)["means3D"].cuda().contiguous()
    pts_b = synth.make_scene(30000, seed=22)["means3D"].cuda().contiguous()
    first = _dist2_noalloc(lib, pts_a)           # first call on the device: does the one-time set-up
    ref_a = _dist2_workspace(lib, pts_a)
    util.assert_bit_equal(first, ref_a, "first call")
    free_before = torch.cuda.mem_get_info()[0]
    second = _dist2_noalloc(lib, pts_a)          # runs on the first call's leftovers
    assert torch.cuda.mem_get_info()[0] == free_before, "the call changed the amount of free device memory"
    util.assert_bit_equal(second, ref_a, "second call")
    third = _dist2_noalloc(lib, pts_b)           # different, smaller input on the same scratch
    util.assert_bit_equal(third, _dist2_workspace(lib, pts_b), "smaller input after a larger one")


def test_dist2_noalloc_rejects_oversized_input():
    lib = _lib.load()
    torch.cuda.set_device(0)
    too_many = 131072 + 1
    pts = torch.zeros((too_many, 3), device="cuda")
    out = torch.zeros((too_many,), device="cuda")
    rc = lib.vegs_dist2_noalloc(too_many, pts.data_ptr(), out.data_ptr(), None)
    assert rc == 3
    msg = lib.vegs_last_error().decode()
    assert "131072" in msg and "vegs_dist2" in msg
    exact = torch.rand((131072, 3), device="cuda")
    util.assert_bit_equal(_dist2_noalloc(lib, exact), _dist2_workspace(lib, exact), "exactly the maximum")
