"""CPU tests of the view-parallel step (vegs_b200/dist.py) with the gloo backend, world size 2:
the all-reduced gradients and densification statistics of a sharded step equal those of the same
views processed by one rank.  The rasterizer itself needs a GPU, so a stand-in per-view function
with the same signature of results (gradients per Gaussian, radii, screen-space gradient) is used:
what is under test is the sharding, packing and reduction logic."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from vegs_b200 import dist as vdist


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _make_params(P, M=4, seed=0):
    g = torch.Generator().manual_seed(seed)
    shapes = dict(means3D=(P, 3), shs=(P, M, 3), opacities=(P, 1), scales=(P, 3), rotations=(P, 4))
    return {k: torch.randn(*s, generator=g).requires_grad_(True) for k, s in shapes.items()}


def _fake_view(params, v):
    """Deterministic stand-in for one view's forward+backward: a loss that touches every parameter,
    a radii vector with culled entries, and a screen-space gradient."""
    P = params["means3D"].shape[0]
    g = torch.Generator().manual_seed(100 + v)
    w = {k: torch.randn(*t.shape, generator=g) for k, t in params.items()}
    loss = sum((params[k] * w[k]).sum() * (v + 1) for k in params)
    loss.backward()
    radii = torch.randint(0, 5, (P,), generator=g, dtype=torch.int32)
    view_grad = torch.randn(P, 3, generator=g)
    return radii, view_grad


def _run_views(params, views):
    stats = vdist.DensificationStats(params["means3D"].shape[0], "cpu")
    for t in params.values():
        t.grad = None
    for v in views:
        radii, vg = _fake_view(params, v)
        stats.add_view(vg, radii)
    return stats


def _worker(rank, world, port, P, n_views, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    params = _make_params(P)
    stats = _run_views(params, vdist.shard_views(n_views, rank, world))
    # the per-tensor route (NCCL's default) on a copy of the same step must give the same result
    params2 = _make_params(P)
    stats2 = _run_views(params2, vdist.shard_views(n_views, rank, world))
    assert vdist.allreduce_step(params2, stats2, per_tensor=True) is None
    flat = vdist.allreduce_step(params, stats)
    assert flat.numel() == vdist.flat_size(params)
    for k in params:
        assert torch.allclose(params[k].grad, params2[k].grad, rtol=1e-6, atol=1e-6), k
    assert torch.equal(stats.denom, stats2.denom) and torch.equal(stats.max_radii2D, stats2.max_radii2D)
    # the single-view SH factor exchange is a CUDA, per-tensor-route option: refused (on every rank,
    # before any collective) on the packed route and for host tensors
    import pytest
    with pytest.raises(ValueError, match="per-tensor"):
        vdist.allreduce_step(params2, stats2, single_view_campos=torch.zeros(world, 3))
    with pytest.raises(ValueError, match="CUDA"):
        vdist.allreduce_step(params2, stats2, per_tensor=True, single_view_campos=torch.zeros(world, 3))
    if rank == 0:
        torch.save(dict(grads={k: v.grad.clone() for k, v in params.items()},
                        accum=stats.xyz_gradient_accum.clone(), denom=stats.denom.clone(),
                        maxr=stats.max_radii2D.clone()), out)
    dist.destroy_process_group()


def test_shard_views_round_robin():
    assert vdist.shard_views(8, 0, 1) == list(range(8))
    assert vdist.shard_views(8, 1, 2) == [1, 3, 5, 7]
    assert vdist.shard_views(8, 3, 8) == [3]
    assert sorted(sum((vdist.shard_views(8, r, 4) for r in range(4)), [])) == list(range(8))


def test_densification_stats_follow_reference():
    """train.py:300-301 and scene/gaussian_model.py:411-413."""
    st = vdist.DensificationStats(4, "cpu")
    radii = torch.tensor([0, 3, 7, 0], dtype=torch.int32)
    vg = torch.tensor([[3.0, 4.0, 9.0], [6.0, 8.0, 9.0], [0.0, 1.0, 9.0], [1.0, 1.0, 1.0]])
    st.add_view(vg, radii)
    st.add_view(vg, torch.tensor([2, 1, 0, 0], dtype=torch.int32))
    assert st.max_radii2D.tolist() == [2.0, 3.0, 7.0, 0.0]
    assert st.denom.view(-1).tolist() == [1.0, 2.0, 1.0, 0.0]
    assert torch.allclose(st.xyz_gradient_accum.view(-1), torch.tensor([5.0, 20.0, 1.0, 0.0]))


def test_pack_unpack_roundtrip():
    params = _make_params(50)
    stats = _run_views(params, [0, 1])
    ref = {k: v.grad.clone() for k, v in params.items()}
    accum, denom = stats.xyz_gradient_accum.clone(), stats.denom.clone()
    flat = vdist.pack(params, stats)
    assert flat.numel() == 152 + 600 + 52 + 152 + 200 + 52 + 52     # segments padded to multiples of 4 floats
    vdist.unpack(flat, params, stats)
    for k in params:
        assert torch.equal(params[k].grad, ref[k])
    assert torch.equal(stats.xyz_gradient_accum, accum) and torch.equal(stats.denom, denom)


def test_two_rank_step_equals_single_rank(tmp_path):
    P, n_views = 200, 8
    out = str(tmp_path / "r0.pt")
    port = _free_port()
    mp.spawn(_worker, args=(2, port, P, n_views, out), nprocs=2, join=True)
    got = torch.load(out)
    params = _make_params(P)
    stats = _run_views(params, list(range(n_views)))
    for k in params:
        assert torch.allclose(got["grads"][k], params[k].grad, rtol=1e-5, atol=1e-5), k
    assert torch.allclose(got["accum"], stats.xyz_gradient_accum, rtol=1e-6, atol=1e-6)
    assert torch.equal(got["denom"], stats.denom)
    assert torch.equal(got["maxr"], stats.max_radii2D)


def test_densification_stats_kwargs_fall_back_for_modules_without_the_extension():
    """rasterizer_kwargs() hands the buffers to a module that offers per-instance extensions and
    leaves add_view as the path for one that does not (the reference's module)."""
    P = 16

    class WithExtension:
        class RasterizerOptions:
            pass

    class Reference:
        pass

    radii = torch.tensor([0, 3] * (P // 2), dtype=torch.int32)
    grad = torch.ones(P, 3)

    s = vdist.DensificationStats(P, "cpu")
    assert s.rasterizer_kwargs(Reference) == {} and s.fused is False
    ids = (id(s.xyz_gradient_accum), id(s.denom), id(s.max_radii2D))
    s.add_view(grad, radii)
    assert ids == (id(s.xyz_gradient_accum), id(s.denom), id(s.max_radii2D)), "tensors keep their identity"
    assert float(s.denom.sum()) == P // 2 and float(s.max_radii2D.max()) == 3.0
    assert torch.allclose(s.xyz_gradient_accum[radii > 0], torch.full((P // 2, 1), 2.0 ** 0.5))

    f = vdist.DensificationStats(P, "cpu")
    kw = f.rasterizer_kwargs(WithExtension)
    assert f.fused and kw["densification_accumulators"][0] is f.xyz_gradient_accum
    assert kw["densification_accumulators"][2] is f.max_radii2D
    f.add_view(grad, radii)                      # the kernel would have done it: no-op
    assert float(f.denom.sum()) == 0.0


def test_rasterizer_options_validate_their_arguments():
    import pytest
    import vegs_b200.diff_gaussian_rasterization as dgr
    cpu = torch.zeros(4, 1)
    with pytest.raises(ValueError):
        dgr.RasterizerOptions(densification_accumulators=(cpu, cpu, torch.zeros(4)))     # not CUDA tensors
    with pytest.raises(ValueError):
        dgr.RasterizerOptions(fused_grad_accumulation="sometimes")
    assert not dgr.RasterizerOptions().any
    assert dgr.RasterizerOptions(True).fused_grad_accumulation == "add"
    assert dgr._accum_target(torch.zeros(3, requires_grad=True), (3,), "add") is None   # leaf without .grad yet
    # extensions are a property of the instance: a plain rasterizer next to an extended one has none
    z = torch.zeros
    rs = dgr.GaussianRasterizationSettings(16, 16, 1.0, 1.0, z(3), 1.0, torch.eye(4), torch.eye(4), 3, z(3), False, False)
    a, b = dgr.GaussianRasterizer(rs, fused_grad_accumulation=True), dgr.GaussianRasterizer(rs)
    assert a.options.any and not b.options.any
    assert not hasattr(dgr, "set_fused_grad_accumulation") and not hasattr(dgr, "set_densification_accumulators")


def _worker_two_steps(rank, world, port, P, n_views, out):
    """Two optimisation steps with a persistent FlatGradBuffer and a running DensificationStats:
    .grad tensors are views of the flat buffer across steps, the per-step statistics are reduced as
    deltas and added to the running accumulator once (not world-size times)."""
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    params = _make_params(P)
    buf = vdist.FlatGradBuffer(params)
    running = vdist.DensificationStats(P, "cpu")
    flat_ptr = buf.flat.data_ptr()
    grads = []
    for step in range(2):
        buf.begin_step()
        kw = buf.rasterizer_kwargs(object, first=True)      # a module without extensions
        assert kw == {}
        for v in vdist.shard_views(n_views, rank, world):
            radii, vg = _fake_view(params, v + 10 * step)
            buf.step_stats.add_view(vg, radii)
        assert params["shs"].grad.data_ptr() == buf.grad_views["shs"].data_ptr()
        buf.allreduce(accumulate_into=running)
        assert buf.flat.data_ptr() == flat_ptr
        grads.append({k: v.grad.clone() for k, v in params.items()})
    # the compatibility path with a reused flat buffer whose slices are already the .grad tensors
    stats = vdist.DensificationStats(P, "cpu")
    flat = vdist.allreduce_step(params, stats, flat=None)
    for t in params.values():
        t.grad.zero_()
    _fake_view(params, 99)
    flat2 = vdist.allreduce_step(params, stats, flat=flat)     # inputs alias `flat`: must not raise
    assert flat2.data_ptr() == flat.data_ptr()
    if rank == 0:
        torch.save(dict(grads=grads, accum=running.xyz_gradient_accum.clone(), denom=running.denom.clone(),
                        maxr=running.max_radii2D.clone()), out)
    dist.destroy_process_group()


def test_two_steps_with_persistent_buffers(tmp_path):
    P, n_views = 120, 4
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker_two_steps, args=(2, _free_port(), P, n_views, out), nprocs=2, join=True)
    got = torch.load(out)
    params = _make_params(P)
    running = vdist.DensificationStats(P, "cpu")
    for step in range(2):
        stats = _run_views(params, [v + 10 * step for v in range(n_views)])
        for k in params:
            assert torch.allclose(got["grads"][step][k], params[k].grad, rtol=1e-5, atol=1e-5), (step, k)
        running.xyz_gradient_accum += stats.xyz_gradient_accum
        running.denom += stats.denom
        running.max_radii2D = torch.maximum(running.max_radii2D, stats.max_radii2D)
    assert torch.allclose(got["accum"], running.xyz_gradient_accum, rtol=1e-6, atol=1e-6)
    assert torch.equal(got["denom"], running.denom)
    assert torch.equal(got["maxr"], running.max_radii2D)


def test_sh_factor_capture_keeps_each_pass_and_leaves_autograd_alone():
    """The hook behind allreduce_step(sh_factor=...): every backward pass leaves grad[:, 0, :] / SH_C0 in
    the next slot, `.grad` accumulates as if the hook were not there, a pass too many is refused,
    and nothing is captured outside begin_step() .. end_step()."""
    import pytest
    g = torch.Generator().manual_seed(5)
    shs = torch.randn(50, 4, 3, generator=g).requires_grad_(True)
    w = [torch.randn(50, 4, 3, generator=g) for _ in range(3)]
    cap = vdist.ShFactorCapture(shs, slots=2)
    (shs * w[0]).sum().backward()                       # before begin_step: not captured
    assert cap.n == 0 and float(cap.buf.abs().max()) == 0.0
    shs.grad = None
    cap.begin_step()
    (shs * w[0]).sum().backward()
    (shs * w[1]).sum().backward()
    assert cap.n == 2
    assert torch.allclose(cap.buf[0], w[0][:, 0, :] / vdist.SH_C0) and torch.allclose(cap.buf[1], w[1][:, 0, :] / vdist.SH_C0)
    assert torch.allclose(shs.grad, w[0] + w[1])
    with pytest.raises(RuntimeError, match="more than 2"):
        (shs * w[2]).sum().backward()
    cap.end_step()
    (shs * w[2]).sum().backward()                       # after end_step: not captured, no complaint
    assert cap.n == 2
    cap.begin_step()
    assert cap.n == 0 and float(cap.buf.abs().max()) == 0.0
    cap.remove()
    with pytest.raises(ValueError):
        vdist.ShFactorCapture(torch.zeros(5, 3, requires_grad=True), slots=1)


def test_slot_campos_follows_the_round_robin_sharding():
    """View j of the batch is rank j % world's (j // world)-th view: its camera centre must sit in
    row (j % world) * slots + j // world of the all-gathered order; rows of unused slots are zero."""
    V, world, slots = 5, 2, 3
    vc = torch.arange(V * 3, dtype=torch.float32).reshape(V, 3) + 1.0
    out = vdist._slot_campos(vc, world, slots, "cpu")
    assert out.shape == (world * slots, 3)
    for r in range(world):
        mine = vdist.shard_views(V, r, world)
        for i in range(slots):
            want = vc[mine[i]] if i < len(mine) else torch.zeros(3)
            assert torch.equal(out[r * slots + i], want), (r, i)
    assert torch.equal(vdist._slot_campos([vc[j] for j in range(V)], world, slots, "cpu"), out)
    import pytest
    with pytest.raises(ValueError):
        vdist._slot_campos(vc, 2, 2, "cpu")
