"""CPU tests of the view-parallel training driver (vegs_b200/parallel_train.py) and the model it
drives (vegs_b200/splat_model.py), gloo backend, world size 2.  The rasterizer needs a GPU, so a
stand-in per-view function with the same shape of results is used: what is under test is the
driver - view sharding, the flat all-reduce buffer across steps, identical densification / pruning /
opacity reset on every rank, the replica check, the rank-0 checkpoint - and the model's structural
operations against the reference's rules (scene/gaussian_model.py:253-257, :309-409)."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from vegs_b200 import synth
from vegs_b200.parallel_train import ViewParallelTrainer
from vegs_b200.splat_model import SplatModel, TrainingArgs


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


ARGS = dict(densify_from_iter=1, densification_interval=2, densify_until_iter=7, opacity_reset_interval=3,
            densify_grad_threshold=7.0, percent_dense=0.05)


def _model(P=240):
    sc = synth.make_scene(P, seed=5)
    sc["scales"][::3] *= 40.0          # some Gaussians above the split limit
    m = SplatModel(sc, device="cpu")
    m.training_setup(TrainingArgs(**ARGS))
    return m


def _fake_render(model, camera, rasterizer_kwargs):
    """Stand-in for render + loss: touches every parameter through its activated getter; the
    screen-space gradient is integer-valued with norms that are whole numbers, so the densification
    statistics are exact whatever the order of the additions (1 rank vs 2)."""
    assert rasterizer_kwargs == {}
    P = model.num_points
    g = torch.Generator().manual_seed(1000 + camera * 7919 + P)
    k = torch.randint(0, 4, (P, 1), generator=g).float()
    w2d = torch.cat([3 * k, 4 * k, torch.ones(P, 1)], 1)
    radii = torch.randint(0, 6, (P,), generator=g, dtype=torch.int32)
    viewspace = torch.zeros(P, 3, requires_grad=True)
    wx = torch.randn(P, 3, generator=g)
    loss = ((model.get_xyz * wx).sum() + (model.get_features * 0.1).sum() * (camera + 1) + model.get_opacity.sum()
            + (model.get_scaling * wx.abs()).sum() + (model.get_rotation * 0.3).sum() + (viewspace * w2d).sum())
    return loss, viewspace, radii


def _run(rank, world, port, iters, out, ckpt_dir):
    if world > 1:
        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
    m = _model()
    tr = ViewParallelTrainer(m, cameras=list(range(6)), render_view=_fake_render, args=TrainingArgs(**ARGS),
                             views_per_step=4, cameras_extent=4.0, seed=3, checkpoint_dir=ckpt_dir,
                             checkpoint_iterations=(4,))
    log = tr.run(iters)
    if rank == 0:
        torch.save(dict(params={k: v.detach().clone() for k, v in m.parameters().items()}, log=log,
                        accum=m.xyz_gradient_accum.clone(), denom=m.denom.clone(), maxr=m.max_radii2D.clone()), out)
    if world > 1:
        dist.destroy_process_group()


def test_two_rank_training_equals_single_rank(tmp_path):
    iters = 6
    one, two = str(tmp_path / "one.pt"), str(tmp_path / "two.pt")
    _run(0, 1, 0, iters, one, str(tmp_path / "ck1"))
    mp.spawn(_run, args=(2, _free_port(), iters, two, str(tmp_path / "ck2")), nprocs=2, join=True)
    a, b = torch.load(one, weights_only=False), torch.load(two, weights_only=False)
    # same batches, disjoint shards, same Gaussian counts after every densification
    for la, lb in zip(a["log"], b["log"]):
        assert la["views"] == lb["views"] and la["P"] == lb["P"]
        assert sorted(lb["local_views"]) == sorted(la["views"][0::2])
        assert la.get("densified_to") == lb.get("densified_to")
    assert any("densified_to" in l for l in a["log"]), "the run must cross a densification"
    assert a["log"][-1]["P"] != a["log"][0]["P"], "densification changed the Gaussian count"
    for k in a["params"]:
        assert a["params"][k].shape == b["params"][k].shape, k
        assert torch.allclose(a["params"][k], b["params"][k], rtol=1e-4, atol=1e-5), k
    assert torch.equal(a["denom"], b["denom"]) and torch.equal(a["maxr"], b["maxr"])
    assert torch.allclose(a["accum"], b["accum"])
    assert os.path.exists(str(tmp_path / "ck2" / "chkpnt4.pth")), "rank 0 writes the checkpoint"
    ck, it = torch.load(str(tmp_path / "ck2" / "chkpnt4.pth"), weights_only=False)
    assert it == 4 and set(ck["params"]) == set(a["params"])


def test_structural_operations_follow_the_reference_rules():
    m = _model(120)
    P0 = m.num_points
    # give every parameter a gradient and one optimiser step so that Adam moments exist
    (m.get_xyz.sum() + m.get_features.sum() + m.get_opacity.sum() + m.get_scaling.sum() + m.get_rotation.sum()).backward()
    m.optimizer.step()
    moments = {g["name"]: m.optimizer.state[g["params"][0]]["exp_avg"].clone() for g in m.optimizer.param_groups}
    keep = torch.ones(P0, dtype=torch.bool); keep[::4] = False
    m.keep_rows(keep)
    assert m.num_points == int(keep.sum()) and m.denom.shape[0] == m.num_points
    for g in m.optimizer.param_groups:
        st = m.optimizer.state[g["params"][0]]
        assert torch.equal(st["exp_avg"], moments[g["name"]][keep]) and g["params"][0] is m.parameters()[g["name"]]
    # appended rows start with zero moments; statistics are cleared
    extra = {n: t.detach()[:5].clone() for n, t in m.parameters().items()}
    m.denom += 1
    m.append_rows(extra)
    assert m.num_points == int(keep.sum()) + 5 and float(m.denom.sum()) == 0
    for g in m.optimizer.param_groups:
        assert float(m.optimizer.state[g["params"][0]]["exp_avg"][-5:].abs().sum()) == 0
    # reset_opacity caps at 0.01 and clears the opacity moments only
    m.reset_opacity()
    assert float(m.get_opacity.detach().max()) <= 0.01 + 1e-6
    st = m.optimizer.state[m._opacity]
    assert float(st["exp_avg"].abs().sum()) == 0 and float(m.optimizer.state[m._xyz]["exp_avg"].abs().sum()) > 0
    # densify: clone where small & high gradient, split (2 for 1) where large & high gradient
    P1 = m.num_points
    m.xyz_gradient_accum[:] = 1.0; m.denom[:] = 1.0
    m.xyz_gradient_accum[10:] = 0.0
    small = (m.get_scaling.max(dim=1).values[:10] <= m.percent_dense * 4.0)
    n_small, n_big = int(small.sum()), int((~small).sum())
    torch.manual_seed(0)
    m.densify_and_prune(0.5, 0.0, 4.0, None)
    assert m.num_points == P1 + n_small + n_big      # +1 per clone, +2 -1 per split


def test_checkpoint_roundtrip():
    m = _model(60)
    (m.get_xyz.sum() + m.get_opacity.sum()).backward()
    m.optimizer.step()
    ck = m.capture()
    m2 = _model(60)
    m2.restore(ck, TrainingArgs(**ARGS))
    for k, v in m.parameters().items():
        assert torch.equal(v, m2.parameters()[k])
    assert torch.equal(m.optimizer.state[m._xyz]["exp_avg"], m2.optimizer.state[m2._xyz]["exp_avg"])
