"""Hardware multi-GPU equivalence: the view-parallel step on 2 GPUs over NCCL with
the REAL rasterizer equals the same 8-view batch accumulated on one GPU, gradients within 1e-4 of
each tensor's max norm, integer statistics exactly.  Both routes are covered: the plain drop-in
step (autograd accumulation + vegs_b200.dist.allreduce_step), the fused one (FlatGradBuffer:
.grad tensors and step statistics as views of the all-reduce buffer, written by the backward
kernel) and the fused one with the deferred SH gradient (an all-gather of each view's dL/dRGB
instead of the all-reduce of the SH gradient), and the drop-in step of ONE view per rank whose SH
gradient is exchanged as its rank-one factor (allreduce_step(single_view_campos=...)), and the same
for several views per rank with the factors kept by a hook on the SH leaf (ShFactorCapture).
Skipped on a box with fewer than 2 GPUs (run it with `gpurun --gpus 2`)."""
import os
import socket

import numpy as np
import pytest
import torch

import util
from vegs_b200 import synth

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")]

P, N_VIEWS, SCALE = 150_000, 8, 0.5


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _view_step(dgr, params, cam, tgt, ext, dev):
    rs = dgr.GaussianRasterizationSettings(cam.H, cam.W, cam.tanfovx, cam.tanfovy, torch.zeros(3, device=dev), 1.0,
                                           cam.viewmatrix, cam.projmatrix, 3, cam.campos, False, False)
    m2 = torch.zeros_like(params["means3D"], requires_grad=True)
    out = dgr.GaussianRasterizer(rs, **ext)(means3D=params["means3D"], means2D=m2, opacities=params["opacities"],
                                            shs=params["shs"], scales=params["scales"], rotations=params["rotations"])
    loss = (out[0] - tgt).abs().mean() + 0.05 * (out[2].abs().mean() + out[3].abs().mean()) + 0.1 * out[1].mean()
    loss.backward()
    return m2.grad, out[5]


def _targets(cams):
    g = torch.Generator().manual_seed(77)
    return [torch.rand(3, c.H, c.W, generator=g) for c in cams]


def _run_step(rank, world, fused, dev):
    """One step (8 views; 2 for the single-view-per-rank route) for `rank` of `world`; returns
    gradients and statistics (CPU tensors)."""
    import torch.distributed as dist
    import vegs_b200.diff_gaussian_rasterization as dgr
    from vegs_b200 import dist as vdist
    sc = util.scene_with_specials(P)
    params = {k: v.to(dev).requires_grad_(True) for k, v in sc.items()}
    n_views = 2 if fused == "single" else N_VIEWS
    cams = [synth.make_camera(k, scale=SCALE, device=dev) for k in range(n_views)]
    tg = [t.to(dev) for t in _targets(cams)]
    mine = vdist.shard_views(n_views, rank, world)
    if fused == "single":
        # plain module, autograd accumulation; with 2 ranks each has one view and the SH gradient
        # travels as dL/dRGB (on one GPU the same call is the plain 2-view accumulation)
        stats = vdist.DensificationStats(params["means3D"].shape[0], dev)
        for v in mine:
            vg, radii = _view_step(dgr, params, cams[v], tg[v], {}, dev)
            stats.add_view(vg, radii)
        assert world == 1 or len(mine) == 1
        vdist.allreduce_step(params, stats, single_view_campos=torch.stack([c.campos for c in cams]) if world > 1 else None)
    elif fused == "hook":
        # plain module, autograd accumulation over the rank's 4 views; a hook on the SH leaf keeps each
        # view's dL/dRGB and the SH gradient travels as those (8 x 3 words per Gaussian instead of 48)
        stats = vdist.DensificationStats(params["means3D"].shape[0], dev)
        cap = vdist.ShFactorCapture(params["shs"], slots=len(mine)) if world > 1 else None
        if cap is not None:
            cap.begin_step()
        for v in mine:
            vg, radii = _view_step(dgr, params, cams[v], tg[v], {}, dev)
            stats.add_view(vg, radii)
        vdist.allreduce_step(params, stats, sh_factor=cap, view_campos=[c.campos for c in cams] if cap is not None else None)
        if cap is not None:
            assert cap.n == len(mine) and not cap.enabled
            cap.remove()
    elif fused:
        defer = fused == "defer"
        buf = vdist.FlatGradBuffer(params, defer_sh=("shs",) if defer else (), views_per_step=N_VIEWS)
        buf.begin_step([c.campos for c in cams] if defer else None)
        for i, v in enumerate(mine):
            vg, radii = _view_step(dgr, params, cams[v], tg[v], buf.rasterizer_kwargs(dgr, first=(i == 0), slot=i), dev)
            buf.step_stats.add_view(vg, radii)
        buf.allreduce()
        stats = buf.step_stats
    else:
        stats = vdist.DensificationStats(params["means3D"].shape[0], dev)
        for v in mine:
            vg, radii = _view_step(dgr, params, cams[v], tg[v], {}, dev)
            stats.add_view(vg, radii)
        vdist.allreduce_step(params, stats)
    torch.cuda.synchronize(dev)
    return dict(grads={k: v.grad.detach().cpu() for k, v in params.items()}, accum=stats.xyz_gradient_accum.cpu(),
                denom=stats.denom.cpu(), maxr=stats.max_radii2D.cpu())


def _worker(rank, world, port, fused, out):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    res = _run_step(rank, world, fused, dev)
    if rank == 0:
        torch.save(res, out)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("fused", [False, True, "defer", "single", "hook"])
def test_two_gpu_step_equals_one_gpu_accumulation(fused, tmp_path):
    import torch.multiprocessing as mp
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker, args=(2, _free_port(), fused, out), nprocs=2, join=True)
    two = torch.load(out, weights_only=False)
    one = _run_step(0, 1, fused, torch.device("cuda", 0))
    rec = util.parity_record(f"multi_gpu_2x_{'fused_' + str(fused) if fused else 'dropin'}_{P}_{2 if fused == 'single' else N_VIEWS}views")
    for k in one["grads"]:
        assert float(one["grads"][k].abs().max()) > 0, k
        rec[k] = float(util.grad_close(two["grads"][k].numpy(), one["grads"][k].numpy(), f"2-GPU all-reduced grad {k}", rtol=1e-4))
    util.grad_close(two["accum"].numpy(), one["accum"].numpy(), "xyz_gradient_accum", rtol=1e-4)
    assert torch.equal(two["denom"], one["denom"]) and torch.equal(two["maxr"], one["maxr"])
    assert float(one["denom"].max()) > 1, "several views saw the same Gaussians"
