#!/usr/bin/env python
"""bench.py — forward+backward views/s through GaussianRasterizer at LiDAR density.

Workload (BASELINE.json config 3): 6 M Gaussians, 1408x376, a step = one 8-view batch of the
synthetic drive split round-robin over the ranks (strong scaling: the batch is fixed), per view a
forward and a backward through the drop-in GaussianRasterizer with a loss on colour + covariance
maps, local gradient accumulation, per-view densification statistics, then one all-reduce
(vegs_b200/dist.py).  `--impl reference` runs the same step with the reference's own CUDA
extension from baseline/_ref (falls back to the CPU oracle port when that cannot be imported); that
process never imports vegs_b200's rasterizer module nor maps libvegs_b200.so.

Both arms time the SAME step: the plain drop-in module, autograd's per-view gradient add, the
caller's densification-statistic passes (`value`, `e2e`).  The vegs_b200 arm additionally reports
`value_fused` / `e2e_fused`: the same step with this package's two per-instance extensions
(gradient accumulation and statistics inside the backward kernel, .grad tensors as views of the
all-reduce buffer), which the reference's callers do not have.

One JSON line on stdout (rank 0).  See DESIGN.md "Measurement" for how each field is obtained.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "fwd+bwd views/s @6M Gaussians 1408x376"
UNIT = "views/s"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "200", "-i", str(gpu_index)], stdout=self.f, stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush(); self.f.seek(0)
        sm, mx, reasons = [], [], set()
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 9:
                continue
            try:
                sm.append(float(c[1])); mx.append(float(c[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if sm:
            out = {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                   "samples": len(sm)}
        try:
            os.unlink(self.f.name)
        except OSError:
            pass
        return out


def load_rasterizer(impl: str):
    """(module, kind): the GaussianRasterizer implementation under test."""
    if impl == "ours":
        import vegs_b200.diff_gaussian_rasterization as m
        return m, "vegs_b200"
    import util
    if util.reference_available():
        try:
            return util.import_reference()["diff_gaussian_rasterization"], "reference-cuda"
        except Exception as ex:  # noqa: BLE001
            sys.stderr.write(f"reference extension not importable: {ex}\n")
    return None, "oracle-port"


class Workload:
    def __init__(self, args, device, rank, world):
        from vegs_b200 import synth, dist as vdist
        self.synth, self.vdist = synth, vdist
        self.device, self.rank, self.world = device, rank, world
        self.P, self.n_views = args.P, args.views
        sc = synth.make_scene(args.P, seed=7)
        self.params = {k: v.to(device).requires_grad_(True) for k, v in sc.items()}
        self.cams = [synth.make_camera(k, scale=args.scale, device=device) for k in range(args.views)]
        self.W, self.H = self.cams[0].W, self.cams[0].H
        g = torch.Generator().manual_seed(123)
        H, W = self.H, self.W
        # per-view step inputs: target image, quaternion-map weights, scale-map weights (10 planes)
        self.host_targets = [torch.rand(10, H, W, generator=g).pin_memory() for _ in range(args.views)]
        self.dev_targets = [t.to(device) for t in self.host_targets]
        self.bg = torch.zeros(3, device=device)
        self.my_views = vdist.shard_views(args.views, rank, world)
        self.flat = None
        self.buf = None
        self.factor_sh = False           # set by main(): vegs_b200 arm at N > 1 - the SH gradient crosses as dL/dRGB
        self.capture = None              # ... kept per view by a hook on the SH leaf when a rank has several views
        self.campos_all = torch.stack([c.campos for c in self.cams]).contiguous()
        self.h2d_bytes = sum(self.host_targets[v].numel() * 4 for v in range(args.views))

    def settings(self, mod, cam):
        return mod.GaussianRasterizationSettings(
            image_height=cam.H, image_width=cam.W, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, bg=self.bg,
            scale_modifier=1.0, viewmatrix=cam.viewmatrix, projmatrix=cam.projmatrix, sh_degree=3,
            campos=cam.campos, prefiltered=False, debug=False)

    def _stage_target(self, v):
        """Host -> device copy of view v's target planes on the copy stream; (tensor, ready event)."""
        if not hasattr(self, "copy_stream"):
            self.copy_stream = torch.cuda.Stream(self.device)
        with torch.cuda.stream(self.copy_stream):
            t = self.host_targets[v % len(self.host_targets)].to(self.device, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.copy_stream)
        t.record_stream(torch.cuda.current_stream(self.device))
        return t, ev

    def step(self, mod, resident: bool, reduce: bool = True, fused: bool = False, views=None):
        """One optimisation step's rasterizer work: local views fwd+bwd, stats, all-reduce.

        fused=False  the plain drop-in module (either arm): autograd adds each view's gradient to
                     .grad, the caller computes the densification statistics per view
                     (train.py:300-301), the all-reduce buffer is packed from the .grad tensors.
        fused=True   vegs_b200's per-instance extensions: .grad tensors and the step's statistics
                     are views of one persistent flat buffer, the backward kernel assigns (first
                     view) / adds (later views) into them and updates the statistics itself.
        """
        p = self.params
        my_views = self.my_views if views is None else views
        if fused:
            if self.buf is None:
                self.buf = self.vdist.FlatGradBuffer(p, defer_sh=("shs",), sh_degree=3, views_per_step=self.n_views)
            # camera centres of the step's views in batch order (the deferred SH gradient is rebuilt from them)
            self.buf.begin_step([self.cams[v].campos for v in range(self.n_views)] if views is None else None)
            stats = self.buf.step_stats
        else:
            for t in p.values():
                t.grad = None
            stats = self.vdist.DensificationStats(self.P, self.device)
            if self.capture is not None and views is None and reduce:
                self.capture.begin_step()
        total = torch.zeros((), device=self.device)
        # e2e leg: each view's target planes come from pinned host memory inside the step; the copy of
        # view i+1 runs on a side stream while view i is rendered (what a data loader would do)
        pending = None
        if not resident and my_views:
            pending = self._stage_target(my_views[0])
        for i, v in enumerate(my_views):
            if resident:
                tgt = self.dev_targets[v % len(self.dev_targets)]
            else:
                tgt, ev = pending
                torch.cuda.current_stream(self.device).wait_event(ev)
                pending = self._stage_target(my_views[i + 1]) if i + 1 < len(my_views) else None
            means2D = torch.zeros_like(p["means3D"], requires_grad=True)
            ext = self.buf.rasterizer_kwargs(mod, first=(i == 0), slot=i) if fused else {}
            rast = mod.GaussianRasterizer(self.settings(mod, self.cams[v % len(self.cams)]), **ext)
            color, depth, quat, scale, alpha, radii = rast(
                means3D=p["means3D"], means2D=means2D, opacities=p["opacities"], shs=p["shs"],
                scales=p["scales"], rotations=p["rotations"])
            loss = (color - tgt[0:3]).abs().mean() + 0.05 * ((quat * tgt[3:7]).sum(0).abs().mean()
                                                             + (scale * tgt[7:10]).sum(0).abs().mean())
            loss.backward()
            stats.add_view(means2D.grad, radii)
            total = total + loss.detach()
        if self.world > 1 and reduce:
            if fused:
                self.buf.allreduce()
            else:
                # one view per rank (vegs_b200 arm only): the SH gradient crosses NVLink as its rank-one
                # factor dL/dRGB and is rebuilt by the library's kernel (vegs_b200/dist.py)
                use = self.factor_sh and views is None
                single = self.campos_all if (use and self.capture is None) else None
                self.flat = self.vdist.allreduce_step(p, stats, flat=self.flat, single_view_campos=single,
                                                      sh_factor=self.capture if use else None,
                                                      view_campos=self.campos_all if (use and self.capture is not None) else None)
        if not resident:
            return float(total.item())   # device -> host read of the step's result
        return total


def timed(fn, steps, warmup, device, world):
    import torch.distributed as dist
    for _ in range(warmup):
        fn()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize(device)
    if world > 1:
        dist.barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device=device)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item())


def kernel_groups(prof):
    """Per-view milliseconds of each stage group from the library's event profile."""
    groups = {"preprocess_fwd": ["preprocess_fwd"],
              "binning": ["radix_hist", "radix_rowscan", "radix_scatter", "tile_count", "tile_spine", "emit",
                          "tile_ranges", "tile_order", "memset_ranges", "memset_sort", "num_rendered_d2h"],
              "composite_fwd": ["composite_fwd"],
              "composite_bwd": ["composite_bwd", "memset_bwd_acc"],
              "preprocess_bwd": ["preprocess_bwd"]}
    return {g: sum(prof.get(n, (0.0, 0))[0] for n in names) for g, names in groups.items()}


def workload_stats(wl, mod):
    """P, V, R, R_f, Npix for view 0 .. n-1 (mean), from the library's exported state."""
    import vegs_b200.diff_gaussian_rasterization as dgr
    Vs, Rs, Rfs = [], [], []
    p = wl.params
    for cam in wl.cams[:2]:
        with torch.no_grad():
            res = dgr._forward_impl(wl.bg, p["means3D"], torch.Tensor([]), p["opacities"], p["scales"],
                                    p["rotations"], 1.0, torch.Tensor([]), cam.viewmatrix, cam.projmatrix,
                                    cam.tanfovx, cam.tanfovy, cam.H, cam.W, p["shs"], 3, cam.campos, False,
                                    False, keep_scratch=True)
        st = dgr.export_state(res[6], res[5])
        gx, gy = (cam.W + 15) // 16, (cam.H + 15) // 16
        nc = st["n_contrib"].to(torch.int64)
        pad = torch.zeros((gy * 16, gx * 16), dtype=torch.int64, device=nc.device)
        pad[:cam.H, :cam.W] = nc
        tmax = pad.view(gy, 16, gx, 16).amax(dim=(1, 3)).reshape(-1)
        rl = (st["ranges"][:, 1] - st["ranges"][:, 0]).to(torch.int64)
        rf = torch.minimum(rl, 256 * ((tmax + 255) // 256)).sum()
        Vs.append(int((res[5] > 0).sum())); Rs.append(int(st["num_rendered"])); Rfs.append(int(rf))
        del res, st
    return dict(P=wl.P, V=float(np.mean(Vs)), R=float(np.mean(Rs)), R_f=float(np.mean(Rfs)), Npix=wl.W * wl.H)


def algorithmic_bytes(s):
    """SURVEY.md §8(d) per-view byte model."""
    P, V, R, Rf, N = s["P"], s["V"], s["R"], s["R_f"], s["Npix"]
    return {"preprocess_fwd": 20 * P + 299 * V, "binning": 44 * R, "composite_fwd": 56 * Rf + 52 * N,
            "composite_bwd": 72 * Rf + 72 * N + 80 * V, "preprocess_bwd": 307 * V + 284 * P}


def cpu_baseline_oracle(args, n_views_sample=1):
    """The oracle port (CPU restatement of the reference algorithm) timed on the host cores on a
    bounded sample of the same workload: `n_views_sample` views of the same scene."""
    import util
    from oracle import oracle
    from vegs_b200 import synth
    sc = synth.make_scene(args.P, seed=7)
    t_total = 0.0
    for k in range(n_views_sample):
        cam = synth.make_camera(k, scale=args.scale)
        cot = util.np_cotangents(cam.W, cam.H)
        t0 = time.time()
        f = util.run_oracle(sc, cam, cot=cot)
        t_total += time.time() - t0
        del f
    return dict(value=n_views_sample / t_total, unit=UNIT, cores=oracle.num_threads(), kind="port",
                sample=f"{n_views_sample} view(s) fwd+bwd of the {args.P}-Gaussian scene at "
                       f"{int(1408 * args.scale)}x{int(376 * args.scale)}, oracle/vegs_oracle.c with OpenMP")


def cpu_baseline_dense_torch(n_chunks=8, chunk=256):
    """The second baseline BASELINE.json names: the dense PyTorch-CPU compositing evaluator
    (oracle/dense_torch.py, every Gaussian x every pixel, autograd backward) at config 1 (100 k
    Gaussians, 352x94), all host cores.  O(P * Npix): a bounded sample of n_chunks * chunk pixels
    spread over the image is timed (fwd+bwd per chunk) and scaled to whole views."""
    from oracle import dense_torch
    from vegs_b200 import synth
    torch.set_num_threads(os.cpu_count() or 1)
    sc = synth.make_scene(100_000, seed=7)
    cam = synth.make_camera(0, scale=0.25)
    npix = cam.W * cam.H
    t_total = 0.0
    for k in range(n_chunks):
        t = {n: v.clone().requires_grad_(True) for n, v in sc.items()}
        s0 = (npix // (n_chunks + 1)) * (k + 1)
        t0 = time.time()
        out = dense_torch.render_dense(
            t["means3D"], t["opacities"], shs=t["shs"], scales=t["scales"], rotations=t["rotations"],
            viewmatrix=cam.viewmatrix, projmatrix=cam.projmatrix, campos=cam.campos, bg=torch.zeros(3),
            W=cam.W, H=cam.H, tanfovx=cam.tanfovx, tanfovy=cam.tanfovy, sh_degree=3, quat_map=False,
            pixel_chunk=chunk, pixel_range=(s0, s0 + chunk))
        (out["color"].sum() + out["depth"].sum() + out["alpha"].sum() + out["cov_scale"].sum()).backward()
        t_total += time.time() - t0
    frac = n_chunks * chunk / npix
    return dict(value=frac / t_total, unit="views/s", cores=torch.get_num_threads(), kind="port",
                config="BASELINE config 1: 100000 Gaussians SH-3, 352x94, fwd+bwd",
                sample=f"{n_chunks} x {chunk} of the {npix} pixels (every Gaussian against every sampled pixel, "
                       f"float32, autograd backward, quaternion map left out), scaled by {1 / frac:.1f} to a whole view")


def emit(line: dict) -> None:
    """The ONE JSON line on the real stdout (everything else, e.g. NCCL's version banner, was
    redirected to stderr at start-up)."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


_REAL_STDOUT = os.dup(1)
os.dup2(2, 1)   # anything a library prints to stdout goes to stderr


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--P", type=int, default=6_000_000)
    ap.add_argument("--views", type=int, default=8)
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-fused-leg", action="store_true",
                    help="skip the extra value_fused / e2e_fused measurement of the vegs_b200 arm")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    mod, kind = load_rasterizer(args.impl)
    if args.impl == "reference" and mod is None:
        # no CUDA reference on this box: time the CPU port, rank 0 only
        if rank != 0:
            return
        cb = cpu_baseline_oracle(args, n_views_sample=max(1, min(args.steps, 3)))
        line = dict(metric=METRIC, value=cb["value"], unit=UNIT, n_gpus=args.gpus, steps=args.steps, warmup=args.warmup,
                    ms_per_step=1000.0 * args.views / cb["value"], higher_is_better=True, scaling="strong",
                    vs_baseline=None, dtype="f32", data="synthetic", impl="reference",
                    config=dict(workload=f"{args.P} Gaussians, {args.views}-view batch, CPU oracle port"),
                    cpu_baseline=cb, e2e=dict(value=cb["value"], unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0),
                    gpu_launches=0)
        emit(line); return

    assert torch.cuda.is_available(), "bench.py needs a GPU (there is no CPU path)"
    device = torch.device("cuda", local_rank)
    torch.cuda.set_device(device)
    if world > 1:
        import torch.distributed as dist
        import datetime
        dist.init_process_group("nccl", device_id=device, timeout=datetime.timedelta(seconds=180))
    ours = args.impl == "ours"
    _lib = None
    if ours:
        # only the vegs_b200 arm may map libvegs_b200.so: the reference process must stay free of it
        from vegs_b200 import _lib
    wl = Workload(args, device, rank, world)
    wl.factor_sh = ours and world > 1
    if wl.factor_sh and world != args.views:          # some rank renders several views (or none)
        wl.capture = wl.vdist.ShFactorCapture(wl.params["shs"], slots=(args.views + world - 1) // world)

    launches0 = _lib.launch_count() if ours else 0
    sampler = ClockSampler(local_rank) if rank == 0 else None
    ms = timed(lambda: wl.step(mod, True), args.steps, args.warmup, device, world)
    clocks = sampler.stop() if sampler else None
    launches = (_lib.launch_count() - launches0) * args.steps // (args.steps + args.warmup) if ours else 0
    ms_step = ms / args.steps
    value = args.views / (ms_step / 1000.0)

    ms_e2e = timed(lambda: wl.step(mod, False), args.steps, 3, device, world)
    e2e_value = args.views / (ms_e2e / args.steps / 1000.0)
    h2d = wl.h2d_bytes * len(wl.my_views) // args.views * world

    line = dict(metric=METRIC, value=value, unit=UNIT, n_gpus=world, steps=args.steps, warmup=args.warmup,
                ms_per_step=ms_step, higher_is_better=True, scaling="strong", vs_baseline=None, dtype="f32",
                data="synthetic",
                config=dict(workload=f"LiDAR-dense street scene, {args.P} Gaussians SH-3, {args.views}-view batch at "
                                     f"{wl.W}x{wl.H} split round-robin over {world} GPU(s), fwd+bwd through "
                                     "GaussianRasterizer + grad/densification-stat all-reduce (BASELINE config 3)",
                            P=args.P, views_per_step=args.views, image=[wl.W, wl.H], parallelism=f"view-parallel x{world}",
                            l2="inputs (1.4 GB of parameters per view) exceed the 126 MB L2; no explicit flush",
                            grad_accumulation="autograd add per view, densification statistics by the caller "
                                              "(the reference's own calling convention, identical in both arms)"),
                rasterizer=kind,
                reduction=("none (one GPU)" if world == 1 else
                           "vegs_b200.dist.allreduce_step: NCCL all-reduce of the .grad tensors and statistics" +
                           ("; the SH gradient is exchanged as every view's dL/dRGB (all-gather of 3 words per Gaussian "
                            "and view, " + ("kept by a hook on the SH leaf" if wl.capture is not None else
                                            "one view per rank: read off the DC row of .grad") +
                            ") and rebuilt by vegs_sh_grad_from_views" if wl.factor_sh else "")),
                e2e=dict(value=e2e_value, unit=UNIT, h2d_bytes_per_step=h2d, d2h_bytes_per_step=4 * world),
                gpu_launches=int(launches), clocks=clocks)

    if ours and not args.no_fused_leg:
        # the same step with this package's per-instance extensions (fused gradient accumulation into
        # views of the all-reduce buffer + fused densification statistics): reported beside, not
        # instead of, the drop-in figure, because the reference's callers cannot use them
        ms_f = timed(lambda: wl.step(mod, True, fused=True), args.steps, 3, device, world)
        line["value_fused"] = args.views / (ms_f / args.steps / 1000.0)
        ms_fe = timed(lambda: wl.step(mod, False, fused=True), args.steps, 2, device, world)
        line["e2e_fused"] = dict(value=args.views / (ms_fe / args.steps / 1000.0), unit=UNIT,
                                 h2d_bytes_per_step=h2d, d2h_bytes_per_step=4 * world)
        line["fused_what"] = ("GaussianRasterizer(settings, fused_grad_accumulation=..., densification_accumulators=..., "
                              "deferred_sh_grad=...): .grad tensors and the step statistics are views of one flat "
                              "all-reduce buffer; each view leaves dL/dRGB [P,3] instead of dL/dSH, the ranks all-gather "
                              "those and rebuild the SH gradient (vegs_sh_grad_from_views)")
    if world > 1:
        # weak scaling beside the fixed-batch figure: every rank renders 8 views per step (8 x N per step)
        wv = list(range(rank * args.views, (rank + 1) * args.views))
        ms_w = timed(lambda: wl.step(mod, True, views=wv), max(2, args.steps // 4), 1, device, world)
        line["weak_scaling"] = dict(views_per_step=args.views * world,
                                    value=args.views * world / (ms_w / max(2, args.steps // 4) / 1000.0), unit=UNIT)
    if args.impl == "reference":
        line["impl"] = "reference"
        line["cpu_baseline"] = dict(value=value, unit=UNIT, cores=0, kind="reference",
                                    sample="the reference's own CUDA extension (baseline/_ref, unmodified) on the GPU: "
                                           "the reference has no CPU implementation of this path")
    if rank == 0 and args.impl == "ours":
        # ---- roofline leg: per-kernel CUDA-event times from the library, byte model of SURVEY §8d
        peak, peak_src = peaks()
        _lib.profile_enable(True)
        wl.step(mod, True, reduce=False)      # rank 0 only: no collective in this extra step
        torch.cuda.synchronize(device)
        prof = _lib.profile_collect()
        _lib.profile_enable(False)
        nv = max(1, len(wl.my_views))
        groups = {k: v / nv for k, v in kernel_groups(prof).items()}
        stats = workload_stats(wl, mod)
        bts = algorithmic_bytes(stats)
        dom = max(groups, key=groups.get)
        per = {}
        for k in groups:
            gbs = bts[k] / (groups[k] * 1e-3) / 1e9 if groups[k] > 0 else 0.0
            per[k] = dict(ms_per_view=groups[k], algorithmic_bytes=bts[k], gbs=gbs, frac=gbs / peak)
        # DRAM bytes per launch of the dominant kernel from the committed ncu --set full capture
        # (profiles/ncu_traffic.json, written by scripts/ncu_summary.py); None when there is none
        traffic, traffic_src = None, None
        try:
            with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "ncu_traffic.json")) as f:
                tj = json.load(f)
            if dom in tj.get("bytes_per_launch", {}):
                traffic, traffic_src = tj["bytes_per_launch"][dom], tj.get("source")
        except (OSError, ValueError):
            pass
        line["roofline"] = dict(bound="hbm", kernel=dom, achieved=per[dom]["gbs"], peak=peak, unit="GB/s",
                                frac=per[dom]["frac"], traffic=traffic, traffic_source=traffic_src,
                                peak_source=peak_src,
                                how="CUDA events recorded by the library after every launch on the launching stream, "
                                    "one extra step after the timed region",
                                note=("the dominant kernel is not HBM-bound: its algorithmic bytes are gathered "
                                      "per tile-list entry; ncu shows 54 % issue-slot utilisation, no pipe above "
                                      "30 %, barrier waits as the top stall (profiles/r02_ncu_full_composite.md, "
                                      "DESIGN.md §8); the HBM-bound kernels are in `kernels`: preprocess_fwd and "
                                      "preprocess_bwd") if dom.startswith("composite") else None)
        line["kernels"] = per
        line["kernel_ms_per_view_raw"] = {k: v[0] / nv for k, v in prof.items()}
        line["workload_stats"] = stats
        view_bytes = sum(bts.values())
        line["view_roofline"] = dict(algorithmic_bytes=view_bytes, ms_per_view=ms_step / nv,
                                     frac=view_bytes / (ms_step / nv * 1e-3) / 1e9 / peak)
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline_oracle(args, 1)
            line["cpu_baseline_dense_torch"] = cpu_baseline_dense_torch()
    if rank == 0:
        emit(line)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
