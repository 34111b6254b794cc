#!/usr/bin/env python
"""Measures the all-reduce that bounds the view-parallel step: SUM over the 61 words per Gaussian
of the 6 M scene (1.46 GB fp32) and MAX over max_radii2D (24 MB), NCCL, CUDA events, max over ranks.
Launch: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/nccl_busbw.py
Prints one JSON line on rank 0: milliseconds, algorithm bandwidth (bytes / time) and bus bandwidth
(algbw * 2 (N-1) / N, what each link carries in a ring / tree all-reduce)."""
import json
import os

import torch
import torch.distributed as dist

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
P = 6_000_000
flat = torch.ones(61 * P, device=dev)
parts = [torch.ones(n * P, device=dev) for n in (3, 48, 1, 3, 4, 1, 1)]
maxr = torch.ones(P, device=dev)


def timed(fn, n=10):
    for _ in range(3):
        fn()
    dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / n], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def one_flat():
    dist.all_reduce(flat); dist.all_reduce(maxr, op=dist.ReduceOp.MAX)


def per_tensor():
    for p in parts:
        dist.all_reduce(p)
    dist.all_reduce(maxr, op=dist.ReduceOp.MAX)


# the exchange with the SH gradient sent as its factor: all-gather of every view's dL/dRGB (8 views
# over the ranks), all-reduce of the other 13 + 2 words per Gaussian, MAX of the radii
V = 8
drgb_local = torch.ones(V // world if V >= world else 1, P, 3, device=dev)
drgb_all = torch.empty(drgb_local.shape[0] * world, P, 3, device=dev)
rest = torch.ones(15 * P, device=dev)


def gather_only():
    dist.all_gather_into_tensor(drgb_all, drgb_local)


def rest_only():
    dist.all_reduce(rest); dist.all_reduce(maxr, op=dist.ReduceOp.MAX)


res = {}
for name, fn in (("flat_buffer", one_flat), ("per_tensor", per_tensor)):
    ms = timed(fn)
    nbytes = 4 * 62 * P
    res[name] = dict(ms=ms, algbw_GBs=nbytes / ms / 1e6, busbw_GBs=nbytes / ms / 1e6 * 2 * (world - 1) / world)
ms_g, ms_r = timed(gather_only), timed(rest_only)
res["factor_exchange"] = dict(all_gather_ms=ms_g, all_gather_bytes_out=4 * drgb_all.numel(),
                              all_gather_busbw_GBs=4 * drgb_all.numel() * (world - 1) / world / ms_g / 1e6,
                              all_reduce_15_words_ms=ms_r)
if rank == 0:
    print(json.dumps(dict(world=world, bytes=4 * 62 * P, **res)))
dist.barrier()
dist.destroy_process_group()
