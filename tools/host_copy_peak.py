"""Pinned-copy ceiling of the box: every rank streams the H2D + D2H traffic of one bench step (0.49 GB of PCM in, 1.18 GB
of encoder states out, 64-clip chunks, both directions at once on two streams) with NO kernels running, for ~2 s.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/host_copy_peak.py

Prints per-rank and aggregate GB/s for each direction alone and for both together: the host-side ceiling the
host-facing path (`e2e`) of bench.py runs against when N ranks share one host's PCIe roots and memory."""
from __future__ import annotations

import json
import os
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
from yoho_b200 import dist as yd  # noqa: E402

rank, local_rank, world = yd.init_process_group()
torch.cuda.set_device(local_rank)
dev = torch.device("cuda", local_rank)
clips, chunk = 256, 64
pcm_h = torch.empty((clips, 480000), dtype=torch.float32, pin_memory=True)
out_h = torch.empty((clips, 1500, 768), dtype=torch.float32, pin_memory=True)
pcm_d = torch.empty((chunk, 480000), dtype=torch.float32, device=dev)
out_d = torch.empty((chunk, 1500, 768), dtype=torch.float32, device=dev)
s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()


def step(h2d: bool, d2h: bool):
    for b0 in range(0, clips, chunk):
        if h2d:
            with torch.cuda.stream(s_in):
                pcm_d.copy_(pcm_h[b0:b0 + chunk], non_blocking=True)
        if d2h:
            with torch.cuda.stream(s_out):
                out_h[b0:b0 + chunk].copy_(out_d, non_blocking=True)


def timed(h2d, d2h, iters=12):
    step(h2d, d2h)
    torch.cuda.synchronize()
    yd.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        step(h2d, d2h)
    s_in.synchronize(); s_out.synchronize()
    b.record()
    torch.cuda.synchronize()
    yd.barrier()
    return yd.max_over_ranks(a.elapsed_time(b) / iters, dev)


res = {}
for name, (h, d) in {"h2d_alone": (True, False), "d2h_alone": (False, True), "both": (True, True)}.items():
    ms = timed(h, d)
    nbytes = (pcm_h.numel() * 4 if h else 0) + (out_h.numel() * 4 if d else 0)
    res[name] = {"ms_per_step": round(ms, 3), "gbs_per_rank": round(nbytes / ms / 1e6, 1), "gbs_all_ranks": round(world * nbytes / ms / 1e6, 1)}
if rank == 0:
    print(json.dumps({"n_gpus": world, "bytes_per_step": {"h2d": pcm_h.numel() * 4, "d2h": out_h.numel() * 4}, **res,
                      "cpu_count": os.cpu_count()}))
try:
    import torch.distributed as dist
    if dist.is_initialized():
        dist.destroy_process_group()
except Exception:
    pass
