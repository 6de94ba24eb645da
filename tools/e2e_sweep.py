"""Times ClipEncoder.encode_host (pinned host PCM -> pinned host encoder states) for several chunk plans next to the
device-resident step.   python tools/e2e_sweep.py [--model small] [--clips 256]"""
from __future__ import annotations

import argparse
import sys
import time
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

from yoho_b200 import MODEL_SHAPES, pipeline  # noqa: E402
from yoho_b200.nn import init_encoder_params  # noqa: E402
from yoho_b200.pipeline import ClipEncoder  # noqa: E402
from yoho_b200.synth import synth_batch  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="small")
ap.add_argument("--clips", type=int, default=256)
ap.add_argument("--iters", type=int, default=4)
args = ap.parse_args()

n_mels, dims, heads, layers = MODEL_SHAPES[args.model]
params = init_encoder_params(n_mels, dims, layers, seed=1)
enc = ClipEncoder((n_mels, dims, heads, layers), params)
B = args.clips
pcm_host, out_host = enc.host_buffers(B)
pcm_host.copy_(torch.from_numpy(synth_batch(B, 480000, unique=4)))
pcm_dev = pcm_host.cuda()
out_dev = torch.empty((B, enc.seq_len, dims), device="cuda")


def timed(fn):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(args.iters):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / args.iters * 1e3


print(f"device-resident: {timed(lambda: enc.encode(pcm_dev, out_dev)):.2f} ms")
orig_plan = pipeline.chunk_plan
for chunk in (32, 64, 128):
    for taper in (True, False):
        ms = timed(lambda: enc.encode_host(pcm_host, out_host, chunk_clips=chunk, taper=taper))
        print(f"chunk {chunk:3d} taper {taper!s:5s}: {ms:7.2f} ms  plan {[n for _, n in orig_plan(B, min(chunk, (B + 1) // 2), taper)]}")
for sizes in ([16, 48, 64, 64, 32, 16, 16], [16, 48, 64, 64, 32, 16, 8, 8], [16, 32, 64, 64, 48, 16, 8, 8], [8, 24, 64, 64, 64, 16, 8, 8], [16, 48, 96, 64, 16, 8, 8], [32, 64, 64, 64, 16, 8, 8], [16, 48, 64, 64, 40, 12, 6, 6], [16, 16, 64, 64, 64, 16, 8, 8], [16, 48, 128, 32, 16, 8, 8]):
    if sum(sizes) != B:
        continue
    def plan(batch, chunk, taper=True, sizes=sizes):
        out, b0 = [], 0
        for n in sizes:
            out.append((b0, n)); b0 += n
        return out
    pipeline.chunk_plan = plan
    ms = timed(lambda: enc.encode_host(pcm_host, out_host, chunk_clips=max(sizes)))
    print(f"plan {sizes}: {ms:7.2f} ms")
pipeline.chunk_plan = orig_plan
