"""Short, fixed workload for `ncu --set full` captures: the log-mel kernel over 128 clips and one
whisper-small encoder forward over 32 clips (same kernels and tile shapes as bench.py, fewer launches).

    python tools/prof_case.py [frontend] [encoder] [--model small] [--clips 32]
"""
from __future__ import annotations

import argparse
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

from yoho_b200 import MODEL_SHAPES  # noqa: E402
from yoho_b200.nn import init_encoder_params  # noqa: E402
from yoho_b200.pipeline import ClipEncoder  # noqa: E402
from yoho_b200.synth import synth_batch  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("what", nargs="*", default=["frontend", "encoder"])
ap.add_argument("--model", default="small")
ap.add_argument("--clips", type=int, default=32)
ap.add_argument("--iters", type=int, default=2)
args = ap.parse_args()

n_mels, dims, heads, layers = MODEL_SHAPES[args.model]
layers = min(layers, 2)                       # two blocks are enough: every kernel shape appears
params = init_encoder_params(n_mels, dims, layers, seed=1)
enc = ClipEncoder((n_mels, dims, heads, layers), params)
n_front = max(args.clips, 128)
pcm = torch.from_numpy(synth_batch(n_front, 480000, unique=4)).cuda()
mel = torch.empty((n_front, enc.frames, n_mels), device="cuda")
out = torch.empty((args.clips, enc.seq_len, dims), device="cuda")
for _ in range(args.iters):
    if "frontend" in args.what:
        enc.logmel(pcm, mel)
    if "encoder" in args.what:
        if "frontend" not in args.what:
            enc.logmel(pcm, mel)
        enc.encoder(mel[: args.clips], out)
torch.cuda.synchronize()
print("ok", float(out.abs().mean()) if "encoder" in args.what else float(mel.abs().mean()))
