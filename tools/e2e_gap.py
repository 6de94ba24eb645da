"""Where does the host-facing path lose time against the device-resident step?  Times, for one batch:

  device, whole batch           one kernel sequence over all clips (the `value` of bench.py)
  device, chunked like e2e      the same chunk plan as encode_host, no copies (chunking cost: launches, wave tails)
  e2e                           encode_host (copies inside)
  e2e without D2H / without H2D encode_host with one direction stubbed out (copy interference per direction)
  copies alone                  the H2D and D2H traffic of one step with no kernels running (PCIe time)

    python tools/e2e_gap.py [--model small] [--clips 256]"""
from __future__ import annotations

import argparse
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

from yoho_b200 import MODEL_SHAPES  # noqa: E402
from yoho_b200.nn import init_encoder_params  # noqa: E402
from yoho_b200.pipeline import ClipEncoder, chunk_plan  # noqa: E402
from yoho_b200.synth import synth_batch  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--model", default="small")
ap.add_argument("--clips", type=int, default=256)
ap.add_argument("--iters", type=int, default=5)
ap.add_argument("--chunk", type=int, default=64)
args = ap.parse_args()

n_mels, dims, heads, layers = MODEL_SHAPES[args.model]
params = init_encoder_params(n_mels, dims, layers, seed=1)
enc = ClipEncoder((n_mels, dims, heads, layers), params)
B = args.clips
pcm_host, out_host = enc.host_buffers(B)
pcm_host.copy_(torch.from_numpy(synth_batch(B, 480000, unique=4)))
pcm_dev = pcm_host.cuda()
out_dev = torch.empty((B, enc.seq_len, dims), device="cuda")


def timed(fn, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(args.iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / args.iters


plan = chunk_plan(B, args.chunk, True)
print(f"plan {[n for _, n in plan]}")
print(f"device, whole batch:        {timed(lambda: enc.encode(pcm_dev, out_dev)):8.2f} ms")


def device_chunked():
    for b0, n in plan:
        enc.encoder(enc.logmel(pcm_dev[b0:b0 + n], enc._mel_buf(n)), out_dev[b0:b0 + n])


print(f"device, chunked like e2e:   {timed(device_chunked):8.2f} ms")
for chunk in (32, 64, 128):
    even = [(b0, min(chunk, B - b0)) for b0 in range(0, B, chunk)]

    def device_even(even=even):
        for b0, n in even:
            enc.encoder(enc.logmel(pcm_dev[b0:b0 + n], enc._mel_buf(n)), out_dev[b0:b0 + n])
    print(f"device, even chunks of {chunk:3d}: {timed(device_even):8.2f} ms")
# alternate chunks on two compute streams (own workspaces): the wave tails / launch bubbles of one chunk's kernels are
# filled by the other chunk's kernels
streams2 = [torch.cuda.Stream(), torch.cuda.Stream()]


def device_two_streams(pl):
    cur = torch.cuda.current_stream()
    for st in streams2:
        st.wait_stream(cur)
    for i, (b0, n) in enumerate(pl):
        with torch.cuda.stream(streams2[i & 1]):
            mel = enc.logmel(pcm_dev[b0:b0 + n], enc._mel_buf(n, slot=10 + (i & 1)), ws_tag=f"logmel{i & 1}")
            enc.encoder(mel, out_dev[b0:b0 + n], ws_tag=f"encoder{i & 1}")
    for st in streams2:
        cur.wait_stream(st)


print(f"device, e2e plan, 2 streams:{timed(lambda: device_two_streams(plan)):8.2f} ms")
for chunk in (32, 64):
    even = [(b0, min(chunk, B - b0)) for b0 in range(0, B, chunk)]
    print(f"device, even {chunk:3d}, 2 streams:{timed(lambda: device_two_streams(even)):8.2f} ms")
print(f"e2e:                        {timed(lambda: enc.encode_host(pcm_host, out_host, chunk_clips=args.chunk)):8.2f} ms")

# one direction stubbed out: the "host" buffer of that direction is a device tensor (the copy becomes a cheap D2D)
print(f"e2e, states stay on device: {timed(lambda: enc.encode_host(pcm_host, out_dev, chunk_clips=args.chunk)):8.2f} ms")
print(f"e2e, PCM already on device: {timed(lambda: enc.encode_host(pcm_dev, out_host, chunk_clips=args.chunk)):8.2f} ms")


def copies_only():
    s1, s2 = enc._streams[0], enc._streams[2]
    with torch.cuda.stream(s1):
        pcm_dev.copy_(pcm_host, non_blocking=True)
    with torch.cuda.stream(s2):
        out_host.copy_(out_dev, non_blocking=True)
    torch.cuda.current_stream().wait_stream(s1)
    torch.cuda.current_stream().wait_stream(s2)


h2d, d2h = pcm_host.numel() * 4, out_host.numel() * out_host.element_size()
ms = timed(copies_only)
print(f"copies alone (both ways):   {ms:8.2f} ms  ({h2d / 1e9:.2f} GB in, {d2h / 1e9:.2f} GB out; {max(h2d, d2h) / ms / 1e6:.1f} GB/s on the busier direction)")
