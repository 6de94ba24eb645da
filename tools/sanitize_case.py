"""Small pass over every hand-written kernel family for compute-sanitizer (memcheck / racecheck / synccheck).

    compute-sanitizer --tool racecheck python tools/sanitize_case.py

Shapes are chosen so that every kernel runs several work items per CTA, a ragged tail and (for the tcgen05 kernels) more
than one accumulator / pipeline round, while the whole script stays within minutes under the sanitizer's ~100x slowdown.
Results are also checked against the oracle, so a run that "passes" the sanitizer with wrong numbers is still caught.
One tool per gpurun call (B200_PROFILING.md)."""
from __future__ import annotations

import ctypes as C
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

from oracle import encoder as eo  # noqa: E402
from oracle import frontend as fo  # noqa: E402
from yoho_b200 import preprocessing as pp  # noqa: E402
from yoho_b200._lib import check  # noqa: E402
from yoho_b200.nn import AudioEncoder, init_encoder_params  # noqa: E402
from yoho_b200.runtime import get_runtime  # noqa: E402
from yoho_b200.synth import synth_batch  # noqa: E402

rt = get_runtime()
lib = rt.lib


def bf16(a):
    return torch.from_numpy(np.ascontiguousarray(a, np.float32)).to(torch.bfloat16).to(torch.float32).numpy()


# ---- frontend: 3 clips x 1.3 s (4 tiles per clip, ragged last tile), all normalisers, float32 and int16 ----
x = synth_batch(3, 20333)
for norm, mode in (("reference", 1), ("whisper", 2)):
    got = pp.log_mel_spectogram(x, norm=norm)
    err = float(np.abs(got - fo.log_mel(x, norm_mode=mode)).max())
    print(f"[sanitize] logmel f32 {norm}: max err {err:.2e}")
    assert err <= 1e-3
got = pp.log_mel_spectogram(np.round(x * 32767).astype(np.int16), n_mels=128, norm="none")
print(f"[sanitize] logmel s16 / 128 mels: {got.shape}")

# ---- tcgen05 GEMM: both tile widths, GELU epilogue, ragged M / N ----
for m, n, k, epi in ((700, 768, 192, 1), (300, 160, 136, 0)):
    rng = np.random.default_rng(m)
    a = rng.standard_normal((m, k)).astype(np.float32)
    w = (rng.standard_normal((k, n)) / np.sqrt(k)).astype(np.float32)
    bias = rng.standard_normal(n).astype(np.float32)
    want = bf16(a).astype(np.float64) @ bf16(w).astype(np.float64) + bias
    if epi:
        want = 0.5 * want * (1 + np.tanh(np.sqrt(2 / np.pi) * (want + 0.044715 * want ** 3)))
    ad, wd, bd = (torch.from_numpy(v).cuda() for v in (a, w, bias))
    out = torch.full((m, n), float("nan"), device="cuda")
    nbytes = lib.yoho_test_gemm_workspace_bytes(m, n, k)
    ws = rt.workspace("test", nbytes)
    check(lib.yoho_test_gemm_bf16(rt.ctx, C.c_void_p(ad.data_ptr()), C.c_void_p(wd.data_ptr()), C.c_void_p(bd.data_ptr()),
                                  C.c_void_p(out.data_ptr()), m, n, k, epi, C.c_void_p(rt.aligned_ptr(ws)), nbytes, rt.stream_ptr()))
    torch.cuda.synchronize()
    err = float(np.abs(out.cpu().numpy() - want).max())
    print(f"[sanitize] gemm {m}x{n}x{k} epi={epi}: max err {err:.2e}")
    assert err <= 2e-3 * max(1.0, float(np.abs(want).max()))

# ---- tcgen05 attention: both kernels, ragged last key step, several items per CTA ----
for tiles in (2, 3):
    batch, t, heads = 3, 200, 2
    rng = np.random.default_rng(tiles)
    qkv = rng.standard_normal((batch, t, 3, heads, 64)).astype(np.float32)
    r = bf16(qkv).astype(np.float64)
    q, k, v = (r[:, :, i].transpose(0, 2, 1, 3) for i in range(3))
    s = q @ k.transpose(0, 1, 3, 2) / 8.0
    p = np.exp(s - s.max(-1, keepdims=True))
    p /= p.sum(-1, keepdims=True)
    want = (p @ v).transpose(0, 2, 1, 3).reshape(batch, t, heads * 64)
    qd = torch.from_numpy(qkv).cuda()
    out = torch.full((batch, t, heads * 64), float("nan"), device="cuda")
    nbytes = lib.yoho_test_attention_workspace_bytes(batch, t, heads)
    ws = rt.workspace("test", nbytes)
    prev = lib.yoho_test_attention_tiles(tiles)
    check(lib.yoho_test_attention_bf16(rt.ctx, C.c_void_p(qd.data_ptr()), C.c_void_p(out.data_ptr()), batch, t, heads,
                                       C.c_void_p(rt.aligned_ptr(ws)), nbytes, rt.stream_ptr()))
    torch.cuda.synchronize()
    lib.yoho_test_attention_tiles(prev)
    err = float(np.abs(out.cpu().numpy() - want).max() / np.abs(want).max())
    print(f"[sanitize] attention ({tiles} query tiles per item) {batch}x{t}x{heads}: max rel err {err:.2e}")
    assert err <= 1.5e-2

# ---- whole encoder, both compute modes (conv stem, LayerNorm, residual epilogues) ----
params = init_encoder_params(80, 128, 2, seed=3)
mel = np.random.default_rng(0).standard_normal((2, 199, 80)).astype(np.float32)
want = eo.audio_encoder(params, mel, 2, np.float64)
model = AudioEncoder(100, 128, 2, 2)
e32 = float(np.abs(model.apply({"params": params}, mel, compute="fp32") - want).max())
e16 = float(np.abs(model.apply({"params": params}, mel, compute="bf16") - want).max() / np.abs(want).max())
print(f"[sanitize] encoder: fp32 max err {e32:.2e}, bf16 max rel err {e16:.2e}")
assert e32 <= 1e-4 and e16 <= 2e-2
print("[sanitize] done")
