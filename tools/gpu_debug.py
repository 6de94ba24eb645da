"""Stage-by-stage numerical diagnostics on the GPU box (prints error metrics, not just pass/fail).

    python tools/gpu_debug.py <stage> [...]     stages: frontend fp32 gemm attention bf16 perf

Each stage is meant to run in its own process under `timeout` (see tools/gpu_debug.sh) so that a hung
kernel cannot take the whole gpurun call with it.
"""
from __future__ import annotations

import ctypes as C
import sys
import time
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

from oracle import encoder as eo  # noqa: E402
from oracle import frontend as fe  # noqa: E402
from yoho_b200._lib import check  # noqa: E402
from yoho_b200.runtime import get_runtime  # noqa: E402
from yoho_b200.synth import synth_batch  # noqa: E402

GOLD = ROOT / "tests" / "golden"


def bf16(a):
    return torch.from_numpy(np.ascontiguousarray(a, np.float32)).to(torch.bfloat16).to(torch.float32).numpy()


def timed(fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def _hbm_peak_gbs() -> float:
    """MEASURED_PEAKS.json (driver-written) if present, else the profiling recipe's fallback."""
    import json
    try:
        return float(json.loads((Path(__file__).resolve().parents[1] / "MEASURED_PEAKS.json").read_text())["hbm_gbs"])
    except Exception:
        return 6500.0


HBM_PEAK_GBS = _hbm_peak_gbs()


def stage_frontend():
    from yoho_b200 import preprocessing as pp
    g = np.load(GOLD / "frontend_golden.npz")
    x = g["audio"]
    for m in (80, 128):
        p = pp.mel_spectogram(x, 400, 160, 16000, m)
        w = g[f"m{m}_power"]
        rel = np.abs(p[:2] - w[:2]) / np.maximum(np.abs(w[:2]), 1e-30)
        print(f"[frontend] M={m} power: max rel err {rel.max():.3e}  zero-clip max {np.abs(p[2]).max():.3e}")
        for norm in ("reference", "whisper"):
            o = pp.log_mel_spectogram(x, n_mels=m, norm=norm)
            e = np.abs(o - g[f"m{m}_{norm}"])
            print(f"[frontend] M={m} {norm}: max abs err clips0-2 {e[:3].max():.3e}  tonal clip {e[3].max():.3e}")
    xb = synth_batch(3, 480000)
    want = fe.log_mel(xb, norm_mode=1)
    got = pp.log_mel_spectogram(xb, norm="reference")
    print(f"[frontend] full 30 s x3 reference-norm max abs err {np.abs(got - want).max():.3e}")


def stage_fp32():
    from yoho_b200.nn import AudioEncoder, init_encoder_params
    g = np.load(GOLD / "encoder_golden.npz")
    params = init_encoder_params(80, 128, 2, seed=7)
    out = AudioEncoder(100, 128, 2, 2).apply({"params": params}, g["mel"], compute="fp32")
    print(f"[fp32] golden encoder max abs err {np.abs(out - g['out']).max():.3e} (max|y| {np.abs(g['out']).max():.2f})")


def run_gemm(rt, a, w, bias, epi):
    m, k = a.shape
    n = w.shape[1]
    ad, wd, bd = (torch.from_numpy(v).cuda() for v in (a, w, bias))
    out = torch.full((m, n), float("nan"), device="cuda")
    nbytes = rt.lib.yoho_test_gemm_workspace_bytes(m, n, k)
    ws = rt.workspace("test", nbytes)
    check(rt.lib.yoho_test_gemm_bf16(rt.ctx, C.c_void_p(ad.data_ptr()), C.c_void_p(wd.data_ptr()), C.c_void_p(bd.data_ptr()),
                                     C.c_void_p(out.data_ptr()), m, n, k, epi, C.c_void_p(rt.aligned_ptr(ws)), nbytes, rt.stream_ptr()))
    torch.cuda.synchronize()
    return out.cpu().numpy()


def stage_gemm():
    rt = get_runtime()
    for (m, n, k) in [(128, 128, 64), (128, 256, 64), (128, 128, 128), (256, 256, 256), (1000, 384, 192), (4099, 1152, 384), (129, 160, 72)]:
        rng = np.random.default_rng(m + n + k)
        a = rng.standard_normal((m, k)).astype(np.float32)
        w = (rng.standard_normal((k, n)) / np.sqrt(k)).astype(np.float32)
        bias = rng.standard_normal(n).astype(np.float32)
        want = bf16(a).astype(np.float64) @ bf16(w).astype(np.float64) + bias
        got = run_gemm(rt, a, w, bias, 0)
        err = np.abs(got - want)
        bad = ~np.isfinite(got)
        print(f"[gemm] {m}x{n}x{k}: max err {np.nanmax(err):.3e} nan {int(bad.sum())} (max|c| {np.abs(want).max():.2f})")
        if np.nanmax(err) > 1e-2 or bad.any():
            # structure of the error: per 32-row x 32-col block
            e = np.where(bad, 1e9, err)[: min(m, 128), : min(n, 256)]
            blk = e[: e.shape[0] // 32 * 32, : e.shape[1] // 32 * 32].reshape(e.shape[0] // 32, 32, e.shape[1] // 32, 32).max(axis=(1, 3))
            print("[gemm]   block max err (32x32 blocks of the first tile):\n", np.array2string(blk, precision=2, max_line_width=200))
            print("[gemm]   got[0,:8]", got[0, :8], "\n[gemm]   want[0,:8]", want[0, :8])
            # does it match a k-truncated product? (descriptor advance bugs)
            for kk in (16, 32, 48, 64):
                if kk <= k:
                    part = bf16(a[:, :kk]).astype(np.float64) @ bf16(w[:kk]).astype(np.float64) + bias
                    print(f"[gemm]   err vs first-{kk}-k partial: {np.nanmax(np.abs(got - part)):.3e}")


def stage_attention():
    rt = get_runtime()
    for (batch, t, heads, scale) in [(1, 128, 1, 1.0), (1, 256, 1, 1.0), (2, 100, 2, 1.0), (1, 64, 1, 1.0), (2, 192, 2, 1.0), (3, 700, 3, 1.0),
                                     (1, 1500, 2, 1.0), (2, 1500, 2, 3.0), (200, 300, 3, 1.0)]:
        rng = np.random.default_rng(t + heads)
        qkv = rng.standard_normal((batch, t, 3, heads, 64)).astype(np.float32)
        qkv[:, :, :2] *= scale          # scale > 1: score spread far beyond 2^8, exercises the lazy O rescale
        r = bf16(qkv).astype(np.float64)
        q, k, v = (r[:, :, i].transpose(0, 2, 1, 3) for i in range(3))
        s = q @ k.transpose(0, 1, 3, 2) / 8.0
        p = np.exp(s - s.max(-1, keepdims=True))
        p /= p.sum(-1, keepdims=True)
        want = (p @ v).transpose(0, 2, 1, 3).reshape(batch, t, heads * 64)
        qd = torch.from_numpy(qkv).cuda()
        out = torch.full((batch, t, heads * 64), float("nan"), device="cuda")
        nbytes = rt.lib.yoho_test_attention_workspace_bytes(batch, t, heads)
        ws = rt.workspace("test", nbytes)
        check(rt.lib.yoho_test_attention_bf16(rt.ctx, C.c_void_p(qd.data_ptr()), C.c_void_p(out.data_ptr()), batch, t, heads,
                                              C.c_void_p(rt.aligned_ptr(ws)), nbytes, rt.stream_ptr()))
        torch.cuda.synchronize()
        got = out.cpu().numpy()
        err = np.abs(got - want)
        print(f"[attention] B{batch} T{t} H{heads} x{scale}: max err {np.nanmax(err):.3e} nan {int((~np.isfinite(got)).sum())} (max|o| {np.abs(want).max():.2f})")
        if np.nanmax(err) > 3e-2:
            vmean = v.mean(axis=2)  # uniform softmax would give this
            print("[attention]   err vs uniform-softmax:", np.abs(got[0, :, :64] - vmean[0, 0]).max())
            print("[attention]   got[0,0,:6]", got[0, 0, :6], " want", want[0, 0, :6])
            rows = err[0].max(axis=1)
            print("[attention]   per-row err (first 8, last 8):", rows[:8], rows[-8:])


def stage_bf16():
    from yoho_b200.nn import AudioEncoder, init_encoder_params
    g = np.load(GOLD / "encoder_golden.npz")
    params = init_encoder_params(80, 128, 2, seed=7)
    model = AudioEncoder(100, 128, 2, 2)
    out = model.apply({"params": params}, g["mel"], compute="bf16")
    rel = np.abs(out - g["out"]).max() / np.abs(g["out"]).max()
    print(f"[bf16] golden encoder max-rel err {rel:.3e} nan {int((~np.isfinite(out)).sum())}")
    n_mels, dims, heads, layers = 80, 384, 6, 4
    from yoho_b200.synth import synth_clip
    mel = fe.log_mel(synth_clip(0)[None], norm_mode=1).astype(np.float32)
    params = init_encoder_params(n_mels, dims, layers, seed=42)
    want = eo.audio_encoder(params, mel, heads, np.float64)
    model = AudioEncoder(1500, dims, heads, layers)
    o32 = model.apply({"params": params}, mel, compute="fp32")
    o16 = model.apply({"params": params}, mel, compute="bf16")
    print(f"[bf16] tiny full clip: fp32 max abs err {np.abs(o32 - want).max():.3e}; bf16 max-rel {np.abs(o16 - want).max() / np.abs(want).max():.3e}")


def stage_perf():
    from yoho_b200 import preprocessing as pp
    from yoho_b200.nn import AudioEncoder, init_encoder_params
    B = 256
    pcm = torch.from_numpy(synth_batch(B, 480000, unique=8)).cuda()
    out = torch.empty((B, 2999, 80), device="cuda")
    for norm in ("none", "reference"):
        ms = timed(lambda: pp.log_mel_spectogram(pcm, norm=norm, out=out))
        gbs = B * 2879680 / ms / 1e6
        print(f"[perf] frontend B={B} norm={norm}: {ms:.3f} ms  {gbs:.0f} GB/s algorithmic ({gbs / HBM_PEAK_GBS * 100:.1f}% of measured HBM peak)")
    mel = pp.log_mel_spectogram(pcm, norm="reference")
    for name, (n_mels, dims, heads, layers), b in (("base", (80, 512, 8, 6), 64), ("small", (80, 768, 12, 12), 64)):
        params = init_encoder_params(n_mels, dims, layers, seed=1)
        model = AudioEncoder(1500, dims, heads, layers)
        x = mel[:b].contiguous()
        o = torch.empty((b, 1500, dims), device="cuda")
        ms = timed(lambda: model.apply({"params": params}, x, compute="bf16", out=o), iters=3, warm=1)
        gf = {"base": 87.37, "small": 344.16}[name]
        print(f"[perf] encoder {name} B={b} bf16: {ms:.2f} ms  {b * gf / ms:.1f} TFLOP/s  {b * 30 / ms * 1000:.0f}x realtime")


def stage_logmelperf():
    """Frontend alone: Whisper shape (80 mels, 30 s) in float32 and int16, YOHO shape (128 mels, 10.26 s)."""
    from yoho_b200 import preprocessing as pp
    for B, N, M, F in ((256, 480000, 80, 2999), (512, 164080, 128, 1024)):
        f32 = torch.from_numpy(synth_batch(B, N, unique=8)).cuda()
        s16 = (f32 * 32767.0).round().clamp(-32768, 32767).to(torch.int16)
        out = torch.empty((B, F, M), device="cuda")
        for name, pcm in (("f32", f32), ("s16", s16)):
            for norm in ("none", "reference"):
                ms = timed(lambda: pp.log_mel_spectogram(pcm, n_mels=M, norm=norm, out=out), iters=20, warm=3)
                gbs = B * (N * pcm.element_size() + F * M * 4) / ms / 1e6
                print(f"[logmelperf] B={B} N={N} M={M} {name} norm={norm}: {ms:.3f} ms  {gbs:.0f} GB/s algorithmic "
                      f"({gbs / HBM_PEAK_GBS * 100:.1f}% of measured HBM peak)  {B * N / 16000.0 / ms * 1000:.0f}x realtime")


def stage_logmelerr():
    """Where the log-mel error sits: worst element of the fused reference-normalised output, the mel power error at
    that element, and the error of the float32 oracle (numpy / scipy in float32) against the float64 one beside it."""
    from yoho_b200 import preprocessing as pp
    x = synth_batch(3, 480000)
    want = fe.log_mel(x, norm_mode=1)
    want32 = fe.log_mel(x, norm_mode=1, dtype=np.float32)
    got = pp.log_mel_spectogram(x, norm="reference")
    pw = fe.mel_spectogram(x.astype(np.float64), 400, 160, 16000, 80)
    pg = pp.mel_spectogram(x, 400, 160, 16000, 80)
    err = np.abs(got - want)
    b, f, m = np.unravel_index(err.argmax(), err.shape)
    print(f"[logmelerr] max abs err {err.max():.3e} at clip {b} frame {f} mel {m}: got {got[b, f, m]:.6f} want {want[b, f, m]:.6f}; "
          f"float32 oracle err {np.abs(want32 - want).max():.3e}")
    rel = np.abs(pg - pw) / np.maximum(pw, 1e-30)
    print(f"[logmelerr] mel power there: got {pg[b, f, m]:.6e} want {pw[b, f, m]:.6e} rel {rel[b, f, m]:.3e}; frame max power {pw[b, f].max():.3e}")
    big = pw > 1e-12
    rb, rf, rm = np.unravel_index(np.where(big, rel, 0).argmax(), rel.shape)
    print(f"[logmelerr] worst relative power error (power > 1e-12): {rel[rb, rf, rm]:.3e} at clip {rb} frame {rf} mel {rm} (power {pw[rb, rf, rm]:.3e}, frame max {pw[rb, rf].max():.3e})")
    for q in (50, 90, 99, 99.9, 99.99):
        print(f"[logmelerr]   percentile {q}: abs err {np.percentile(err, q):.3e}   rel power err {np.percentile(rel[big], q):.3e}")
    frames = np.unique(np.argwhere(err > 5e-5)[:, 1])
    print(f"[logmelerr] frames with an element above 5e-5: {frames[:40].tolist()} ({len(frames)} frames)")


def stage_decperf():
    """Text decoder (whisper-base shape: 51865 x 512, 8 heads, 6 layers, 448 positions, 1500 encoder positions): greedy
    generation with the KV cache, plain launches vs one replayed CUDA graph, and the teacher-forced forward."""
    from yoho_b200.nn import TextDecoder, init_decoder_params
    V, L, D, H, layers, T = 51865, 448, 512, 8, 6, 1500
    p = init_decoder_params(V, L, D, layers, seed=1)
    dec = TextDecoder(V, L, D, H, layers)
    w_bytes = layers * (4 * D * D + 2 * D * D + 8 * D * D) * 4 + V * D * 4            # weights read per token (self q,kv,o; cross q,o; mlp; logits)
    n_steps = 64
    for B in (1, 8, 64):
        states = torch.randn((B, T, D), device="cuda")
        buf = torch.zeros((B, L), dtype=torch.int32, device="cuda")
        buf[:, 0] = 50258
        kv_bytes = B * layers * (T * 2 * D * 4)                                        # cross k|v read per token
        for graph in (False, True):
            dec.generate({"params": p}, states, buf, n_steps=4, use_graph=graph)
            torch.cuda.synchronize()
            t0 = time.time()
            dec.generate({"params": p}, states, buf, n_steps=n_steps, use_graph=graph)
            torch.cuda.synchronize()
            dt = time.time() - t0
            print(f"[decperf] B={B:3d} graph={int(graph)}: {dt / n_steps * 1e6:8.1f} us/step (incl. cross k/v setup)  {B * n_steps / dt:9.0f} tokens/s  "
                  f"weights+cross-kv {(w_bytes + kv_bytes) / 1e6:.0f} MB/step -> {(w_bytes + kv_bytes) * n_steps / dt / 1e9:.0f} GB/s")
        lib = get_runtime().lib
        lib.yoho_prof_enable(1)
        dec.generate({"params": p}, states, buf, n_steps=16)
        torch.cuda.synchronize()
        ncls = lib.yoho_prof_num_classes()
        ms, work, cnt = (C.c_double * ncls)(), (C.c_double * ncls)(), (C.c_longlong * ncls)()
        check(lib.yoho_prof_collect(ms, work, cnt, ncls))
        lib.yoho_prof_enable(0)
        for i in range(ncls):
            if cnt[i]:
                print(f"[decperf]     {lib.yoho_prof_class_name(i).decode():26s} {cnt[i] / 16:6.1f} launches/step  {ms[i] / 16 * 1e3:8.1f} us/step  "
                      f"{ms[i] / cnt[i] * 1e3:6.1f} us each  {work[i] / ms[i] / 1e6:7.0f} GB/s")
    for B, S in ((8, 64), (8, 448)):
        states = torch.randn((B, T, D), device="cuda")
        toks = torch.randint(0, V, (B, S), device="cuda", dtype=torch.int32)
        ms = timed(lambda: dec.apply({"params": p}, toks, states), iters=3, warm=1)
        fl = B * (S * layers * (2 * 14 * D * D + 4 * S * D + 4 * T * D) + layers * T * 2 * 2 * D * D + S * 2 * D * V)
        print(f"[decperf] teacher-forced B={B} S={S}: {ms:.2f} ms  {fl / ms / 1e9:.1f} TFLOP/s fp32 SIMT")


def stage_yohoperf():
    """YOHO's own config (train/config.toml:7-18): 164080-sample clips -> (B, 1024, 128) log-mel -> own encoder (384/6/3/4)."""
    from yoho_b200 import preprocessing as pp
    from yoho_b200.nn.model import AudioEncoder, init_yoho_encoder_params
    B, N = 512, 164080
    pcm = torch.from_numpy(synth_batch(B, N, unique=8)).cuda()
    mel = torch.empty((B, 1024, 128), device="cuda")
    ms_f = timed(lambda: pp.log_mel_spectogram(pcm, n_mels=128, norm="reference", out=mel))
    params = init_yoho_encoder_params(128, 384, 6, 3, 4, seed=1)
    enc = AudioEncoder(6, 3, 384, 1024, 4)
    out = torch.empty((B, 512, 384), device="cuda")
    ms_e = timed(lambda: enc.apply({"params": params}, mel, compute="bf16", out=out), iters=5, warm=2)
    # per clip: conv0 2*1024*384*384, conv1 2*512*1152*384, per block qkv 2*512*384*576, attention 4*512^2*192,
    # out 2*512*192*384, w12 2*512*384*2304, w3 2*512*1152*384
    gf = (2 * 1024 * 384 * 384 + 2 * 512 * 1152 * 384 + 4 * (2 * 512 * 384 * 576 + 4 * 512 * 512 * 192 + 2 * 512 * 192 * 384
                                                           + 2 * 512 * 384 * 2304 + 2 * 512 * 1152 * 384)) / 1e9
    secs = N / 16000.0
    print(f"[yohoperf] B={B} clips of {secs:.2f} s: log-mel {ms_f:.3f} ms, encoder {ms_e:.2f} ms ({B * gf / ms_e:.1f} TFLOP/s, {gf:.2f} GFLOP/clip), "
          f"{B * secs / (ms_f + ms_e) * 1000:.0f}x realtime")


def stage_gemmperf():
    """Times the tcgen05 GEMM alone on the encoder's shapes (whisper-small, 64 clips = 96000 rows)."""
    rt = get_runtime()
    from yoho_b200 import _lib
    lib = rt.lib
    M = 96000
    for (name, n, k, epi) in [("qkv", 2304, 768, 0), ("out", 768, 768, 0), ("fc1", 3072, 768, 1), ("fc2", 768, 3072, 0), ("conv0", 768, 256, 1)]:
        a = torch.randn((M, k), device="cuda")
        w = torch.randn((k, n), device="cuda") / k ** 0.5
        bias = torch.randn(n, device="cuda")
        out = torch.empty((M, n), device="cuda")
        nbytes = lib.yoho_test_gemm_workspace_bytes(M, n, k)
        ws = rt.workspace("test", nbytes)
        lib.yoho_prof_enable(1)
        for _ in range(6):
            check(lib.yoho_test_gemm_bf16(rt.ctx, C.c_void_p(a.data_ptr()), C.c_void_p(w.data_ptr()), C.c_void_p(bias.data_ptr()),
                                          C.c_void_p(out.data_ptr()), M, n, k, epi, C.c_void_p(rt.aligned_ptr(ws)), nbytes, rt.stream_ptr()))
        torch.cuda.synchronize()
        ncls = lib.yoho_prof_num_classes()
        ms, work, cnt = (C.c_double * ncls)(), (C.c_double * ncls)(), (C.c_longlong * ncls)()
        check(lib.yoho_prof_collect(ms, work, cnt, ncls))
        lib.yoho_prof_enable(0)
        g = 2  # gemm_tcgen05 class
        print(f"[gemmperf] {name:6s} M={M} N={n} K={k}: {ms[g] / cnt[g] * 1e3:8.1f} us  {work[g] / ms[g] / 1e9:7.1f} TFLOP/s (f32 out; avg of {cnt[g]})")


def stage_attnperf():
    """Times the tcgen05 attention kernel alone (whisper-small heads, 64 clips)."""
    rt = get_runtime()
    lib = rt.lib
    for (batch, t, heads) in [(64, 1500, 12), (64, 1536, 12), (16, 1500, 20)]:
        qkv = torch.randn((batch, t, 3, heads, 64), device="cuda")
        out = torch.empty((batch, t, heads * 64), device="cuda")
        nbytes = lib.yoho_test_attention_workspace_bytes(batch, t, heads)
        ws = rt.workspace("test", nbytes)
        lib.yoho_prof_enable(1)
        for _ in range(5):
            check(lib.yoho_test_attention_bf16(rt.ctx, C.c_void_p(qkv.data_ptr()), C.c_void_p(out.data_ptr()), batch, t, heads,
                                               C.c_void_p(rt.aligned_ptr(ws)), nbytes, rt.stream_ptr()))
        torch.cuda.synchronize()
        ncls = lib.yoho_prof_num_classes()
        ms, work, cnt = (C.c_double * ncls)(), (C.c_double * ncls)(), (C.c_longlong * ncls)()
        check(lib.yoho_prof_collect(ms, work, cnt, ncls))
        lib.yoho_prof_enable(0)
        a = 3
        tiles = batch * heads * ((t + 127) // 128) ** 2
        us = ms[a] / cnt[a] * 1e3
        print(f"[attnperf] B{batch} T{t} H{heads}: {us:8.1f} us  {work[a] / ms[a] / 1e9:7.1f} TFLOP/s  "
              f"{us * 1e-6 * 1.9e9 * 148 / tiles:6.0f} SM-cycles@1.9GHz per 128x128 tile pair")


def stage_attnsus():
    """Attention kernel in the sustained (power-capped) regime: 256 clips of whisper-small, ~1.5 s of back-to-back
    launches before the timed ones, so the SM clock is where the bench step finds it."""
    import time
    rt = get_runtime()
    lib = rt.lib
    batch, t, heads = 256, 1500, 12
    qkv = torch.randn((batch, t, 3, heads, 64), device="cuda")
    out = torch.empty((batch, t, heads * 64), device="cuda")
    nbytes = lib.yoho_test_attention_workspace_bytes(batch, t, heads)
    ws = rt.workspace("test", nbytes)

    def run(n):
        for _ in range(n):
            check(lib.yoho_test_attention_bf16(rt.ctx, C.c_void_p(qkv.data_ptr()), C.c_void_p(out.data_ptr()), batch, t, heads,
                                               C.c_void_p(rt.aligned_ptr(ws)), nbytes, rt.stream_ptr()))
    t0 = time.time()
    while time.time() - t0 < 1.5:
        run(20)
        torch.cuda.synchronize()
    lib.yoho_prof_enable(1)
    run(100)
    torch.cuda.synchronize()
    ncls = lib.yoho_prof_num_classes()
    ms, work, cnt = (C.c_double * ncls)(), (C.c_double * ncls)(), (C.c_longlong * ncls)()
    check(lib.yoho_prof_collect(ms, work, cnt, ncls))
    lib.yoho_prof_enable(0)
    a = 3
    print(f"[attnsus] B{batch} T{t} H{heads} sustained: {ms[a] / cnt[a] * 1e3:8.1f} us  {work[a] / ms[a] / 1e9:7.1f} TFLOP/s  ({cnt[a]} launches)")


def stage_attnphases():
    """Debug build only (YOHO_EXTRA_NVCC_FLAGS=-DYOHO_ATT_TIMING): clock64 totals of the softmax phases of CTA 0 / warp 0."""
    rt = get_runtime()
    lib = rt.lib
    raw = C.CDLL(str(_lib_path()))
    lib.yoho_test_attention_tiles(2)          # the phase timers live in the two-tile kernel
    batch, t, heads = 64, 1500, 12
    qkv = torch.randn((batch, t, 3, heads, 64), device="cuda")
    out = torch.empty((batch, t, heads * 64), device="cuda")
    nbytes = lib.yoho_test_attention_workspace_bytes(batch, t, heads)
    ws = rt.workspace("test", nbytes)
    for _ in range(3):
        check(lib.yoho_test_attention_bf16(rt.ctx, C.c_void_p(qkv.data_ptr()), C.c_void_p(out.data_ptr()), batch, t, heads,
                                           C.c_void_p(rt.aligned_ptr(ws)), nbytes, rt.stream_ptr()))
    torch.cuda.synchronize()
    ph16 = (C.c_longlong * 16)()
    assert raw.yoho_debug_att_phases(ph16) == 0
    ph = list(ph16)[:8]
    names = ["loop top", "wait s_full", "tcgen05.ld x64 + wait", "row max + vote (+ rescale)", "scale, exp2, sum, pack", "tcgen05.st + wait + arrive",
             "wait last P V (per item)", "O read-out + store (per item)"]
    items = (batch * heads * ((t + 255) // 256) + 147) // 148
    steps = items * ((t + 63) // 64)
    total = sum(ph)
    for n, v in zip(names, ph):
        print(f"[attnphases] {n:32s} {v:10d} clk  {100.0 * v / total:5.1f} %   {v / steps:7.1f} clk/step")
    mma_names = ["loop top", "wait p_full", "wait o_free / v_full", "issue P V (4 MMA + commits)", "-", "issue Q K^T (g+2)", "-", "-"]
    for n, v in zip(mma_names, list(ph16)[8:]):
        print(f"[attnphases] MMA warp A: {n:32s} {v:10d} clk  {v / steps:7.1f} clk/step")
    print(f"[attnphases] total {total} clk over {items} items / {steps} steps = {total / steps:.0f} clk/step (SFU floor 512 per warp, 1024 per sub-partition)")


def _lib_path():
    from yoho_b200 import _lib
    return _lib.LIB_PATH


if __name__ == "__main__":
    torch.cuda.init()
    print("device:", torch.cuda.get_device_name(0))
    for st in sys.argv[1:]:
        t0 = time.time()
        globals()[f"stage_{st}"]()
        print(f"[{st}] done in {time.time() - t0:.1f}s", flush=True)
