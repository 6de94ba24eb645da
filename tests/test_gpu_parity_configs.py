"""Direct-oracle parity of the bf16 / tcgen05 path on the BASELINE.json configurations.

The other GPU tests hold the tensor-core path against the float64 oracle only at tiny sizes and check the larger
configurations through the repo's own float32 kernels.  Here the *benchmarked* launch geometry is compared with
``oracle.encoder.audio_encoder`` (float64, reference whisper.py:117-135) itself:

* configs[1] base / 64 clips and configs[2] small / 256 clips: PCM -> log-mel -> encoder through ``ClipEncoder``; two
  probe clips sit *inside* the full batch (not at its edges) and are compared with the oracle chain on those clips;
* configs[3] large-v3 (128 mels, D = 1280, 20 heads, 32 layers): one full 2999-frame clip, with the drift reported at
  several depths (the same weights truncated to their first k blocks + final LayerNorm against the oracle's taps);
* the tcgen05 GEMM at the large-v3 shapes (K, N) in {(1280, 3840), (1280, 5120), (5120, 1280)} with M >= 96 000.

Tolerances (BASELINE.json north_star): log-mel 1e-4 absolute, encoder 2e-2 max-relative (max|ours - oracle| / max|oracle|).
Every measured error is appended to ``gpurun_out/r02_parity.txt`` (copied to ``profiles/`` by hand after a run).
"""
import ctypes as C
import os
import time
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL_MEL = 1e-4
TOL_BF16 = 2e-2
ROOT = Path(__file__).resolve().parents[1]


def _log(line: str) -> None:
    print(line)
    try:
        out = ROOT / "gpurun_out"
        out.mkdir(exist_ok=True)
        with open(out / "r02_parity.txt", "a") as f:
            f.write(line + "\n")
    except OSError:
        pass


@pytest.fixture(scope="module")
def rt(built_library):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from yoho_b200.runtime import get_runtime
    return get_runtime()


def _max_rel(got, want):
    return float(np.abs(got - want).max() / np.abs(want).max())


def _bf16(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a, np.float32)).to(torch.bfloat16).to(torch.float32).numpy()


# ------------------------------------------------------------------ configs[1] / configs[2]: probes inside the bench batch
@pytest.mark.parametrize("model,batch,probes", [("base", 64, (21, 62)), ("small", 256, (37, 201))])
def test_bench_batch_vs_oracle(rt, model, batch, probes):
    import torch
    from oracle import encoder as eo
    from oracle import frontend as fo
    from yoho_b200 import MODEL_SHAPES
    from yoho_b200.nn import init_encoder_params
    from yoho_b200.pipeline import ClipEncoder
    from yoho_b200.synth import synth_batch, synth_clip
    n_mels, dims, heads, layers = MODEL_SHAPES[model]
    params = init_encoder_params(n_mels, dims, layers, seed=42)
    enc = ClipEncoder(model, params)
    pcm = synth_batch(batch, 480000, unique=8)                     # the bench's own batch recipe (8 distinct clips x gains)
    for j, slot in enumerate(probes):                              # two clips found nowhere else in the batch
        pcm[slot] = synth_clip(1000 + j)
    dev = torch.from_numpy(pcm).cuda()
    mel = enc.logmel(dev).cpu().numpy()
    out = enc.encode(dev).cpu().numpy()
    out_bf = ClipEncoder(model, params, states_dtype="bf16").encode(dev)
    assert out_bf.dtype == torch.bfloat16
    out_bf = out_bf.float().cpu().numpy()
    rt._weights.clear()
    assert np.isfinite(out).all()
    t0 = time.time()
    probe_pcm = pcm[list(probes)]
    want_mel = fo.log_mel(probe_pcm, n_mels=n_mels, norm_mode=fo.NORM_REFERENCE)              # float64
    want = eo.audio_encoder(params, want_mel, heads, np.float64)
    mel_err = float(np.abs(mel[list(probes)] - want_mel).max())
    rel = _max_rel(out[list(probes)], want)
    rel_bf = _max_rel(out_bf[list(probes)], want)
    mean_rel = float(np.abs(out[list(probes)] - want).mean() / np.abs(want).mean())
    _log(f"[parity] {model} x {batch} clips (BASELINE configs[{1 if model == 'base' else 2}]), probes at {probes}: "
         f"log-mel max-abs {mel_err:.2e} (tol {TOL_MEL:g}); encoder bf16 path max-rel {rel:.3e}, mean-rel {mean_rel:.3e}, "
         f"bf16-states max-rel {rel_bf:.3e} (tol {TOL_BF16:g}); float64 oracle {time.time() - t0:.0f} s")
    assert mel_err <= TOL_MEL
    assert rel <= TOL_BF16
    assert rel_bf <= TOL_BF16
    # the states handed back as bf16 are the float32 states rounded once
    assert np.abs(out_bf - out).max() <= 2.0 ** -8 * np.abs(out).max() + 1e-6


# ------------------------------------------------------------------ configs[3]: large-v3, 32 layers, drift by depth
def test_large_v3_vs_oracle(rt):
    import torch
    from oracle import encoder as eo
    from oracle import frontend as fo
    from yoho_b200.nn import AudioEncoder, init_encoder_params
    from yoho_b200.synth import synth_clip
    n_mels, dims, heads, layers = 128, 1280, 20, 32
    t0 = time.time()
    params = init_encoder_params(n_mels, dims, layers, seed=42)
    mel = fo.log_mel(synth_clip(7)[None], n_mels=n_mels, norm_mode=fo.NORM_REFERENCE).astype(np.float32)
    assert mel.shape == (1, 2999, n_mels)
    taps: dict = {}
    want = eo.audio_encoder(params, mel, heads, np.float64, taps=taps)
    t_oracle = time.time() - t0
    worst = 0.0
    for depth in (1, 2, 4, 8, 16, 24, 32):
        model = AudioEncoder(1500, dims, heads, depth)                      # first `depth` blocks + the final LayerNorm
        got = model.apply({"params": params}, mel, compute="bf16")
        ref = want if depth == layers else eo.layer_norm(taps[f"block{depth - 1}"], params["LayerNorm_0"])
        rel = _max_rel(got, ref)
        mean_rel = float(np.abs(got - ref).mean() / np.abs(ref).mean())
        worst = max(worst, rel)
        _log(f"[parity] large-v3 (BASELINE configs[3]) depth {depth:2d}/32: bf16 path max-rel {rel:.3e}, mean-rel {mean_rel:.3e} "
             f"(tol {TOL_BF16:g}); residual max |x| {float(np.abs(taps[f'block{depth - 1}']).max()):.1f}")
        rt._weights.clear()                                                 # one 3.8 GB blob per depth: release it
        torch.cuda.empty_cache()
        assert np.isfinite(got).all()
        assert rel <= TOL_BF16, (depth, rel)
    _log(f"[parity] large-v3: worst max-rel over depths {worst:.3e}; float64 oracle + init {t_oracle:.0f} s on {os.cpu_count()} cores")


# ------------------------------------------------------------------ tcgen05 GEMM at the large-v3 shapes, M >= 96 000
@pytest.mark.parametrize("m,n,k,epi", [(96000, 3840, 1280, 0), (96001, 5120, 1280, 1), (96000, 1280, 5120, 0)])
def test_tcgen05_gemm_large_shapes(rt, m, n, k, epi):
    """64 clips x 1500 rows of large-v3's q|k|v, Dense_0 (+GELU) and Dense_1.  The float64 product is formed for a sample
    of rows only (the first and last 256-row tiles, the rows around every 65 536 boundary and 2048 random rows)."""
    import torch
    from yoho_b200._lib import check
    rng = np.random.default_rng(m + n + k)
    a = rng.standard_normal((m, k), dtype=np.float32)
    w = (rng.standard_normal((k, n), dtype=np.float32) / np.float32(np.sqrt(k)))
    bias = rng.standard_normal(n).astype(np.float32)
    ad, wd, bd = (torch.from_numpy(v).cuda() for v in (a, w, bias))
    out = torch.full((m, n), float("nan"), device="cuda")
    nbytes = rt.lib.yoho_test_gemm_workspace_bytes(m, n, k)
    ws = rt.workspace("test", nbytes)
    check(rt.lib.yoho_test_gemm_bf16(rt.ctx, C.c_void_p(ad.data_ptr()), C.c_void_p(wd.data_ptr()), C.c_void_p(bd.data_ptr()),
                                     C.c_void_p(out.data_ptr()), m, n, k, epi, C.c_void_p(rt.aligned_ptr(ws)), nbytes, rt.stream_ptr()))
    torch.cuda.synchronize()
    assert bool(torch.isfinite(out).all())                                   # every row was written
    rows = np.unique(np.concatenate([np.arange(256), np.arange(m - 256, m), np.arange(65536 - 128, 65536 + 128),
                                     rng.integers(0, m, 2048)]))
    got = out[torch.from_numpy(rows).cuda()].cpu().numpy()
    want = _bf16(a[rows]).astype(np.float64) @ _bf16(w).astype(np.float64) + bias
    if epi == 1:
        want = 0.5 * want * (1 + np.tanh(np.sqrt(2 / np.pi) * (want + 0.044715 * want ** 3)))
    err = float(np.abs(got - want).max())
    _log(f"[parity] tcgen05 GEMM m={m} n={n} k={k} epilogue={'bias+gelu' if epi else 'bias'}: max-abs {err:.2e} over {len(rows)} sampled rows "
         f"(|want| max {float(np.abs(want).max()):.2f})")
    assert err <= 2e-3 * max(1.0, float(np.abs(want).max()))
