"""CPU tests of the oracle (test infrastructure) against the committed golden vectors, known-answer
inputs and the independent witnesses available offline.  No GPU."""
import numpy as np
import pytest

from oracle import encoder as enc
from oracle import frontend as fe
from yoho_b200.synth import synth_clip


# ------------------------------------------------------------------ F0 filterbank (reference-pinned)
@pytest.mark.parametrize("n_mels", [80, 128])
def test_filterbank_bit_exact_vs_reference_output(golden_dir, n_mels):
    ref = np.load(golden_dir / f"mel_filter_banks_{n_mels}.npy")      # written by importing the reference
    ours = fe.mel_filter_banks(16000, 400, n_mels)
    assert ours.dtype == np.float32 and ours.shape == (n_mels, 201)
    assert np.array_equal(ours, ref)


def test_filterbank_structure():
    w80, w128 = fe.mel_filter_banks(16000, 400, 80), fe.mel_filter_banks(16000, 400, 128)
    assert int((w80 != 0).sum()) == 393
    assert ((w80 != 0).sum(axis=0) <= 2).all() and ((w128 != 0).sum(axis=0) <= 2).all()
    assert [i for i in range(128) if not w128[i].any()] == [0, 3, 6, 13]   # four empty bands at 128 mels
    for w in (w80, w128):                                                # every band is one contiguous run
        for row in w:
            nz = np.nonzero(row)[0]
            assert nz.size == 0 or nz[-1] - nz[0] + 1 == nz.size


# ------------------------------------------------------------------ F1-F6 STFT
@pytest.mark.parametrize("n,expect", [(480000, (80, 2999)), (480240, (0, 3000)), (164080, (0, 1024)),
                                      (16400, (0, 101)), (400, (0, 1)), (401, (159, 2))])
def test_stft_geometry(n, expect):
    assert fe.stft_geometry(n, 400, 160) == expect


def test_stft_matches_scipy_witness():
    scipy_signal = pytest.importorskip("scipy.signal")
    x = np.stack([synth_clip(3, 32000), synth_clip(4, 32000)])
    ours = fe.stft(x, 400, 160, np.float64)
    _, _, z = scipy_signal.stft(x.astype(np.float64), nperseg=400, noverlap=240, boundary=None)
    assert ours.shape == z.shape == (2, 201, 199)
    assert np.abs(ours - z).max() < 1e-7          # float32-rounded vs float64 window


def test_stft_known_answers():
    n = 400 + 160 * 4
    w = fe.hann_periodic(400, np.float64)
    assert abs(w.sum() - 200.0) < 1e-9
    # DC: X[0] = 1, X[1] = -1/2 (Hann leakage), X[k>=2] = 0
    s = fe.stft(np.ones((1, n), np.float32), 400, 160)
    assert np.allclose(s[0, 0], 1.0, atol=1e-6) and np.allclose(s[0, 1], -0.5, atol=1e-6)
    assert np.abs(s[0, 2:]).max() < 1e-6
    # bin-centred sinusoid (k0 = 25 -> 1 kHz): |X[k0]| = 1/2, neighbours 1/4
    t = np.arange(n)
    s = fe.stft(np.cos(2 * np.pi * 25 * t / 400)[None].astype(np.float32), 400, 160)
    assert np.allclose(np.abs(s[0, 25]), 0.5, atol=1e-6) and np.allclose(np.abs(s[0, 24]), 0.25, atol=1e-6)
    # unit impulse at sample 500: frame f sees it at n = 500 - 160 f with weight w[n] / 200, flat magnitude
    x = np.zeros((1, n), np.float32)
    x[0, 500] = 1.0
    s = fe.stft(x, 400, 160)
    for f in range(5):
        pos = 500 - 160 * f
        want = w[pos] / 200.0 if 0 <= pos < 400 else 0.0
        assert np.allclose(np.abs(s[0, :, f]), want, atol=1e-6)


def test_tail_padding_is_zero_extension():
    x = synth_clip(5, 16000 + 37)[None]
    n_pad, frames = fe.stft_geometry(x.shape[1], 400, 160)
    padded = np.concatenate([x, np.zeros((1, n_pad), np.float32)], axis=1)
    a, b = fe.mel_spectogram(x, 400, 160, 16000, 80), fe.mel_spectogram(padded, 400, 160, 16000, 80)
    assert a.shape == (1, frames, 80) and np.array_equal(a, b)


# ------------------------------------------------------------------ F7-F9'
def test_frontend_golden_fixture(golden_dir):
    g = np.load(golden_dir / "frontend_golden.npz")
    x = g["audio"]
    for m in (80, 128):
        for mode, name in ((0, "power"), (1, "reference"), (2, "whisper")):
            got = fe.log_mel(x, n_mels=m, norm_mode=mode)
            want = g[f"m{m}_{name}"]
            assert got.shape == want.shape == (4, 149, m)
            assert np.allclose(got, want, rtol=1e-6, atol=1e-7), (m, name)


def test_normalisers_edge_cases():
    zeros = np.zeros((1, 7, 5))
    assert np.array_equal(fe.normalize_spectogram(zeros), np.zeros_like(zeros))       # std == 0 -> divide by 1
    assert np.allclose(fe.whisper_normalize_spectogram(zeros), (-10.0 + 4.0) / 4.0)
    spec = np.full((1, 4, 3), 1e-3)
    spec[0, 0, 0] = 1e3                                        # peak 3 -> floor at -5, the rest clamp up from -3? no: -3 > -5
    out = fe.whisper_normalize_spectogram(spec)
    assert np.isclose(out[0, 0, 0], (3 + 4) / 4) and np.isclose(out[0, 1, 1], (-3 + 4) / 4)
    spec[0, 1, 1] = 1e-9                                       # -9 < 3 - 8 -> clamped to -5
    assert np.isclose(fe.whisper_normalize_spectogram(spec)[0, 1, 1], (-5 + 4) / 4)
    rng = np.random.default_rng(0)
    s = fe.normalize_spectogram(rng.uniform(1e-6, 1.0, (3, 50, 8)))
    assert np.allclose(s.mean(axis=(1, 2)), 0, atol=1e-12) and np.allclose(s.std(axis=(1, 2)), 1, atol=1e-12)


def test_empty_mel_bands_floor_at_minus_13():
    x = synth_clip(6, 8000)[None]
    logmel = np.log10(np.maximum(fe.mel_spectogram(x, 400, 160, 16000, 128), 1e-13))
    assert np.array_equal(logmel[..., [0, 3, 6, 13]], np.full(logmel.shape[:2] + (4,), -13.0))


def test_float32_oracle_error_envelope():
    """The reference computes in float32; its own rounding error vs exact arithmetic is ~5e-6, far inside
    the 1e-4 tolerance the GPU kernels are held to."""
    x = np.stack([synth_clip(0, 48000), synth_clip(1, 48000)])
    for mode in (1, 2):
        err = np.abs(fe.log_mel(x, norm_mode=mode, dtype=np.float32) - fe.log_mel(x, norm_mode=mode)).max()
        assert err < 5e-5


# ------------------------------------------------------------------ encoder
def test_sinusoid_table():
    t = enc.sinusoid_table(1500, 384)
    assert t.shape == (1500, 384) and t.dtype == np.float32
    assert np.array_equal(t[0, :192], np.zeros(192, np.float32)) and np.array_equal(t[0, 192:], np.ones(192, np.float32))
    assert np.allclose(t[1, 0], np.sin(1.0), atol=1e-6) and np.allclose(t[1, 191], np.sin(1e-4), rtol=1e-4)


def test_conv_lengths():
    assert enc.conv_out_len(2999) == 1500 and enc.conv_out_len(3000) == 1500 and enc.conv_out_len(1024) == 512


def test_encoder_golden_fixture(golden_dir):
    g = np.load(golden_dir / "encoder_golden.npz")
    params = enc.init_params(int(g["n_mels"]), int(g["dims"]), int(g["n_layers"]), seed=int(g["seed"]))
    taps = {}
    y = enc.audio_encoder(params, g["mel"], int(g["n_heads"]), taps=taps)
    assert np.allclose(y, g["out"], atol=2e-6) and np.allclose(taps["conv0"], g["conv0"], atol=2e-6)
    y32 = enc.audio_encoder(params, g["mel"], int(g["n_heads"]), np.float32)
    assert np.abs(y32 - y).max() < 5e-5
    yt = enc.audio_encoder_torch_cpu(params, g["mel"], int(g["n_heads"]))      # the timed CPU baseline computes the same thing
    assert np.abs(yt - y).max() < 5e-5


def test_gelu_is_tanh_form_and_ln_eps():
    x = np.linspace(-4, 4, 101)
    from math import erf
    exact = np.array([0.5 * v * (1 + erf(v / np.sqrt(2))) for v in x])
    assert 1e-4 < np.abs(enc.gelu_tanh(x) - exact).max() < 1e-3          # the two GELUs differ by ~4.7e-4: parity-critical
    p = {"scale": np.ones(4), "bias": np.zeros(4)}
    tiny = np.array([[1e-4, -1e-4, 1e-4, -1e-4]])
    assert np.allclose(enc.layer_norm(tiny, p), tiny / np.sqrt(1e-8 + 1e-6))   # eps = 1e-6 (flax), not 1e-5 (torch)


def test_conv_is_cross_correlation_with_zero_pad():
    rng = np.random.default_rng(2)
    x = rng.standard_normal((1, 7, 3))
    p = {"kernel": rng.standard_normal((3, 3, 2)), "bias": rng.standard_normal(2)}
    y1, y2 = enc.conv1d_k3(x, p, 1), enc.conv1d_k3(x, p, 2)
    assert y1.shape == (1, 7, 2) and y2.shape == (1, 4, 2)
    want0 = x[0, 0] @ p["kernel"][1] + x[0, 1] @ p["kernel"][2] + p["bias"]          # t=0: left tap hits the pad
    assert np.allclose(y1[0, 0], want0) and np.allclose(y2[0, 0], want0)
    assert np.allclose(y2[0, 3], x[0, 5] @ p["kernel"][0] + x[0, 6] @ p["kernel"][1] + p["bias"])
    assert np.allclose(y2[0, 1], y1[0, 2])
