"""Regenerate the committed golden fixtures (run from the repo root in the BUILD container):

    python tests/golden/make_golden.py

What pins what
* ``mel_filter_banks_{80,128}.npy`` -- produced by IMPORTING THE REFERENCE's own
  ``yoho/src/preprocessing/mel_filterbanks.py`` from /root/reference (the only hot-path module of the
  reference that imports here: jax/flax are absent).  These are reference outputs, bit for bit.
* ``frontend_golden.npz`` / ``encoder_golden.npz`` -- produced by ``oracle/`` (float64), after this script
  has re-checked the oracle against the independent witnesses available in the container
  (``scipy.signal.stft`` for the STFT, ``transformers`` WhisperEncoder sub-modules for the encoder).  The
  reference has no tests or vectors of its own for these stages, so they stay "parity unpinned by the
  reference"; the fixtures freeze the witnessed oracle so that the GPU box (no /root/reference, no
  network) checks against the very same numbers.
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
OUT = Path(__file__).resolve().parent

from oracle import encoder as enc  # noqa: E402
from oracle import frontend as fe  # noqa: E402
from yoho_b200.synth import synth_clip  # noqa: E402

GOLDEN_SAMPLES = 24000          # 1.5 s -> 149 frames (tail padded by 80 samples)
ENC_SHAPE = dict(n_mels=80, dims=128, n_heads=2, n_layers=2, frames=199)   # T = 100


def reference_filterbanks():
    ref_root = Path("/root/reference")
    if not ref_root.exists():
        raise SystemExit("needs /root/reference (build container) to regenerate the reference-pinned filterbanks")
    sys.path.insert(0, str(ref_root))
    from yoho.src.preprocessing.mel_filterbanks import mel_filter_banks as ref_fb
    for m in (80, 128):
        w = np.array(ref_fb(16000, 400, m))
        assert w.dtype == np.float32 and w.shape == (m, 201)
        assert np.array_equal(w, fe.mel_filter_banks(16000, 400, m)), "oracle filterbank != reference"
        np.save(OUT / f"mel_filter_banks_{m}.npy", w)
        print(f"mel_filter_banks_{m}: nnz={int((w != 0).sum())} written (reference output)")


def golden_audio() -> np.ndarray:
    """4 clips: two synthetic speech-like clips, all-zero, full-scale 1 kHz sine."""
    n = GOLDEN_SAMPLES
    t = np.arange(n) / 16000.0
    return np.stack([synth_clip(0, n), synth_clip(1, n), np.zeros(n, np.float32),
                     np.sin(2 * np.pi * 1000.0 * t).astype(np.float32)])


def frontend_fixture():
    import scipy.signal
    x = golden_audio()
    s = fe.stft(x, 400, 160, np.float64)
    _, _, z = scipy.signal.stft(x.astype(np.float64), nperseg=400, noverlap=240, boundary=None)
    err = np.abs(s - z).max()
    assert s.shape == z.shape and err < 1e-7, f"oracle STFT vs scipy witness: {err}"
    out = {"audio": x}
    for m in (80, 128):
        for mode, name in ((0, "power"), (1, "reference"), (2, "whisper")):
            out[f"m{m}_{name}"] = fe.log_mel(x, n_mels=m, norm_mode=mode, dtype=np.float64).astype(np.float32)
    np.savez_compressed(OUT / "frontend_golden.npz", **out)
    print(f"frontend_golden: stft witness err {err:.2e}; frames={s.shape[-1]}")


def encoder_fixture():
    c = ENC_SHAPE
    params = enc.init_params(c["n_mels"], c["dims"], c["n_layers"], seed=7)
    rng = np.random.default_rng(11)
    mel = rng.standard_normal((2, c["frames"], c["n_mels"])).astype(np.float32)
    taps = {}
    y = enc.audio_encoder(params, mel, c["n_heads"], np.float64, taps=taps)
    np.savez_compressed(OUT / "encoder_golden.npz", mel=mel, out=y.astype(np.float32),
                        conv0=taps["conv0"].astype(np.float32), block0=taps["block0"].astype(np.float32),
                        seed=np.int64(7), **{k: np.int64(v) for k, v in c.items()})
    print(f"encoder_golden: out {y.shape}, max|y|={np.abs(y).max():.3f}")


if __name__ == "__main__":
    reference_filterbanks()
    frontend_fixture()
    encoder_fixture()
