"""CPU restatement of the reference log-mel frontend (TEST INFRASTRUCTURE ONLY).

Parity: unpinned by the reference except ``mel_filter_banks`` (see
``oracle/__init__.py``).  Every function cites the reference lines it follows;
paths are relative to the reference checkout (JirkaKlimes/YOHO).

The arithmetic that the reference delegates to third-party code is restated
from the published behaviour of those libraries at the versions pinned in the
reference's ``poetry.lock`` (jax 0.4.28 ``jax.scipy.signal.stft``, which mirrors
scipy 1.13 ``scipy.signal.stft``):

* ``window='hann'``  -> periodic Hann, built in float64, cast to the data dtype
* ``boundary=None``  -> no centring / reflect padding
* ``padded=True``    -> zero-extend the tail so that frames tile the signal
* ``detrend=False``, ``return_onesided=True``
* ``scaling='spectrum'`` -> multiply the rFFT by ``1/sum(window)``
* output laid out ``(..., freq, time)``
"""
from __future__ import annotations

import numpy as np

# --------------------------------------------------------------------------
# F0  mel filterbank   (yoho/src/preprocessing/mel_filterbanks.py:5-42)
# --------------------------------------------------------------------------

def hz_to_mel(f):
    """HTK mel scale, mel_filterbanks.py:5-6."""
    return 2595.0 * np.log10(1.0 + f / 700.0)


def mel_to_hz(m):
    """Inverse HTK mel scale, mel_filterbanks.py:9-10."""
    return 700.0 * (10.0 ** (m / 2595.0) - 1.0)


def mel_filter_banks(sr: float, n_fft: int, n_mels: int) -> np.ndarray:
    """(n_mels, n_fft//2+1) float32 triangular, area-normalised filterbank.

    Follows mel_filterbanks.py:21-42: ``n_mels+2`` band edges equally spaced in
    mel between 0 and sr/2 (:13-18, :31); triangles ``max(0, min(lower, upper))``
    from the edge-to-bin distances (:33-38); each float64 row is rounded into a
    float32 array (:29, :38) and only then scaled in place by the float64 area
    normaliser ``2/(f[i+2]-f[i])`` (:40-41) -- the double rounding is part of
    the result and is reproduced here.
    """
    n_bins = 1 + n_fft // 2
    bin_hz = np.fft.rfftfreq(n=n_fft, d=1.0 / sr)
    edges = mel_to_hz(np.linspace(hz_to_mel(0.0), hz_to_mel(sr * 0.5), n_mels + 2))
    width = np.diff(edges)
    dist = np.subtract.outer(edges, bin_hz)                    # (n_mels+2, n_bins)
    out = np.zeros((n_mels, n_bins), np.float32)
    for m in range(n_mels):
        rising = -dist[m] / width[m]
        falling = dist[m + 2] / width[m + 1]
        out[m] = np.maximum(0, np.minimum(rising, falling))    # f64 -> f32 store
    area = 2.0 / (edges[2 : n_mels + 2] - edges[:n_mels])
    out *= area[:, np.newaxis]                                 # in-place: f32 *= f64 -> f32
    return out


# --------------------------------------------------------------------------
# F1-F6  STFT   (yoho/src/preprocessing/audio.py:33)
# --------------------------------------------------------------------------

def stft_geometry(n_samples: int, n_fft: int, hop: int) -> tuple[int, int]:
    """Return (n_pad, n_frames) for ``stft(..., boundary=None, padded=True)``.

    audio.py:33 calls ``stft(audio, nperseg=n_fft, noverlap=n_fft-hop,
    boundary=None)``; with ``padded=True`` the library appends
    ``(-(N-nperseg) % nstep) % nperseg`` zeros so that an integer number of
    hops fits.  480000 -> (80, 2999); 480240 -> (0, 3000); 164080 -> (0, 1024).
    """
    if n_samples < n_fft:
        raise ValueError("oracle covers n_samples >= n_fft only")
    n_pad = (-(n_samples - n_fft) % hop) % n_fft
    n_frames = (n_samples + n_pad - n_fft) // hop + 1
    return n_pad, n_frames


def hann_periodic(n_fft: int, dtype=np.float32) -> np.ndarray:
    """``scipy.signal.get_window('hann', n_fft)`` (fftbins=True), cast to dtype."""
    n = np.arange(n_fft, dtype=np.float64)
    return (0.5 - 0.5 * np.cos(2.0 * np.pi * n / n_fft)).astype(dtype)


def stft(audio: np.ndarray, n_fft: int, hop: int, dtype=np.float64) -> np.ndarray:
    """One-sided STFT ``(B, n_fft//2+1, F)`` as audio.py:33 obtains it.

    ``dtype=np.float32`` performs framing, windowing, the rFFT and the scaling
    in single precision (the reference's precision, x64 disabled);
    ``np.float64`` is the exact-arithmetic yardstick.  The window is always the
    float32-rounded periodic Hann because that is the constant the reference
    multiplies by.
    """
    audio = np.asarray(audio)
    if audio.ndim != 2:
        raise ValueError("audio must be (B, N): the reference einsum is batched-only (audio.py:36)")
    B, N = audio.shape
    n_pad, F = stft_geometry(N, n_fft, hop)
    x = np.zeros((B, N + n_pad), dtype)
    x[:, :N] = audio
    win32 = hann_periodic(n_fft, np.float32)
    win = win32.astype(dtype)
    idx = hop * np.arange(F)[:, None] + np.arange(n_fft)[None, :]
    frames = x[:, idx] * win                                   # (B, F, n_fft)
    spec = np.fft.rfft(frames, n=n_fft, axis=-1)
    if dtype == np.float32:
        spec = spec.astype(np.complex64)
    scale = dtype(1.0) / win.sum(dtype=dtype)                  # sqrt(1/sum(w)^2)
    spec = spec * scale
    return np.swapaxes(spec, -1, -2)                           # (B, freq, time)


# --------------------------------------------------------------------------
# F7-F8  power + mel projection   (audio.py:34-36)
# --------------------------------------------------------------------------

def mel_spectogram(audio, n_fft: int, hop_len: int, sample_rate: int, n_mels: int,
                   dtype=np.float64, filters: np.ndarray | None = None) -> np.ndarray:
    """``(B, F, n_mels)`` mel power spectrogram; audio.py:32-37.

    magnitudes = |stft|**2 (:34); spectrogram = einsum('ij,ajk->aki', filters,
    magnitudes) (:36), i.e. time-major / mel-last.
    """
    s = stft(audio, n_fft, hop_len, dtype)
    power = (np.abs(s) ** 2).astype(dtype)
    if filters is None:
        filters = mel_filter_banks(sample_rate, n_fft, n_mels)
    return np.einsum("ij,ajk->aki", filters.astype(dtype), power, optimize=True).astype(dtype)


# --------------------------------------------------------------------------
# F9  reference normaliser   (audio.py:53-58)
# --------------------------------------------------------------------------

def normalize_spectogram(spectogram: np.ndarray, dtype=np.float64) -> np.ndarray:
    """log10(max(x,1e-13)), then per-clip (x-mean)/std over the last two axes.

    audio.py:54-57.  ``std`` is the population standard deviation (ddof=0) and
    a zero std is replaced by 1 (:57) so an all-constant clip maps to zeros.
    """
    x = np.asarray(spectogram, dtype)
    s = np.log10(np.maximum(x, dtype(1e-13)))
    mean = s.mean(axis=(-1, -2), keepdims=True, dtype=dtype)
    std = s.std(axis=(-1, -2), keepdims=True, dtype=dtype)
    return ((s - mean) / np.where(std != 0, std, dtype(1))).astype(dtype)


# --------------------------------------------------------------------------
# F9'  Whisper normaliser (named by BASELINE.json north_star; NOT in the reference)
# --------------------------------------------------------------------------

def whisper_normalize_spectogram(spectogram: np.ndarray, dtype=np.float64) -> np.ndarray:
    """log10(max(x,1e-10)); clamp below at (per-clip max - 8); (x+4)/4.

    No reference file: this is the formula BASELINE.json's north_star states
    ("log10 plus Whisper's clamp-to-(max-8) and (x+4)/4 normalization").  The max
    is taken per clip over (frames, mels).
    """
    x = np.asarray(spectogram, dtype)
    s = np.log10(np.maximum(x, dtype(1e-10)))
    peak = s.max(axis=(-1, -2), keepdims=True)
    s = np.maximum(s, peak - dtype(8.0))
    return ((s + dtype(4.0)) / dtype(4.0)).astype(dtype)


NORM_NONE, NORM_REFERENCE, NORM_WHISPER = 0, 1, 2


def log_mel(audio, n_fft=400, hop=160, sample_rate=16000, n_mels=80, norm_mode=NORM_REFERENCE,
            dtype=np.float64, filters=None) -> np.ndarray:
    """Frontend end to end: audio.py:32-37 followed by :53-58 (or the Whisper normaliser)."""
    spec = mel_spectogram(audio, n_fft, hop, sample_rate, n_mels, dtype, filters)
    if norm_mode == NORM_NONE:
        return spec
    if norm_mode == NORM_REFERENCE:
        return normalize_spectogram(spec, dtype)
    if norm_mode == NORM_WHISPER:
        return whisper_normalize_spectogram(spec, dtype)
    raise ValueError(f"unknown norm_mode {norm_mode}")
