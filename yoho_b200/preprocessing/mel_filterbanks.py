"""Host-side mel filterbank constant: drop-in for the reference
``yoho/src/preprocessing/mel_filterbanks.py:21-42`` (same name, arguments, dtype, and -- checked with
``==`` against fixtures generated from the reference itself -- the same bits).

HTK mel scale; ``n_mels + 2`` band edges equally spaced in mel over [0, sr/2]; triangular filters with
area normalisation ``2 / (f[i+2] - f[i])``.  The reference rounds each float64 triangle into a float32
row and then scales the float32 array in place, so the two roundings are kept in that order.
The matrix is a device *constant* of the frontend (uploaded once as bands by yoho_filterbank_create).
"""
from functools import lru_cache

import numpy as np


def _hz_to_mel(hz):
    return 2595.0 * np.log10(1.0 + hz / 700.0)


def _mel_to_hz(mel):
    return 700.0 * (10.0 ** (mel / 2595.0) - 1.0)


@lru_cache(maxsize=None)
def mel_filter_banks(sr: float, n_fft: int, n_mels: int) -> np.ndarray:
    n_bins = n_fft // 2 + 1
    centers = np.fft.rfftfreq(n=n_fft, d=1.0 / sr)
    edges = _mel_to_hz(np.linspace(_hz_to_mel(0.0), _hz_to_mel(0.5 * sr), n_mels + 2))
    gaps = np.diff(edges)
    offsets = np.subtract.outer(edges, centers)
    bank = np.zeros((n_mels, n_bins), dtype=np.float32)
    for band in range(n_mels):
        up = -offsets[band] / gaps[band]
        down = offsets[band + 2] / gaps[band + 1]
        bank[band] = np.maximum(0, np.minimum(up, down))
    bank *= (2.0 / (edges[2:] - edges[:-2]))[:, None]
    bank.setflags(write=False)
    return bank
