"""Deterministic synthetic 16 kHz PCM used by the tests, ``smoke()`` and ``bench.py``.

There is no dataset on the build or GPU machines, so every measurement runs on
this generator: per clip, white Gaussian noise (sigma 0.1) plus three
linear chirps sweeping 100 Hz -> 7 kHz with amplitudes in [0.05, 0.2] and random
phase, plus a 0.5 s digital-silence gap at a random offset (exercises the
1e-13 / 1e-10 log floors and Whisper's max-8 clamp), clipped to [-1, 1], float32.
"""
from __future__ import annotations

import numpy as np

SAMPLE_RATE = 16000
CLIP_SECONDS = 30
CLIP_SAMPLES = SAMPLE_RATE * CLIP_SECONDS          # 480000 -> 2999 frames


def synth_clip(index: int, n_samples: int = CLIP_SAMPLES, sample_rate: int = SAMPLE_RATE) -> np.ndarray:
    rng = np.random.default_rng(1234 + index)
    t = np.arange(n_samples, dtype=np.float64) / sample_rate
    dur = n_samples / sample_rate
    x = 0.1 * rng.standard_normal(n_samples)
    for _ in range(3):
        amp = rng.uniform(0.05, 0.2)
        phase = rng.uniform(0.0, 2.0 * np.pi)
        f0, f1 = 100.0, 7000.0
        if rng.random() < 0.5:
            f0, f1 = f1, f0
        x += amp * np.sin(phase + 2.0 * np.pi * (f0 * t + 0.5 * (f1 - f0) / dur * t * t))
    gap = min(sample_rate // 2, n_samples // 4)
    start = int(rng.integers(0, n_samples - gap + 1))
    x[start : start + gap] = 0.0
    return np.clip(x, -1.0, 1.0).astype(np.float32)


def synth_batch(n_clips: int, n_samples: int = CLIP_SAMPLES, first_index: int = 0,
                unique: int | None = None, mix: bool = False) -> np.ndarray:
    """(n_clips, n_samples) float32.  ``unique`` bounds the number of clips that are synthesised from scratch so that
    large benchmark batches build quickly; the others are gain-scaled repeats of them, or, with ``mix``, DISTINCT clips:
    a synthesised clip plus a quarter of another one rotated by a clip-specific number of samples."""
    unique = n_clips if unique is None else max(1, min(unique, n_clips))
    base = np.stack([synth_clip(first_index + i, n_samples) for i in range(unique)])
    if unique == n_clips:
        return base
    out = np.empty((n_clips, n_samples), np.float32)
    for i in range(n_clips):
        gain = np.float32(1.0 - 0.03125 * ((i // unique) % 16))
        if mix and i >= unique:
            g = first_index + i
            out[i] = base[i % unique] * np.float32(0.75 * gain) + np.float32(0.25) * np.roll(base[(7 * i + 3) % unique], 997 * g + 131)
        else:
            out[i] = base[i % unique] * gain
    return out
