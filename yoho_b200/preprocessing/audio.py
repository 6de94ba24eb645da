"""Drop-in for the hot-path functions of the reference ``yoho/src/preprocessing/audio.py``:

    mel_spectogram(audio, n_fft, hop_len, sample_rate, n_mels)     audio.py:32-37
    get_batched_spectogram(config)                                 audio.py:40-50
    normalize_spectogram(spectogram)                               audio.py:53-58

Same names (including the reference's spelling), argument order, shapes and dtypes:
audio ``(B, N)`` float32 -> ``(B, F, n_mels)`` float32, time-major / mel-last.  Every call runs the
hand-written sm_100a kernels of libyoho_b200 through the C ABI; there is no CPU path.
``load_audio`` / ``save_audio`` (ffmpeg subprocess I/O, audio.py:11-29) are out of scope.

Array types: a CUDA ``torch.Tensor`` is used in place and a CUDA tensor is returned (the analogue of
a traced device array inside ``jax.jit``); a NumPy array / CPU tensor is staged through pinned memory,
copied to the device, and the result is copied back to a NumPy array.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from .._lib import NORM_NONE, NORM_REFERENCE, NORM_WHISPER, check
from ..runtime import Runtime, get_runtime
from .mel_filterbanks import mel_filter_banks

_NORMS = {None: NORM_NONE, "none": NORM_NONE, "reference": NORM_REFERENCE, "whisper": NORM_WHISPER,
          NORM_NONE: NORM_NONE, NORM_REFERENCE: NORM_REFERENCE, NORM_WHISPER: NORM_WHISPER}


def _is_host(x) -> bool:
    return not (isinstance(x, torch.Tensor) and x.is_cuda)


def _runtime_for(x) -> Runtime:
    if isinstance(x, torch.Tensor) and x.is_cuda:
        return get_runtime(x.device.index)
    return get_runtime()


def _finish(out: torch.Tensor, host: bool):
    if not host:
        return out
    return out.cpu().numpy()       # synchronising D2H: the caller asked for a host array


def _logmel(audio, n_fft: int, hop_len: int, sample_rate: int, n_mels: int, norm: int, out: torch.Tensor | None = None):
    if getattr(audio, "ndim", None) != 2:
        # the reference's einsum 'ij,ajk->aki' is batched-only (audio.py:36)
        raise ValueError(f"audio must be (batch, samples); got ndim={getattr(audio, 'ndim', None)}")
    host = _is_host(audio)
    rt = _runtime_for(audio)
    int16 = (audio.dtype == torch.int16) if isinstance(audio, torch.Tensor) else (np.asarray(audio).dtype == np.int16)
    pcm = rt.to_device(audio, "pcm", torch.int16 if int16 else torch.float32)
    B, N = pcm.shape
    frames = check(rt.lib.yoho_stft_frames(N, n_fft, hop_len))
    fb = rt.filterbank(mel_filter_banks(sample_rate, n_fft, n_mels))
    if out is None:
        out = torch.empty((B, frames, n_mels), dtype=torch.float32, device=rt.device)
    if B == 0:
        return _finish(out, host)
    ws_bytes = rt.lib.yoho_logmel_workspace_bytes(B, N, n_fft, hop_len, n_mels)
    ws = rt.workspace("logmel", ws_bytes)
    fn = rt.lib.yoho_logmel_s16 if int16 else rt.lib.yoho_logmel_f32
    with torch.cuda.device(rt.device):
        check(fn(rt.ctx, fb, C.c_void_p(pcm.data_ptr()), B, N, n_fft, hop_len, C.c_void_p(out.data_ptr()), norm,
                 C.c_void_p(rt.aligned_ptr(ws)), ws_bytes, rt.stream_ptr()))
    return _finish(out, host)


def mel_spectogram(audio, n_fft: int, hop_len: int, sample_rate: int, n_mels: int):
    """Mel power spectrogram, reference audio.py:32-37: periodic-Hann STFT without centring
    (``boundary=None``, tail zero-padded), ``|.|**2``, HTK-mel projection.  ``(B, N) -> (B, F, n_mels)``."""
    return _logmel(audio, n_fft, hop_len, sample_rate, n_mels, NORM_NONE)


def get_batched_spectogram(config):
    """Closure over ``config.n_fft, stft_hop, sample_rate, n_mel_bands``; reference audio.py:40-50."""

    def func(audio):
        return mel_spectogram(audio, config.n_fft, config.stft_hop, config.sample_rate, config.n_mel_bands)

    return func


def _normalize(spectogram, norm: int):
    if getattr(spectogram, "ndim", 0) < 2:
        raise ValueError("spectogram needs at least (frames, mels)")
    host = _is_host(spectogram)
    rt = _runtime_for(spectogram)
    spec = rt.to_device(spectogram, "spec", torch.float32)
    shape = tuple(spec.shape)
    frames, mels = shape[-2], shape[-1]
    batch = int(np.prod(shape[:-2])) if len(shape) > 2 else 1     # statistics over the last two axes
    out = torch.empty_like(spec)
    ws_bytes = rt.lib.yoho_normalize_workspace_bytes(batch, frames, mels)
    ws = rt.workspace("normalize", ws_bytes)
    with torch.cuda.device(rt.device):
        check(rt.lib.yoho_normalize_f32(rt.ctx, C.c_void_p(spec.data_ptr()), batch, frames, mels, C.c_void_p(out.data_ptr()),
                                        norm, C.c_void_p(rt.aligned_ptr(ws)), ws_bytes, rt.stream_ptr()))
    return _finish(out, host)


def normalize_spectogram(spectogram):
    """``log10(max(x, 1e-13))`` then per-clip ``(x - mean) / std`` over (frames, mels), population std,
    zero std replaced by 1; reference audio.py:53-58."""
    return _normalize(spectogram, NORM_REFERENCE)


def whisper_normalize_spectogram(spectogram):
    """Whisper's normaliser named by BASELINE.json's north_star (not in the reference):
    ``log10(max(x, 1e-10))``, clamp below at per-clip ``max - 8``, ``(x + 4) / 4``."""
    return _normalize(spectogram, NORM_WHISPER)


def log_mel_spectogram(audio, n_fft: int = 400, hop_len: int = 160, sample_rate: int = 16000, n_mels: int = 80,
                       norm="reference", out: torch.Tensor | None = None):
    """Fused ``normalize_spectogram(mel_spectogram(...))`` (the composition at reference
    transcription_pretrain_no_voiceprints.py:172-179 and whisper.py:255-256) in one kernel pass;
    ``norm`` is ``"reference"``, ``"whisper"`` or ``"none"``.  Accepts float32 or int16 PCM."""
    return _logmel(audio, n_fft, hop_len, sample_rate, n_mels, _NORMS[norm], out)


def log_mel_spectogram_chunked(recording, clip_samples: int = 480000, n_fft: int = 400, hop_len: int = 160, sample_rate: int = 16000,
                               n_mels: int = 80, norm="reference", out: torch.Tensor | None = None):
    """Long-recording ingest: one mono recording ``(L,)`` (int16 as ``load_audio`` returns it, audio.py:11-18, or
    float32) is cut into ``ceil(L / clip_samples)`` clips, the last one zero padded (whisper.py:251-253), and each
    clip goes through the fused log-mel kernel -- the int16 -> float conversion (``/ 32768``, whisper.py:249), the
    padding and the framing all happen in the kernel's load, no ``(clips, clip_samples)`` array is materialised.
    Returns ``(clips, frames, n_mels)`` float32."""
    if getattr(recording, "ndim", None) != 1:
        raise ValueError(f"recording must be (samples,); got ndim={getattr(recording, 'ndim', None)}")
    host = _is_host(recording)
    rt = _runtime_for(recording)
    int16 = (recording.dtype == torch.int16) if isinstance(recording, torch.Tensor) else (np.asarray(recording).dtype == np.int16)
    pcm = rt.to_device(recording, "pcm", torch.int16 if int16 else torch.float32)
    total = int(pcm.shape[0])
    clips = check(rt.lib.yoho_chunk_count(total, clip_samples))
    frames = check(rt.lib.yoho_stft_frames(clip_samples, n_fft, hop_len))
    fb = rt.filterbank(mel_filter_banks(sample_rate, n_fft, n_mels))
    if out is None:
        out = torch.empty((clips, frames, n_mels), dtype=torch.float32, device=rt.device)
    if clips == 0:                                      # empty recording: nothing to launch (the C ABI agrees: batch 0 is a no-op)
        return _finish(out, host)
    ws_bytes = rt.lib.yoho_logmel_workspace_bytes(max(clips, 1), clip_samples, n_fft, hop_len, n_mels)
    ws = rt.workspace("logmel", ws_bytes)
    fn = rt.lib.yoho_logmel_chunked_s16 if int16 else rt.lib.yoho_logmel_chunked_f32
    with torch.cuda.device(rt.device):
        check(fn(rt.ctx, fb, C.c_void_p(pcm.data_ptr()), total, clip_samples, n_fft, hop_len, C.c_void_p(out.data_ptr()),
                 _NORMS[norm], C.c_void_p(rt.aligned_ptr(ws)), ws_bytes, rt.stream_ptr()))
    return _finish(out, host)
