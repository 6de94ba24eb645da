"""CPU restatement of YOHO's own audio encoder (TEST INFRASTRUCTURE ONLY).

What ``Model.encode_audio`` runs in training (``yoho/src/nn/model.py:9-29, 82-83`` with the
blocks of ``yoho/src/nn/layers.py:7-116``; call site
``train/stages/transcription_pretrain_no_voiceprints.py:198-203``).  Parity: **unpinned** -- the
reference ships no tests and needs flax; this file follows the source line by line and encodes the
flax 0.8.4 defaults it relies on:

* ``nn.Conv(dims, 3)``: kernel ``(3, in, out)``, bias, stride 1, padding ``SAME`` (one zero row each
  side), cross-correlation;
* ``nn.RMSNorm()``: ``x * rsqrt(mean(x^2, -1) + 1e-6) * scale``;
* ``nn.max_pool(x, (1, 2), (1, 2))`` on a rank-3 ``(B, S, D)`` input: flax treats the two window
  dims as (B, S) (no batch dim left, it adds a singleton), padding ``VALID`` -> max over time pairs,
  ``(B, S // 2, D)``;
* ``nn.DenseGeneral((heads, head_dim), use_bias=False)``: kernel ``(in, heads, head_dim)``.

Flax auto-names of the ``@nn.compact`` bodies::

    params = {
      "Conv_0": {"kernel": (3, M, D), "bias": (D,)}, "RMSNorm_0": {"scale": (D,)},
      "Conv_1": {"kernel": (3, D, D), "bias": (D,)}, "RMSNorm_1": {"scale": (D,)},
      "EncoderBlock_i": {
          "RMSNorm_0": {"scale"},
          "GroupedQueryAttention_0": {
              "DenseGeneral_0": {"kernel": (D, q_heads, hd)},     # q
              "DenseGeneral_1": {"kernel": (D, kv_heads, hd)},    # k
              "DenseGeneral_2": {"kernel": (D, kv_heads, hd)},    # v
              "Dense_0": {"kernel": (kv_heads * hd, D)}},          # out   (RoPE_0 has no params)
          "RMSNorm_1": {"scale"},
          "SwiGLU_0": {"Dense_0": {"kernel": (D, 3D)}, "Dense_1": {"kernel": (D, 3D)},
                       "Dense_2": {"kernel": (3D, D)}}},
    }
"""
from __future__ import annotations

import math

import numpy as np

# [yoho] section of the reference's train/config.toml:7-18
YOHO_CONFIG = dict(n_mel_bands=128, max_audio_len=1024, dims=384, n_audio_heads=6, n_audio_blocks=4)


def rms_norm(x, p, eps=1e-6):
    """flax nn.RMSNorm defaults (model.py:18,23; layers.py:102,107)."""
    ms = np.mean(np.square(x), axis=-1, keepdims=True)
    return x / np.sqrt(ms + x.dtype.type(eps)) * np.asarray(p["scale"], x.dtype)


def silu(x):
    return x / (1.0 + np.exp(-x))


def conv1d_k3_same(x, p):
    """nn.Conv(dims, 3) (model.py:17,22): stride 1, SAME padding, cross-correlation."""
    k = np.asarray(p["kernel"], x.dtype)
    b = np.asarray(p["bias"], x.dtype)
    B, S, _ = x.shape
    xp = np.zeros((B, S + 2, x.shape[2]), x.dtype)
    xp[:, 1:-1] = x
    y = np.zeros((B, S, k.shape[2]), x.dtype)
    for tap in range(3):
        y += xp[:, tap:tap + S] @ k[tap]
    return y + b


def max_pool_time2(x):
    """nn.max_pool(x, (1, 2), (1, 2)) on (B, S, D) (model.py:20): VALID, max over time pairs."""
    B, S, D = x.shape
    s2 = S // 2
    return x[:, :2 * s2].reshape(B, s2, 2, D).max(axis=2)


def rope_tables(seq_len: int, head_dim: int):
    """cos / sin of layers.py:31-33: ``thetas = 10 ** -arange(dim//2)`` and positions 1..seq_len, the angle
    product taken in float32 (x64 is disabled in the reference), returned in float64."""
    thetas = np.power(np.float32(10.0), -np.arange(head_dim // 2, dtype=np.float32)).astype(np.float32)
    angles = np.outer(np.arange(1, seq_len + 1).astype(np.float32), thetas).astype(np.float32).astype(np.float64)
    return np.cos(angles), np.sin(angles)


def rope(x):
    """layers.py:25-43 on (..., s, d): pairs (x[2i], x[2i+1]) rotated by angle (s+1) * 10^-i."""
    s, d = x.shape[-2:]
    cos, sin = rope_tables(s, d)
    cos, sin = cos.astype(x.dtype), sin.astype(x.dtype)
    re, im = x[..., ::2], x[..., 1::2]
    out = np.empty_like(x)
    out[..., ::2] = re * cos - im * sin
    out[..., 1::2] = re * sin + im * cos
    return out


def grouped_query_attention(x, p, q_heads: int, kv_heads: int):
    """layers.py:46-92 with kv = q = x and no mask (the encoder's call, layers.py:103).

    Faithful to the einsum at :81: ``"... g h s d, ... h a d -> ... h s a"`` has no ``g`` on the right,
    so the scores are summed over the query group as well as over d."""
    B, S, D = x.shape
    wq = np.asarray(p["DenseGeneral_0"]["kernel"], x.dtype)   # (D, q_heads, hd)
    wk = np.asarray(p["DenseGeneral_1"]["kernel"], x.dtype)   # (D, kv_heads, hd)
    wv = np.asarray(p["DenseGeneral_2"]["kernel"], x.dtype)
    hd = wq.shape[2]
    groups = q_heads // kv_heads
    q = np.einsum("bsi,ihd->bshd", x, wq)                      # :65
    k = np.einsum("bsi,ihd->bshd", x, wk)
    v = np.einsum("bsi,ihd->bshd", x, wv)
    q = q.reshape(B, S, kv_heads, groups, hd).transpose(0, 3, 2, 1, 4)   # :69  s (h g) d -> g h s d
    k = k.transpose(0, 2, 1, 3)                                          # :70
    v = v.transpose(0, 2, 1, 3)
    q = rope(q)                                                          # :74
    k = rope(k)
    scores = np.einsum("bghsd,bhad->bhsa", q, k)                         # :77
    scores = scores / x.dtype.type(math.sqrt(hd))                        # :83-84
    scores = scores - scores.max(axis=-1, keepdims=True)
    w = np.exp(scores)
    w = w / w.sum(axis=-1, keepdims=True)
    out = np.einsum("bhsa,bhad->bhsd", w, v)                             # :86
    out = out.transpose(0, 2, 1, 3).reshape(B, S, kv_heads * hd)         # :87
    return out @ np.asarray(p["Dense_0"]["kernel"], x.dtype)             # :88


def swiglu(x, p):
    """layers.py:7-20."""
    w1 = np.asarray(p["Dense_0"]["kernel"], x.dtype)
    w2 = np.asarray(p["Dense_1"]["kernel"], x.dtype)
    w3 = np.asarray(p["Dense_2"]["kernel"], x.dtype)
    return (silu(x @ w1) * (x @ w2)) @ w3


def encoder_block(x, p, q_heads: int, kv_heads: int):
    """layers.py:95-116."""
    x = x + grouped_query_attention(rms_norm(x, p["RMSNorm_0"]), p["GroupedQueryAttention_0"], q_heads, kv_heads)
    return x + swiglu(rms_norm(x, p["RMSNorm_1"]), p["SwiGLU_0"])


def audio_encoder(params: dict, mel: np.ndarray, q_heads: int, kv_heads: int, dtype=np.float64, return_stages=False):
    """model.py:16-29: (B, S, M) -> (B, S // 2, D)."""
    x = np.asarray(mel, dtype)
    stages = {}
    x = silu(rms_norm(conv1d_k3_same(x, params["Conv_0"]), params["RMSNorm_0"]))
    x = max_pool_time2(x)
    stages["stem0"] = x
    x = silu(rms_norm(conv1d_k3_same(x, params["Conv_1"]), params["RMSNorm_1"]))
    stages["stem1"] = x
    i = 0
    while f"EncoderBlock_{i}" in params:
        x = encoder_block(x, params[f"EncoderBlock_{i}"], q_heads, kv_heads)
        stages[f"block{i}"] = x
        i += 1
    return (x, stages) if return_stages else x


def init_params(n_mels: int, dims: int, q_heads: int, kv_heads: int, n_blocks: int, seed: int = 42) -> dict:
    """Random float32 pytree (no checkpoint is available offline): kernels N(0, 1/fan_in), biases and RMSNorm
    scales non-trivial so that every fused term matters.  Same generator order as yoho_b200.nn.model.init_yoho_encoder_params."""
    rng = np.random.default_rng(seed)
    f32 = np.float32
    hd = dims // q_heads

    def kern(fan_in, *shape):
        return (rng.standard_normal(shape) / math.sqrt(fan_in)).astype(f32)

    def vec(n, mean=0.0):
        return (mean + 0.1 * rng.standard_normal(n)).astype(f32)

    p = {"Conv_0": {"kernel": kern(3 * n_mels, 3, n_mels, dims), "bias": vec(dims)}, "RMSNorm_0": {"scale": vec(dims, 1.0)},
         "Conv_1": {"kernel": kern(3 * dims, 3, dims, dims), "bias": vec(dims)}, "RMSNorm_1": {"scale": vec(dims, 1.0)}}
    for i in range(n_blocks):
        p[f"EncoderBlock_{i}"] = {
            "RMSNorm_0": {"scale": vec(dims, 1.0)},
            "GroupedQueryAttention_0": {
                "DenseGeneral_0": {"kernel": kern(dims, dims, q_heads, hd)},
                "DenseGeneral_1": {"kernel": kern(dims, dims, kv_heads, hd)},
                "DenseGeneral_2": {"kernel": kern(dims, dims, kv_heads, hd)},
                "Dense_0": {"kernel": kern(kv_heads * hd, kv_heads * hd, dims)},
            },
            "RMSNorm_1": {"scale": vec(dims, 1.0)},
            "SwiGLU_0": {"Dense_0": {"kernel": kern(dims, dims, 3 * dims)}, "Dense_1": {"kernel": kern(dims, dims, 3 * dims)},
                         "Dense_2": {"kernel": kern(3 * dims, 3 * dims, dims)}},
        }
    return p
