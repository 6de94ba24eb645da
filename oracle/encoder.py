"""CPU restatement of the reference Whisper audio encoder (TEST INFRASTRUCTURE ONLY).

Parity: unpinned by the reference (it has no tests and needs flax, which is not
installed); cross-checked against ``transformers`` ``WhisperEncoder`` sub-modules
in ``tests/test_oracle_witness.py``.  Citations are into
``yoho/src/nn/whisper.py`` of the reference checkout.

Library defaults the reference relies on (flax 0.8.4, pinned in poetry.lock):
``nn.gelu`` is the tanh approximation, ``nn.LayerNorm`` has epsilon 1e-6 and
computes the variance as E[x^2]-E[x]^2 clipped at 0, ``nn.Dense`` kernels are
``(in, out)``, ``nn.Conv`` kernels are ``(k, in, out)`` applied as a
cross-correlation, ``nn.softmax`` reduces over the last axis.

Parameters use the Flax auto-generated names of the reference modules, so a
pytree pickled from the reference (whisper.py:244-247) drops in unchanged::

    params = {
      "Conv_0": {"kernel": (3, M, D), "bias": (D,)},
      "Conv_1": {"kernel": (3, D, D), "bias": (D,)},
      "EncoderBlock_i": {
          "LayerNorm_0": {"scale", "bias"},
          "MultiHeadAttention_0": {
              "query_proj": {"kernel", "bias"}, "key_proj": {"kernel"},
              "value_proj": {"kernel", "bias"}, "out_proj": {"kernel", "bias"}},
          "LayerNorm_1": {"scale", "bias"},
          "Dense_0": {"kernel": (D, 4D), "bias"}, "Dense_1": {"kernel": (4D, D), "bias"}},
      "LayerNorm_0": {"scale", "bias"},
    }
"""
from __future__ import annotations

import math

import numpy as np

MODEL_SHAPES = {
    # name: (n_mels, dims, n_heads, n_layers)   -- whisper.py:215-224 gives the base row
    "tiny": (80, 384, 6, 4),
    "base": (80, 512, 8, 6),
    "small": (80, 768, 12, 12),
    "large-v3": (128, 1280, 20, 32),
}


def conv_out_len(n_frames: int) -> int:
    """Length after Conv(k=3, padding=1, strides=2): whisper.py:127."""
    return (n_frames + 2 - 3) // 2 + 1


# ---------------------------------------------------------------- E3
def sinusoid_table(length: int, dim: int, max_timescale: int = 10000) -> np.ndarray:
    """(length, dim) float32 table of whisper.py:11-17, evaluated in float32.

    The reference runs with x64 disabled, so every step is single precision:
    ``log(max_timescale)/(dim//2-1)`` (:14), ``e ** (-inc * arange(dim//2))``
    (:15), ``arange(length)[:,None] * inv[None,:]`` (:16), then
    ``concat([sin, cos], axis=1)`` (:17).  float32-vs-float64 evaluation differs
    by ~1e-4, so the table is treated as a constant that the host builds once
    and hands to every implementation.
    """
    assert dim % 2 == 0
    f32 = np.float32
    inc = f32(np.log(f32(max_timescale))) / f32(dim // 2 - 1)
    inv = np.power(f32(np.e), (-inc * np.arange(dim // 2, dtype=f32)).astype(f32)).astype(f32)
    st = (np.arange(length, dtype=f32)[:, None] * inv[None, :]).astype(f32)
    return np.concatenate([np.sin(st), np.cos(st)], axis=1).astype(f32)


# ---------------------------------------------------------------- primitives
def gelu_tanh(x):
    """flax ``nn.gelu`` default (approximate=True); whisper.py:84,126,128."""
    c = math.sqrt(2.0 / math.pi)
    return 0.5 * x * (1.0 + np.tanh(c * (x + 0.044715 * (x * x * x))))


def layer_norm(x, p, eps=1e-6):
    """flax ``nn.LayerNorm()``: whisper.py:77,82,134."""
    mean = x.mean(axis=-1, keepdims=True)
    mean2 = (x * x).mean(axis=-1, keepdims=True)
    var = np.maximum(0.0, mean2 - mean * mean)
    y = (x - mean) / np.sqrt(var + eps)
    return y * p["scale"].astype(x.dtype) + p["bias"].astype(x.dtype)


def dense(x, p):
    """flax ``nn.Dense``: y = x @ kernel (+ bias); whisper.py:32-35,83,85."""
    y = x @ p["kernel"].astype(x.dtype)
    if "bias" in p:
        y = y + p["bias"].astype(x.dtype)
    return y


def conv1d_k3(x, p, stride: int):
    """flax ``nn.Conv(D, 3, padding=1, strides=stride)`` over axis -2; whisper.py:125,127.

    y[b,t,o] = bias[o] + sum_k sum_c x[b, stride*t + k - 1, c] * W[k, c, o] with
    zero padding of one row on each side.
    """
    B, L, C = x.shape
    W = p["kernel"].astype(x.dtype)
    T = (L + 2 - 3) // stride + 1
    xp = np.zeros((B, L + 2, C), x.dtype)
    xp[:, 1 : L + 1] = x
    y = np.zeros((B, T, W.shape[2]), x.dtype)
    for k in range(3):
        y += xp[:, k : k + stride * (T - 1) + 1 : stride] @ W[k]
    return y + p["bias"].astype(x.dtype)


def softmax_last(x):
    x = x - x.max(axis=-1, keepdims=True)
    e = np.exp(x)
    return e / e.sum(axis=-1, keepdims=True)


# ---------------------------------------------------------------- E5/E6
def multi_head_attention(x, p, n_head: int):
    """Encoder self-attention ``MultiHeadAttention(x, x)``; whisper.py:37-67.

    q = x Wq + bq (:43); k = x Wk, no bias (:33,45); v = x Wv + bv (:46); q and k
    are each scaled by d_head**-0.25 (:54-57); softmax over keys (:64); heads
    merged (:65); out_proj (:50).
    """
    B, S, D = x.shape
    q = dense(x, p["query_proj"])
    k = dense(x, p["key_proj"])
    v = dense(x, p["value_proj"])
    dh = D // n_head
    scale = x.dtype.type(dh ** -0.25)
    q = q.reshape(B, S, n_head, dh).transpose(0, 2, 1, 3) * scale
    k = k.reshape(B, S, n_head, dh).transpose(0, 2, 3, 1) * scale
    v = v.reshape(B, S, n_head, dh).transpose(0, 2, 1, 3)
    w = softmax_last(q @ k)
    o = (w @ v).transpose(0, 2, 1, 3).reshape(B, S, D)
    return dense(o, p["out_proj"])


# ---------------------------------------------------------------- E4
def encoder_block(x, p, n_head: int):
    """Pre-LN block; whisper.py:75-88."""
    x = x + multi_head_attention(layer_norm(x, p["LayerNorm_0"]), p["MultiHeadAttention_0"], n_head)
    h = gelu_tanh(dense(layer_norm(x, p["LayerNorm_1"]), p["Dense_0"]))
    return x + dense(h, p["Dense_1"])


# ---------------------------------------------------------------- E1,E2,E3,E7
def audio_encoder(params: dict, mel: np.ndarray, n_heads: int, dtype=np.float64,
                  pos_table: np.ndarray | None = None, taps: dict | None = None) -> np.ndarray:
    """``AudioEncoder.__call__``; whisper.py:124-135.  mel (B, F, M) -> (B, T, D).

    ``taps`` (optional dict) receives intermediate activations for debugging a
    device implementation stage by stage.
    """
    x = np.asarray(mel, dtype)
    x = gelu_tanh(conv1d_k3(x, params["Conv_0"], 1))
    if taps is not None:
        taps["conv0"] = x
    x = gelu_tanh(conv1d_k3(x, params["Conv_1"], 2))
    if taps is not None:
        taps["conv1"] = x
    T, D = x.shape[1], x.shape[2]
    if pos_table is None:
        pos_table = sinusoid_table(T, D)
    if pos_table.shape != (T, D):
        raise ValueError(f"positional table {pos_table.shape} must equal (T, D)=({T}, {D}) (whisper.py:23 broadcasts)")
    x = x + pos_table.astype(dtype)
    n_layers = sum(1 for k in params if k.startswith("EncoderBlock_"))
    for i in range(n_layers):
        x = encoder_block(x, params[f"EncoderBlock_{i}"], n_heads)
        if taps is not None:
            taps[f"block{i}"] = x
    return layer_norm(x, params["LayerNorm_0"])


def encode_audio(variables: dict, mel, n_heads: int, dtype=np.float64, **kw):
    """``Whisper.encode_audio`` (whisper.py:194-195) on ``{"params": {"audio_encoder": ...}}``
    or on a bare AudioEncoder tree."""
    params = variables.get("params", variables)
    params = params.get("audio_encoder", params)
    return audio_encoder(params, mel, n_heads, dtype, **kw)


# ---------------------------------------------------------------- weights
def init_params(n_mels: int, dims: int, n_layers: int, seed: int = 42) -> dict:
    """Random-init float32 AudioEncoder pytree: kernels N(0, 1/fan_in),
    biases N(0, 0.1^2), LN scale 1+N(0, 0.1^2) so that every fused epilogue term
    is exercised (Flax's zeros/ones init would hide bias bugs)."""
    rng = np.random.default_rng(seed)
    f32 = np.float32

    def kern(*shape):
        fan_in = int(np.prod(shape[:-1]))
        return (rng.standard_normal(shape) / math.sqrt(fan_in)).astype(f32)

    def vec(n, mean=0.0):
        return (mean + 0.1 * rng.standard_normal(n)).astype(f32)

    def ln():
        return {"scale": vec(dims, 1.0), "bias": vec(dims)}

    p = {
        "Conv_0": {"kernel": kern(3, n_mels, dims), "bias": vec(dims)},
        "Conv_1": {"kernel": kern(3, dims, dims), "bias": vec(dims)},
    }
    for i in range(n_layers):
        p[f"EncoderBlock_{i}"] = {
            "LayerNorm_0": ln(),
            "MultiHeadAttention_0": {
                "query_proj": {"kernel": kern(dims, dims), "bias": vec(dims)},
                "key_proj": {"kernel": kern(dims, dims)},
                "value_proj": {"kernel": kern(dims, dims), "bias": vec(dims)},
                "out_proj": {"kernel": kern(dims, dims), "bias": vec(dims)},
            },
            "LayerNorm_1": ln(),
            "Dense_0": {"kernel": kern(dims, 4 * dims), "bias": vec(4 * dims)},
            "Dense_1": {"kernel": kern(4 * dims, dims), "bias": vec(dims)},
        }
    p["LayerNorm_0"] = ln()
    return p


# ---------------------------------------------------------------- timed CPU baseline
def audio_encoder_torch_cpu(params: dict, mel: np.ndarray, n_heads: int, pos_table=None):
    """Same restatement on torch-CPU float32 with all host threads, used only as the
    *timed* CPU baseline (bench.py ``cpu_baseline`` / ``--impl reference``): numpy's
    fp32 einsum/matmul path is what it is checked against in tests."""
    import torch
    import torch.nn.functional as Fn

    t = lambda a: torch.from_numpy(np.ascontiguousarray(a, np.float32))
    x = t(mel)

    def conv(x, p, stride):
        w = t(p["kernel"]).permute(2, 1, 0).contiguous()       # (out, in, k)
        return Fn.conv1d(x.transpose(1, 2), w, t(p["bias"]), stride=stride, padding=1).transpose(1, 2)

    def ln(x, p):
        mean = x.mean(-1, keepdim=True)
        var = torch.clamp((x * x).mean(-1, keepdim=True) - mean * mean, min=0.0)
        return (x - mean) * torch.rsqrt(var + 1e-6) * t(p["scale"]) + t(p["bias"])

    def lin(x, p):
        y = x @ t(p["kernel"])
        return y + t(p["bias"]) if "bias" in p else y

    gelu = lambda v: Fn.gelu(v, approximate="tanh")
    with torch.no_grad():
        x = gelu(conv(x, params["Conv_0"], 1))
        x = gelu(conv(x, params["Conv_1"], 2))
        B, T, D = x.shape
        x = x + t(pos_table if pos_table is not None else sinusoid_table(T, D))
        dh = D // n_heads
        n_layers = sum(1 for k in params if k.startswith("EncoderBlock_"))
        for i in range(n_layers):
            p = params[f"EncoderBlock_{i}"]
            a = p["MultiHeadAttention_0"]
            h = ln(x, p["LayerNorm_0"])
            q = lin(h, a["query_proj"]).view(B, T, n_heads, dh).transpose(1, 2) * dh ** -0.25
            k = lin(h, a["key_proj"]).view(B, T, n_heads, dh).transpose(1, 2) * dh ** -0.25
            v = lin(h, a["value_proj"]).view(B, T, n_heads, dh).transpose(1, 2)
            w = torch.softmax(q @ k.transpose(-1, -2), dim=-1)
            o = (w @ v).transpose(1, 2).reshape(B, T, D)
            x = x + lin(o, a["out_proj"])
            x = x + lin(gelu(lin(ln(x, p["LayerNorm_1"]), p["Dense_0"])), p["Dense_1"])
        x = ln(x, params["LayerNorm_0"])
    return x.numpy()


# ---------------------------------------------------------------- step after the path
def cross_attention_kv(text_decoder_params: dict, states: np.ndarray, n_layers: int, dtype=np.float64) -> np.ndarray:
    """Keys / values the cross attention of every DecoderBlock takes from the encoder states: whisper.py:104 calls
    ``MultiHeadAttention(q, kv)``, whose ``key_proj`` (no bias, :33) and ``value_proj`` act on ``kv`` (:45-46).
    Returns (n_layers, B, T, 2 * D_text) with [k | v] along the last axis (unscaled: :54-57 scales inside attention)."""
    x = np.asarray(states, dtype)
    out = []
    for i in range(n_layers):
        att = text_decoder_params[f"DecoderBlock_{i}"]["MultiHeadAttention_1"]
        k = x @ np.asarray(att["key_proj"]["kernel"], dtype)
        v = x @ np.asarray(att["value_proj"]["kernel"], dtype) + np.asarray(att["value_proj"]["bias"], dtype)
        out.append(np.concatenate([k, v], axis=-1))
    return np.stack(out)


def init_cross_kv_params(d_in: int, d_out: int, n_layers: int, seed: int = 42) -> dict:
    """Random float32 ``text_decoder`` sub-tree holding only the cross-attention key / value projections."""
    rng = np.random.default_rng(seed)
    p = {}
    for i in range(n_layers):
        p[f"DecoderBlock_{i}"] = {"MultiHeadAttention_1": {
            "key_proj": {"kernel": (rng.standard_normal((d_in, d_out)) / math.sqrt(d_in)).astype(np.float32)},
            "value_proj": {"kernel": (rng.standard_normal((d_in, d_out)) / math.sqrt(d_in)).astype(np.float32),
                           "bias": (0.1 * rng.standard_normal(d_out)).astype(np.float32)}}}
    return p
