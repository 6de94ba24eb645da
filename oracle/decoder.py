"""CPU restatement of the reference Whisper text decoder (TEST INFRASTRUCTURE ONLY).

``TextDecoder.__call__`` (``yoho/src/nn/whisper.py:138-165``), its ``DecoderBlock`` (:91-114) and the
greedy loop of the reference's ``__main__`` (:260-285), which re-runs the whole 448-token decoder for every
generated token (``# TODO: add KV cache``, :27).  SURVEY.md 8f row 4.

Parity: unpinned by the reference (no tests, flax absent); cross-checked against ``transformers``
``WhisperDecoder`` in ``tests/test_oracle_witness.py``.  Library defaults as in ``oracle/encoder.py``.

Flax auto-names of the ``@nn.compact`` body::

    text_decoder = {
      "positional_embedding": (seq_len, D),                      # :148-152, learned
      "Embed_0": {"embedding": (vocab, D)},                      # :153, tied with the logits (:163)
      "DecoderBlock_i": {
          "LayerNorm_0", "MultiHeadAttention_0",                 # masked self attention (:96-99)
          "LayerNorm_1", "MultiHeadAttention_1",                 # cross attention over the encoder states (:101-105)
          "LayerNorm_2", "Dense_0": (D, 4D), "Dense_1": (4D, D)},# MLP (:107-112)
      "LayerNorm_0": {"scale", "bias"},                          # :162
    }
"""
from __future__ import annotations

import math

import numpy as np

from .encoder import dense, gelu_tanh, layer_norm, softmax_last


def attention(q_in, kv_in, p, n_head: int, causal: bool):
    """``MultiHeadAttention(q, kv, mask)`` with ``kv`` given (whisper.py:37-67): q = q_in Wq + bq (:43), k = kv Wk
    (no bias, :33,45), v = kv Wv + bv (:46); q and k scaled by d_head**-0.25 (:54-57); ``mask`` is
    ``triu(full(-inf), 1)`` added to the scores (:59-60, :157); softmax over keys; out_proj."""
    B, S, D = q_in.shape
    T = kv_in.shape[1]
    q = dense(q_in, p["query_proj"])
    k = dense(kv_in, p["key_proj"])
    v = dense(kv_in, p["value_proj"])
    dh = D // n_head
    scale = q_in.dtype.type(dh ** -0.25)
    q = q.reshape(B, S, n_head, dh).transpose(0, 2, 1, 3) * scale
    k = k.reshape(B, T, n_head, dh).transpose(0, 2, 3, 1) * scale
    v = v.reshape(B, T, n_head, dh).transpose(0, 2, 1, 3)
    qk = q @ k
    if causal:
        qk = qk + np.triu(np.full((S, S), -np.inf, q_in.dtype), 1)
    o = (softmax_last(qk) @ v).transpose(0, 2, 1, 3).reshape(B, S, D)
    return dense(o, p["out_proj"])


def decoder_block(q, kv, p, n_head: int):
    """whisper.py:95-114."""
    h = layer_norm(q, p["LayerNorm_0"])
    q = q + attention(h, h, p["MultiHeadAttention_0"], n_head, causal=True)
    x = q + attention(layer_norm(q, p["LayerNorm_1"]), kv, p["MultiHeadAttention_1"], n_head, causal=False)
    h = gelu_tanh(dense(layer_norm(x, p["LayerNorm_2"]), p["Dense_0"]))
    return x + dense(h, p["Dense_1"])


def text_decoder(params: dict, tokens: np.ndarray, states: np.ndarray, n_heads: int, dtype=np.float64, taps: dict | None = None):
    """``TextDecoder.__call__(q=tokens, kv=states)``: tokens (B, S) int, states (B, T, D) -> logits (B, S, vocab)."""
    tokens = np.asarray(tokens)
    emb = np.asarray(params["Embed_0"]["embedding"], dtype)
    S = tokens.shape[1]
    x = emb[tokens] + np.asarray(params["positional_embedding"], dtype)[:S]
    kv = np.asarray(states, dtype)
    n_layers = sum(1 for k in params if k.startswith("DecoderBlock_"))
    for i in range(n_layers):
        x = decoder_block(x, kv, params[f"DecoderBlock_{i}"], n_heads)
        if taps is not None:
            taps[f"block{i}"] = x
    x = layer_norm(x, params["LayerNorm_0"])
    return x @ emb.T


def decode_text(variables: dict, encoded_audio, tokens, n_heads: int, dtype=np.float64):
    """``Whisper.decode_text`` (whisper.py:197-198) on ``{"params": {"text_decoder": ...}}`` or a bare tree."""
    params = variables.get("params", variables)
    params = params.get("text_decoder", params)
    return text_decoder(params, tokens, encoded_audio, n_heads, dtype)


def greedy_decode(params: dict, states: np.ndarray, n_heads: int, tokens: np.ndarray, forced: np.ndarray, n_steps: int,
                  dtype=np.float64) -> np.ndarray:
    """The loop at whisper.py:260-285 for a batch: positions where ``forced`` is set keep the given token, every other
    position ``pos`` in 1..n_steps receives ``argmax(logits[:, pos - 1])`` of a full decoder run over the buffer
    (causality makes the not-yet-written tail irrelevant).  tokens (B, L) is modified in a copy and returned."""
    tokens = np.array(tokens, copy=True)
    for pos in range(1, n_steps + 1):
        if forced[pos]:
            continue
        logits = text_decoder(params, tokens[:, :pos], states, n_heads, dtype)
        tokens[:, pos] = logits[:, pos - 1].argmax(axis=-1)
    return tokens


def init_params(vocab: int, seq_len: int, dims: int, n_layers: int, seed: int = 42, audio_dims: int | None = None) -> dict:
    """Random float32 TextDecoder pytree (no checkpoint offline): kernels N(0, 1/fan_in), embeddings N(0, 1),
    non-trivial biases and LayerNorm affines.  Same generator order as yoho_b200.nn.whisper.init_decoder_params."""
    rng = np.random.default_rng(seed)
    f32 = np.float32
    audio_dims = dims if audio_dims is None else audio_dims

    def kern(*shape):
        return (rng.standard_normal(shape) / math.sqrt(int(np.prod(shape[:-1])))).astype(f32)

    def vec(n, mean=0.0):
        return (mean + 0.1 * rng.standard_normal(n)).astype(f32)

    def ln():
        return {"scale": vec(dims, 1.0), "bias": vec(dims)}

    def mha(d_kv):
        return {"query_proj": {"kernel": kern(dims, dims), "bias": vec(dims)},
                "key_proj": {"kernel": kern(d_kv, dims)},
                "value_proj": {"kernel": kern(d_kv, dims), "bias": vec(dims)},
                "out_proj": {"kernel": kern(dims, dims), "bias": vec(dims)}}

    p = {"positional_embedding": (0.1 * rng.standard_normal((seq_len, dims))).astype(f32),
         "Embed_0": {"embedding": rng.standard_normal((vocab, dims)).astype(f32)}}
    for i in range(n_layers):
        p[f"DecoderBlock_{i}"] = {
            "LayerNorm_0": ln(), "MultiHeadAttention_0": mha(dims),
            "LayerNorm_1": ln(), "MultiHeadAttention_1": mha(audio_dims),
            "LayerNorm_2": ln(),
            "Dense_0": {"kernel": kern(dims, 4 * dims), "bias": vec(4 * dims)},
            "Dense_1": {"kernel": kern(4 * dims, dims), "bias": vec(dims)},
        }
    p["LayerNorm_0"] = ln()
    return p
