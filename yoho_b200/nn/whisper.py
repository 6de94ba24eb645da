"""Drop-in for the audio-encoder half of the reference ``yoho/src/nn/whisper.py``:

    Whisper(...).apply({"params": params}, mel, method=Whisper.encode_audio)     whisper.py:194-195, 258
    AudioEncoder(seq_len, dims, n_heads, n_layers).apply({"params": params}, x)  whisper.py:117-135

The classes keep the reference's constructor fields and the Flax ``apply(variables, *args, method=)``
calling convention, and consume the reference's parameter pytree unchanged (Flax auto-names:
``Conv_0``, ``Conv_1``, ``EncoderBlock_i/{LayerNorm_0, MultiHeadAttention_0/{query_proj,key_proj,
value_proj,out_proj}, LayerNorm_1, Dense_0, Dense_1}``, ``LayerNorm_0``).  The forward pass itself is
libyoho_b200's ``yoho_encoder_fwd`` (tcgen05 kernels in bf16 mode, SIMT fp32 kernels in fp32 mode).
Forward only: the text decoder and the backward pass are out of scope.
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass

import numpy as np
import torch

from .._lib import COMPUTE_BF16, COMPUTE_FP32, CrossKvDesc, EncDesc, check
from ..runtime import get_runtime

_COMPUTE = {"bf16": COMPUTE_BF16, "fp32": COMPUTE_FP32, COMPUTE_BF16: COMPUTE_BF16, COMPUTE_FP32: COMPUTE_FP32}


def sin_positional_encoding(length: int, dim: int, max_timescale: int = 10000) -> np.ndarray:
    """The constant ``SinPositionalEncoding`` adds (whisper.py:11-17), evaluated in float32 in the
    reference's operation order: ``inc = log(max_timescale)/(dim//2-1)``, ``inv = e**(-inc*arange)``,
    ``arange(length)[:,None]*inv[None,:]``, ``concat([sin, cos], axis=1)``."""
    if dim & 1:
        raise AssertionError("Number of dimensions must be even")
    f32 = np.float32
    inc = f32(np.log(f32(max_timescale))) / f32(dim // 2 - 1)
    inv = np.power(f32(np.e), (-inc * np.arange(dim // 2, dtype=f32)).astype(f32)).astype(f32)
    scaled = (np.arange(length, dtype=f32)[:, None] * inv[None, :]).astype(f32)
    return np.concatenate([np.sin(scaled), np.cos(scaled)], axis=1).astype(f32)


def conv_out_len(frames: int) -> int:
    """Length after ``nn.Conv(dims, 3, padding=1, strides=2)`` (whisper.py:127)."""
    return (frames + 2 - 3) // 2 + 1


def _leaf(tree: dict, *path):
    node = tree
    for key in path:
        if key not in node:
            raise KeyError(f"parameter {'/'.join(path)} missing from the pytree (have {sorted(node)})")
        node = node[key]
    return node


def flatten_encoder_params(params: dict, n_layers: int, pos_table: np.ndarray) -> list:
    """Order the pytree leaves as ``yoho_encoder_weights_pack`` expects (include/yoho_b200.h)."""
    out = [_leaf(params, "Conv_0", "kernel"), _leaf(params, "Conv_0", "bias"),
           _leaf(params, "Conv_1", "kernel"), _leaf(params, "Conv_1", "bias"),
           pos_table, _leaf(params, "LayerNorm_0", "scale"), _leaf(params, "LayerNorm_0", "bias")]
    for i in range(n_layers):
        blk = _leaf(params, f"EncoderBlock_{i}")
        att = _leaf(blk, "MultiHeadAttention_0")
        out += [_leaf(blk, "LayerNorm_0", "scale"), _leaf(blk, "LayerNorm_0", "bias"),
                _leaf(att, "query_proj", "kernel"), _leaf(att, "query_proj", "bias"),
                _leaf(att, "key_proj", "kernel"),
                _leaf(att, "value_proj", "kernel"), _leaf(att, "value_proj", "bias"),
                _leaf(att, "out_proj", "kernel"), _leaf(att, "out_proj", "bias"),
                _leaf(blk, "LayerNorm_1", "scale"), _leaf(blk, "LayerNorm_1", "bias"),
                _leaf(blk, "Dense_0", "kernel"), _leaf(blk, "Dense_0", "bias"),
                _leaf(blk, "Dense_1", "kernel"), _leaf(blk, "Dense_1", "bias")]
    return out


def init_encoder_params(n_mels: int, dims: int, n_layers: int, seed: int = 42) -> dict:
    """Random-init float32 AudioEncoder pytree for benchmarks and tests (no checkpoints are available
    offline): kernels N(0, 1/fan_in), non-trivial biases and LayerNorm affines so that every fused
    epilogue term matters.  Same generator order as ``oracle.encoder.init_params``."""
    rng = np.random.default_rng(seed)
    f32 = np.float32

    def kern(*shape):
        return (rng.standard_normal(shape) / math.sqrt(int(np.prod(shape[:-1])))).astype(f32)

    def vec(n, mean=0.0):
        return (mean + 0.1 * rng.standard_normal(n)).astype(f32)

    def ln():
        return {"scale": vec(dims, 1.0), "bias": vec(dims)}

    p = {"Conv_0": {"kernel": kern(3, n_mels, dims), "bias": vec(dims)},
         "Conv_1": {"kernel": kern(3, dims, dims), "bias": vec(dims)}}
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


@dataclass
class AudioEncoder:
    """Reference ``AudioEncoder`` (whisper.py:117-135): conv1d(k3)+GELU, conv1d(k3, stride 2)+GELU,
    sinusoid positions, ``n_layers`` pre-LN blocks, final LayerNorm.  ``(B, F, n_mels) -> (B, T, dims)``."""
    seq_len: int
    dims: int
    n_heads: int
    n_layers: int

    def __post_init__(self):
        self._pos = None

    def pos_table(self) -> np.ndarray:
        if self._pos is None:
            self._pos = sin_positional_encoding(self.seq_len, self.dims)
        return self._pos

    def apply(self, variables: dict, x, compute="bf16", out: torch.Tensor | None = None):
        params = variables["params"] if "params" in variables else variables
        return self._forward(params, x, compute, out)

    def _forward(self, params: dict, mel, compute, out=None):
        if getattr(mel, "ndim", None) != 3:
            raise ValueError(f"mel must be (batch, frames, n_mels); got ndim={getattr(mel, 'ndim', None)}")
        host = not (isinstance(mel, torch.Tensor) and mel.is_cuda)
        rt = get_runtime(None if host else mel.device.index)
        x = rt.to_device(mel, "mel", torch.float32)
        B, F, M = x.shape
        kshape = tuple(np.shape(_leaf(params, "Conv_0", "kernel")))
        if kshape != (3, M, self.dims):
            raise ValueError(f"Conv_0 kernel {kshape} does not match input (3, {M}, {self.dims})")
        T = conv_out_len(F)
        if T != self.seq_len:
            raise ValueError(f"{F} frames give {T} positions; the sinusoid table has {self.seq_len} rows "
                             "and is added by broadcasting (whisper.py:23)")
        desc = EncDesc(M, self.dims, self.n_heads, self.n_layers, self.seq_len)
        blob = rt.packed_weights(params, desc, flatten_encoder_params(params, self.n_layers, self.pos_table()))
        mode = _COMPUTE[compute]
        if out is None:
            out = torch.empty((B, T, self.dims), dtype=torch.float32, device=rt.device)
        ws_bytes = rt.lib.yoho_encoder_workspace_bytes(C.byref(desc), B, F, mode)
        ws = rt.workspace("encoder", ws_bytes)
        with torch.cuda.device(rt.device):
            check(rt.lib.yoho_encoder_fwd(rt.ctx, C.byref(desc), C.c_void_p(rt.aligned_ptr(blob)), C.c_void_p(x.data_ptr()),
                                          B, F, C.c_void_p(out.data_ptr()), mode, C.c_void_p(rt.aligned_ptr(ws)), ws_bytes,
                                          rt.stream_ptr()))
        return out.cpu().numpy() if host else out


@dataclass
class Whisper:
    """Reference ``Whisper`` module (whisper.py:168-203), encoder side only.  The text_* fields are kept
    so that existing construction code runs; ``decode_text`` / ``__call__`` raise (out of scope)."""
    audio_dims: int
    audio_seq_len: int
    audio_heads: int
    audio_layers: int
    text_vocab_size: int = 0
    text_seq_len: int = 0
    text_dims: int = 0
    text_heads: int = 0
    text_layers: int = 0

    def __post_init__(self):
        self.audio_encoder = AudioEncoder(seq_len=self.audio_seq_len, dims=self.audio_dims,
                                          n_heads=self.audio_heads, n_layers=self.audio_layers)

    # `method=Whisper.encode_audio` passes this function object; apply() dispatches on it.
    def encode_audio(self, mel, *, _params=None, compute="bf16", out=None):
        if _params is None:
            raise TypeError("call through model.apply({'params': params}, mel, method=Whisper.encode_audio)")
        return self.audio_encoder._forward(_params, mel, compute, out)

    def decode_text(self, encoded_audio, tokens):
        raise NotImplementedError("the text decoder (whisper.py:138-165) is outside the accelerated hot path")

    def apply(self, variables: dict, *args, method=None, compute="bf16", out=None):
        if method is None or getattr(method, "__name__", None) != "encode_audio":
            raise NotImplementedError("only method=Whisper.encode_audio is accelerated (whisper.py:194-195)")
        params = variables["params"] if "params" in variables else variables
        params = params["audio_encoder"] if "audio_encoder" in params else params
        return self.encode_audio(*args, _params=params, compute=compute, out=out)


def flatten_cross_kv_params(text_decoder_params: dict, n_layers: int) -> list:
    """``DecoderBlock_l/MultiHeadAttention_1/{key_proj.kernel, value_proj.kernel, value_proj.bias}`` in the order
    ``yoho_crosskv_weights_pack`` expects (MultiHeadAttention_1 is the cross attention, whisper.py:104)."""
    out = []
    for i in range(n_layers):
        att = _leaf(text_decoder_params, f"DecoderBlock_{i}", "MultiHeadAttention_1")
        out += [_leaf(att, "key_proj", "kernel"), _leaf(att, "value_proj", "kernel"), _leaf(att, "value_proj", "bias")]
    return out


def cross_attention_kv(text_decoder_params: dict, encoded_audio, n_layers: int, compute="bf16", out: torch.Tensor | None = None):
    """Keys and values every decoder layer's cross attention takes from the encoder states (whisper.py:104 ->
    :45-46), computed once per clip instead of once per generated token (the reference's ``decode_text`` loop,
    :274-281).  ``encoded_audio (B, T, D_audio) -> (n_layers, B, T, 2 * D_text)`` float32 with ``[k | v]`` along the
    last axis; the ``d_head ** -0.25`` scaling of ``k`` stays in the consumer's ``qkv_attention``."""
    if getattr(encoded_audio, "ndim", None) != 3:
        raise ValueError("encoded_audio must be (batch, positions, dims)")
    host = not (isinstance(encoded_audio, torch.Tensor) and encoded_audio.is_cuda)
    rt = get_runtime(None if host else encoded_audio.device.index)
    x = rt.to_device(encoded_audio, "states", torch.float32)
    B, T, D = x.shape
    tensors = flatten_cross_kv_params(text_decoder_params, n_layers)
    kshape = tuple(np.shape(tensors[0])) if tensors else (D, D)
    if kshape[0] != D:
        raise ValueError(f"key_proj kernel {kshape} does not take {D}-wide encoder states")
    desc = CrossKvDesc(D, kshape[1], n_layers)
    blob = rt.packed_weights(text_decoder_params, desc, tensors, family="crosskv")
    mode = _COMPUTE[compute]
    if out is None:
        out = torch.empty((n_layers, B, T, 2 * kshape[1]), dtype=torch.float32, device=rt.device)
    ws_bytes = rt.lib.yoho_crosskv_workspace_bytes(C.byref(desc), B, T, mode)
    ws = rt.workspace("crosskv", ws_bytes)
    with torch.cuda.device(rt.device):
        check(rt.lib.yoho_crosskv_fwd(rt.ctx, C.byref(desc), C.c_void_p(rt.aligned_ptr(blob)), C.c_void_p(x.data_ptr()), B, T,
                                      C.c_void_p(out.data_ptr()), mode, C.c_void_p(rt.aligned_ptr(ws)), ws_bytes, rt.stream_ptr()))
    return out.cpu().numpy() if host else out
