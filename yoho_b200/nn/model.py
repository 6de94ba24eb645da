"""Drop-in for the audio-encoder half of the reference ``yoho/src/nn/model.py`` (YOHO's own model):

    Model(config, vocab_size).apply({"params": params}, spectogram, method=Model.encode_audio)
        -- train/stages/transcription_pretrain_no_voiceprints.py:198-203; model.py:82-83
    AudioEncoder(q_heads, kv_heads, dims, seq_len, blocks).apply({"params": params}, x)      model.py:9-29

Same constructor fields, the Flax ``apply(variables, *args, method=)`` convention and the reference's
parameter pytree (Flax auto-names: ``Conv_0, RMSNorm_0, Conv_1, RMSNorm_1, EncoderBlock_i/{RMSNorm_0,
GroupedQueryAttention_0/{DenseGeneral_0,1,2, Dense_0}, RMSNorm_1, SwiGLU_0/{Dense_0,1,2}}``).  The forward
pass is libyoho_b200's ``yoho_model_encoder_fwd``.  Forward only; the text decoder is out of scope.
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass

import numpy as np
import torch

from .._lib import COMPUTE_BF16, COMPUTE_FP32, ModelDesc, check
from ..runtime import get_runtime
from .whisper import _leaf

_COMPUTE = {"bf16": COMPUTE_BF16, "fp32": COMPUTE_FP32, COMPUTE_BF16: COMPUTE_BF16, COMPUTE_FP32: COMPUTE_FP32}


def rope_table(length: int, head_dim: int = 64) -> np.ndarray:
    """(length, 2, head_dim // 2) float32 {cos, sin} of the angles RoPE uses (layers.py:31-33), in the reference's
    float32 operation order: ``thetas = 10 ** -arange(dim // 2)``, ``outer(arange(1, S + 1), thetas)``."""
    f32 = np.float32
    thetas = np.power(f32(10.0), -np.arange(head_dim // 2, dtype=f32)).astype(f32)
    angles = np.outer(np.arange(1, length + 1).astype(f32), thetas).astype(f32)
    return np.stack([np.cos(angles), np.sin(angles)], axis=1).astype(f32)


def flatten_yoho_encoder_params(params: dict, n_blocks: int, table: np.ndarray) -> list:
    """Order the pytree leaves as ``yoho_model_weights_pack`` expects (include/yoho_b200.h)."""
    out = [_leaf(params, "Conv_0", "kernel"), _leaf(params, "Conv_0", "bias"), _leaf(params, "RMSNorm_0", "scale"),
           _leaf(params, "Conv_1", "kernel"), _leaf(params, "Conv_1", "bias"), _leaf(params, "RMSNorm_1", "scale"), table]
    for i in range(n_blocks):
        blk = _leaf(params, f"EncoderBlock_{i}")
        att, mlp = _leaf(blk, "GroupedQueryAttention_0"), _leaf(blk, "SwiGLU_0")
        out += [_leaf(blk, "RMSNorm_0", "scale"), _leaf(att, "DenseGeneral_0", "kernel"), _leaf(att, "DenseGeneral_1", "kernel"),
                _leaf(att, "DenseGeneral_2", "kernel"), _leaf(att, "Dense_0", "kernel"), _leaf(blk, "RMSNorm_1", "scale"),
                _leaf(mlp, "Dense_0", "kernel"), _leaf(mlp, "Dense_1", "kernel"), _leaf(mlp, "Dense_2", "kernel")]
    return out


def init_yoho_encoder_params(n_mels: int, dims: int, q_heads: int, kv_heads: int, n_blocks: int, seed: int = 42) -> dict:
    """Random-init float32 pytree for benchmarks and tests (kernels N(0, 1/fan_in), non-trivial biases and
    RMSNorm scales).  Same generator order as ``oracle.yoho_model.init_params``."""
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


@dataclass
class AudioEncoder:
    """Reference ``AudioEncoder`` (model.py:9-29): Conv(k3) -> RMSNorm -> SiLU -> max-pool over time pairs ->
    Conv(k3) -> RMSNorm -> SiLU -> ``blocks`` x EncoderBlock (RMSNorm, GQA with RoPE, SwiGLU).
    ``(B, S, n_mels) -> (B, S // 2, dims)``."""
    q_heads: int
    kv_heads: int
    dims: int
    seq_len: int
    blocks: int

    def __post_init__(self):
        self._rope = None

    def rope(self) -> np.ndarray:
        if self._rope is None:
            self._rope = rope_table(self.seq_len // 2, self.dims // self.q_heads)
        return self._rope

    def apply(self, variables: dict, x, compute="bf16", out: torch.Tensor | None = None):
        params = variables["params"] if "params" in variables else variables
        return self._forward(params, x, compute, out)

    def _forward(self, params: dict, mel, compute, out=None):
        if getattr(mel, "ndim", None) != 3:
            raise ValueError(f"audio must be (batch, frames, n_mels); got ndim={getattr(mel, 'ndim', None)}")
        host = not (isinstance(mel, torch.Tensor) and mel.is_cuda)
        rt = get_runtime(None if host else mel.device.index)
        x = rt.to_device(mel, "mel", torch.float32)
        B, S, M = x.shape
        kshape = tuple(np.shape(_leaf(params, "Conv_0", "kernel")))
        if kshape != (3, M, self.dims):
            raise ValueError(f"Conv_0 kernel {kshape} does not match input (3, {M}, {self.dims})")
        if S // 2 > self.seq_len // 2:
            raise ValueError(f"{S} frames exceed seq_len={self.seq_len}")
        desc = ModelDesc(M, self.dims, self.q_heads, self.kv_heads, self.blocks, self.seq_len)
        blob = rt.packed_weights(params, desc, flatten_yoho_encoder_params(params, self.blocks, self.rope()), family="model")
        mode = _COMPUTE[compute]
        if out is None:
            out = torch.empty((B, S // 2, self.dims), dtype=torch.float32, device=rt.device)
        ws_bytes = rt.lib.yoho_model_workspace_bytes(C.byref(desc), B, S, mode)
        ws = rt.workspace("model", ws_bytes)
        with torch.cuda.device(rt.device):
            check(rt.lib.yoho_model_encoder_fwd(rt.ctx, C.byref(desc), C.c_void_p(rt.aligned_ptr(blob)), C.c_void_p(x.data_ptr()),
                                                B, S, C.c_void_p(out.data_ptr()), mode, C.c_void_p(rt.aligned_ptr(ws)), ws_bytes,
                                                rt.stream_ptr()))
        return out.cpu().numpy() if host else out


class Model:
    """Reference ``Model(config, vocab_size)`` (model.py:54-86), encoder side only: ``setup`` builds the encoder with
    ``kv_heads = n_audio_heads // 2`` (model.py:60-66).  ``decode_text`` / ``__call__`` raise (out of scope)."""

    def __init__(self, config, vocab_size: int = 0):
        self.config = config
        self.vocab_size = vocab_size
        self.encoder = AudioEncoder(config.n_audio_heads, config.n_audio_heads // 2, config.dims, config.max_audio_len,
                                    config.n_audio_blocks)

    # `method=Model.encode_audio` passes this function object; apply() dispatches on it.
    def encode_audio(self, audio, *, _params=None, compute="bf16", out=None):
        if _params is None:
            raise TypeError("call through model.apply({'params': params}, audio, method=Model.encode_audio)")
        return self.encoder._forward(_params, audio, compute, out)

    def decode_text(self, text, audio):
        raise NotImplementedError("the text decoder (model.py:32-51) is outside the accelerated hot path")

    def apply(self, variables: dict, *args, method=None, compute="bf16", out=None):
        if method is None or getattr(method, "__name__", None) != "encode_audio":
            raise NotImplementedError("only method=Model.encode_audio is accelerated (model.py:82-83)")
        params = variables["params"] if "params" in variables else variables
        params = params["encoder"] if "encoder" in params else params
        return self.encode_audio(*args, _params=params, compute=compute, out=out)
