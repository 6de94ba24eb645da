"""Drop-in for the reference ``yoho/src/nn/whisper.py``:

    Whisper(...).apply({"params": params}, mel, method=Whisper.encode_audio)     whisper.py:194-195, 258
    AudioEncoder(seq_len, dims, n_heads, n_layers).apply({"params": params}, x)  whisper.py:117-135
    Whisper(...).apply({"params": params}, states, tokens, method=Whisper.decode_text)   whisper.py:197-198
    TextDecoder(vocab_size, seq_len, dims, n_heads, n_layers).apply({"params": p}, tokens, states)   whisper.py:138-165
    TextDecoder.generate(...)   the greedy loop of whisper.py:260-285 with a KV cache (the reference has none, :27)

The classes keep the reference's constructor fields and the Flax ``apply(variables, *args, method=)``
calling convention, and consume the reference's parameter pytree unchanged (Flax auto-names:
``Conv_0``, ``Conv_1``, ``EncoderBlock_i/{LayerNorm_0, MultiHeadAttention_0/{query_proj,key_proj,
value_proj,out_proj}, LayerNorm_1, Dense_0, Dense_1}``, ``LayerNorm_0``).  The forward pass itself is
libyoho_b200's ``yoho_encoder_fwd`` (tcgen05 kernels in bf16 mode, SIMT fp32 kernels in fp32 mode); the text decoder
runs libyoho_b200's float32 kernels (``yoho_decoder_*``).  Forward only: the backward pass is out of scope.
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass

import numpy as np
import torch

from .._lib import COMPUTE_BF16, COMPUTE_FP32, DTYPE_BF16, DTYPE_F32, CrossKvDesc, DecoderDesc, EncDesc, check
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
            check(rt.lib.yoho_encoder_fwd(rt.ctx, C.byref(desc), C.c_void_p(rt.aligned_ptr(blob)), rt.blob_bytes(blob),
                                          C.c_void_p(x.data_ptr()), B, F, C.c_void_p(out.data_ptr()),
                                          DTYPE_BF16 if out.dtype == torch.bfloat16 else DTYPE_F32, mode,
                                          C.c_void_p(rt.aligned_ptr(ws)), ws_bytes, rt.stream_ptr()))
        return out.cpu().numpy() if host else out


@dataclass
class Whisper:
    """Reference ``Whisper`` module (whisper.py:168-203): ``encode_audio``, ``decode_text`` and ``__call__`` through
    ``apply(variables, *args, method=...)`` on ``{"params": {"audio_encoder": ..., "text_decoder": ...}}``."""
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
        self.text_decoder = (TextDecoder(vocab_size=self.text_vocab_size, seq_len=self.text_seq_len, dims=self.text_dims,
                                         n_heads=self.text_heads, n_layers=self.text_layers) if self.text_vocab_size else None)

    # `method=Whisper.encode_audio` passes this function object; apply() dispatches on it.
    def encode_audio(self, mel, *, _params=None, compute="bf16", out=None):
        if _params is None:
            raise TypeError("call through model.apply({'params': params}, mel, method=Whisper.encode_audio)")
        return self.audio_encoder._forward(_params, mel, compute, out)

    def decode_text(self, encoded_audio, tokens, *, _params=None):
        """``Whisper.decode_text(encoded_audio, tokens)`` (whisper.py:197-198): logits (B, S, vocab)."""
        if _params is None:
            raise TypeError("call through model.apply({'params': params}, encoded_audio, tokens, method=Whisper.decode_text)")
        if self.text_decoder is None:
            raise ValueError("this Whisper was built without text_* fields")
        return self.text_decoder._forward(_params, tokens, encoded_audio)

    def apply(self, variables: dict, *args, method=None, compute="bf16", out=None):
        params = variables["params"] if "params" in variables else variables
        name = getattr(method, "__name__", None)
        if name == "encode_audio":
            params = params["audio_encoder"] if "audio_encoder" in params else params
            return self.encode_audio(*args, _params=params, compute=compute, out=out)
        if name == "decode_text":
            params = params["text_decoder"] if "text_decoder" in params else params
            return self.decode_text(*args, _params=params)
        if method is None:                       # __call__(mel, tokens): whisper.py:200-203
            mel, tokens = args
            states = self.encode_audio(mel, _params=params["audio_encoder"], compute=compute)
            return self.decode_text(states, tokens, _params=params["text_decoder"])
        raise NotImplementedError(f"method {name!r} is not part of the reference Whisper module")


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


# ---------------------------------------------------------------------------------------------- text decoder
def flatten_decoder_params(params: dict, n_layers: int) -> list:
    """Order the TextDecoder pytree leaves as ``yoho_decoder_weights_pack`` expects (include/yoho_b200.h)."""
    out = [_leaf(params, "positional_embedding"), _leaf(params, "Embed_0", "embedding"),
           _leaf(params, "LayerNorm_0", "scale"), _leaf(params, "LayerNorm_0", "bias")]

    def mha(att):
        return [_leaf(att, "query_proj", "kernel"), _leaf(att, "query_proj", "bias"), _leaf(att, "key_proj", "kernel"),
                _leaf(att, "value_proj", "kernel"), _leaf(att, "value_proj", "bias"),
                _leaf(att, "out_proj", "kernel"), _leaf(att, "out_proj", "bias")]

    for i in range(n_layers):
        blk = _leaf(params, f"DecoderBlock_{i}")
        out += [_leaf(blk, "LayerNorm_0", "scale"), _leaf(blk, "LayerNorm_0", "bias")] + mha(_leaf(blk, "MultiHeadAttention_0"))
        out += [_leaf(blk, "LayerNorm_1", "scale"), _leaf(blk, "LayerNorm_1", "bias")] + mha(_leaf(blk, "MultiHeadAttention_1"))
        out += [_leaf(blk, "LayerNorm_2", "scale"), _leaf(blk, "LayerNorm_2", "bias"),
                _leaf(blk, "Dense_0", "kernel"), _leaf(blk, "Dense_0", "bias"),
                _leaf(blk, "Dense_1", "kernel"), _leaf(blk, "Dense_1", "bias")]
    return out


def init_decoder_params(vocab: int, seq_len: int, dims: int, n_layers: int, seed: int = 42, audio_dims: int | None = None) -> dict:
    """Random-init float32 TextDecoder pytree for benchmarks and tests.  Same generator order as
    ``oracle.decoder.init_params``."""
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


@dataclass
class TextDecoder:
    """Reference ``TextDecoder`` (whisper.py:138-165): token embedding + learned positions, ``n_layers`` DecoderBlocks
    (masked self attention, cross attention over the encoder states, MLP), LayerNorm, logits against the tied embedding.
    ``apply({"params": p}, tokens (B, S), states (B, T, D_audio)) -> logits (B, S, vocab)`` float32."""
    vocab_size: int
    seq_len: int
    dims: int
    n_heads: int
    n_layers: int

    def apply(self, variables: dict, q, kv):
        params = variables["params"] if "params" in variables else variables
        return self._forward(params, q, kv)

    # ---- plumbing ---------------------------------------------------------------------------
    def _desc(self, params: dict, audio_dims: int, audio_len: int) -> DecoderDesc:
        eshape = tuple(np.shape(_leaf(params, "Embed_0", "embedding")))
        if eshape != (self.vocab_size, self.dims):
            raise ValueError(f"Embed_0.embedding {eshape} does not match (vocab_size, dims) = ({self.vocab_size}, {self.dims})")
        return DecoderDesc(self.vocab_size, self.seq_len, self.dims, self.n_heads, self.n_layers, audio_dims, audio_len)

    def _blob(self, rt, params: dict, desc: DecoderDesc):
        return rt.packed_weights(params, desc, flatten_decoder_params(params, self.n_layers), family="decoder")

    @staticmethod
    def _tokens(rt, tokens) -> torch.Tensor:
        if isinstance(tokens, torch.Tensor) and tokens.is_cuda:
            return tokens.to(torch.int32).contiguous()
        return torch.as_tensor(np.ascontiguousarray(np.asarray(tokens)).astype(np.int32)).to(rt.device)

    # ---- teacher-forced forward -----------------------------------------------------------------
    def _forward(self, params: dict, tokens, states):
        if getattr(tokens, "ndim", None) != 2 or getattr(states, "ndim", None) != 3:
            raise ValueError("tokens must be (batch, positions) and encoded audio (batch, audio positions, dims)")
        host = not (isinstance(states, torch.Tensor) and states.is_cuda)
        rt = get_runtime(None if host else states.device.index)
        kv = rt.to_device(states, "states", torch.float32)
        tok = self._tokens(rt, tokens)
        B, S = tok.shape
        if kv.shape[0] != B:
            raise ValueError(f"{B} token rows but {kv.shape[0]} clips of encoded audio")
        if S > self.seq_len:
            raise ValueError(f"{S} tokens but positional_embedding has {self.seq_len} rows (whisper.py:155)")
        desc = self._desc(params, kv.shape[2], kv.shape[1])
        blob = self._blob(rt, params, desc)
        logits = torch.empty((B, S, self.vocab_size), dtype=torch.float32, device=rt.device)
        ws_bytes = rt.lib.yoho_decoder_workspace_bytes(C.byref(desc), B, S)
        ws = rt.workspace("decoder", ws_bytes)
        with torch.cuda.device(rt.device):
            check(rt.lib.yoho_decoder_fwd(rt.ctx, C.byref(desc), C.c_void_p(rt.aligned_ptr(blob)), C.c_void_p(tok.data_ptr()),
                                          C.c_void_p(kv.data_ptr()), B, S, C.c_void_p(logits.data_ptr()),
                                          C.c_void_p(rt.aligned_ptr(ws)), ws_bytes, rt.stream_ptr()))
        return logits.cpu().numpy() if host else logits

    # ---- greedy generation with a KV cache --------------------------------------------------------
    def generate(self, variables: dict, states, tokens, forced=None, n_steps: int | None = None, eot: int = -1,
                 return_logits: bool = False, use_graph: bool = False):
        """The loop of whisper.py:260-285 for a batch of clips: ``tokens`` (B, seq_len) holds the pre-filled prompt
        (the reference sets positions 0, 2, 3, :266-269), ``forced`` (seq_len,) marks the positions that keep their
        token; every other position ``p`` in 1..n_steps receives ``argmax(logits[:, p - 1])``.  The cross-attention keys /
        values are computed once, the self-attention keys / values are cached, and the position lives on the device:
        ``n_steps`` tokens are generated without a host synchronisation.  ``use_graph`` replays one captured step.
        Returns the token buffer (and the per-step logits (n_steps, B, vocab) when asked)."""
        params = variables["params"] if "params" in variables else variables
        params = params["text_decoder"] if "text_decoder" in params else params
        host = not (isinstance(states, torch.Tensor) and states.is_cuda)
        rt = get_runtime(None if host else states.device.index)
        kv = rt.to_device(states, "states", torch.float32)
        tok = self._tokens(rt, tokens).clone()
        B = tok.shape[0]
        if tok.shape[1] != self.seq_len:
            raise ValueError(f"token buffer must be (batch, {self.seq_len}); got {tuple(tok.shape)}")
        n_steps = self.seq_len - 1 if n_steps is None else int(n_steps)
        if not 0 <= n_steps <= self.seq_len - 1:
            raise ValueError(f"n_steps={n_steps} outside 0..{self.seq_len - 1}")
        desc = self._desc(params, kv.shape[2], kv.shape[1])
        blob = self._blob(rt, params, desc)
        lib = rt.lib
        fm = None
        if forced is not None:
            fm = torch.as_tensor(np.ascontiguousarray(np.asarray(forced)).astype(np.uint8)).to(rt.device)
            if fm.shape != (self.seq_len,):
                raise ValueError(f"forced must be ({self.seq_len},)")
        fptr = C.c_void_p(fm.data_ptr()) if fm is not None else C.c_void_p(0)
        cache_bytes = lib.yoho_decoder_cache_bytes(C.byref(desc), B)
        cache = rt.workspace("decoder_cache", cache_bytes)
        ws_bytes = lib.yoho_decoder_step_workspace_bytes(C.byref(desc), B)
        ws = rt.workspace("decoder_step", ws_bytes)
        bp, cp, wp = C.c_void_p(rt.aligned_ptr(blob)), C.c_void_p(rt.aligned_ptr(cache)), C.c_void_p(rt.aligned_ptr(ws))
        all_logits = torch.empty((n_steps, B, self.vocab_size), dtype=torch.float32, device=rt.device) if return_logits else None
        with torch.cuda.device(rt.device):
            check(lib.yoho_decoder_begin(rt.ctx, C.byref(desc), bp, C.c_void_p(kv.data_ptr()), B, cp, cache_bytes, rt.stream_ptr()))
            if return_logits:
                for i in range(n_steps):
                    check(lib.yoho_decoder_step(rt.ctx, C.byref(desc), bp, C.c_void_p(tok.data_ptr()), B, cp, cache_bytes,
                                                C.c_void_p(all_logits[i].data_ptr()), fptr, eot, wp, ws_bytes, rt.stream_ptr()))
            elif use_graph and n_steps > 0:
                side = torch.cuda.Stream(device=rt.device)
                side.wait_stream(torch.cuda.current_stream(rt.device))
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.stream(side):
                    # warm-up outside capture (function attributes are set on first launch), then restore the state
                    snap_tok, snap_hdr = tok.clone(), cache[:2048].clone()
                    check(lib.yoho_decoder_step(rt.ctx, C.byref(desc), bp, C.c_void_p(tok.data_ptr()), B, cp, cache_bytes, C.c_void_p(0),
                                                fptr, eot, wp, ws_bytes, rt.stream_ptr()))
                    tok.copy_(snap_tok)
                    cache[:2048].copy_(snap_hdr)
                    with torch.cuda.graph(graph, stream=side, capture_error_mode="relaxed"):
                        check(lib.yoho_decoder_step(rt.ctx, C.byref(desc), bp, C.c_void_p(tok.data_ptr()), B, cp, cache_bytes, C.c_void_p(0),
                                                    fptr, eot, wp, ws_bytes, rt.stream_ptr()))
                torch.cuda.current_stream(rt.device).wait_stream(side)
                for _ in range(n_steps):
                    graph.replay()
                self._graph = graph        # keep the executable graph alive until the replays have run
            else:
                check(lib.yoho_decoder_generate(rt.ctx, C.byref(desc), bp, C.c_void_p(tok.data_ptr()), B, cp, cache_bytes, fptr, eot,
                                                n_steps, wp, ws_bytes, rt.stream_ptr()))
        out = tok.cpu().numpy() if host else tok
        if return_logits:
            return out, (all_logits.cpu().numpy() if host else all_logits)
        return out
