"""End-to-end clip pipeline: PCM -> log-mel -> Whisper audio encoder, the composition the reference
runs at transcription_pretrain_no_voiceprints.py:172-179 + :198-203 and whisper.py:255-258
(``mel_spectogram`` -> ``normalize_spectogram`` -> ``model.apply(..., method=Whisper.encode_audio)``), and
``Transcriber``: the same followed by the greedy token loop of whisper.py:260-285 on the KV-cached decoder.

``ClipEncoder.encode`` keeps everything in HBM (device tensors in, device tensor out).
``ClipEncoder.encode_host`` is the host-facing call: pinned host PCM in, pinned host encoder states
out, with the H2D copy of chunk i+1 and the D2H copy of chunk i-1 overlapped with the kernels of
chunk i on three CUDA streams (double-buffered device staging).  All arithmetic is libyoho_b200.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import MODEL_SHAPES
from ._lib import COMPUTE_BF16, COMPUTE_FP32, DTYPE_BF16, DTYPE_F32, NORM_NONE, NORM_REFERENCE, NORM_WHISPER, EncDesc, check
from .nn.whisper import conv_out_len, flatten_encoder_params, sin_positional_encoding
from .preprocessing.mel_filterbanks import mel_filter_banks
from .runtime import get_runtime

_NORM = {"none": NORM_NONE, "reference": NORM_REFERENCE, "whisper": NORM_WHISPER}
_COMPUTE = {"bf16": COMPUTE_BF16, "fp32": COMPUTE_FP32}
_STATES = {"fp32": torch.float32, "float32": torch.float32, "bf16": torch.bfloat16, "bfloat16": torch.bfloat16,
           torch.float32: torch.float32, torch.bfloat16: torch.bfloat16}


def chunk_plan(batch: int, chunk: int, taper: bool = True) -> list[tuple[int, int]]:
    """[(first clip, clips)] of the pipelined host path.  The copies of all but the first chunk's input and
    the last chunk's output hide behind kernels, so with ``taper`` the first chunk is a quarter and the last
    three are a half and two quarters of ``chunk`` (when the batch holds at least two chunks)."""
    sizes: list[int] = []
    q = chunk // 4
    if taper and q >= 1 and batch >= 2 * chunk:
        head, tail = [q, chunk - q], [chunk // 2, chunk - chunk // 2 - q, q]
        body = batch - 2 * chunk
        sizes = head + [chunk] * (body // chunk) + ([body % chunk] if body % chunk else []) + tail
    else:
        sizes = [chunk] * (batch // chunk) + ([batch % chunk] if batch % chunk else [])
    plan, b0 = [], 0
    for n in sizes:
        plan.append((b0, n))
        b0 += n
    assert b0 == batch
    return plan


class ClipEncoder:
    """Log-mel + encoder over batches of fixed-length clips on one GPU.

    model: a name in ``MODEL_SHAPES`` or ``(n_mels, dims, n_heads, n_layers)``; params: the
    reference's AudioEncoder parameter pytree (or ``{"params": {"audio_encoder": ...}}``).
    """

    def __init__(self, model, params: dict, n_samples: int = 480000, norm: str = "reference", compute: str = "bf16",
                 sample_rate: int = 16000, n_fft: int = 400, hop: int = 160, device: int | None = None,
                 states_dtype="fp32"):
        shape = MODEL_SHAPES[model] if isinstance(model, str) else tuple(model)
        self.n_mels, self.dims, self.n_heads, self.n_layers = shape
        self.rt = get_runtime(device)
        lib = self.rt.lib
        self.n_samples, self.n_fft, self.hop = int(n_samples), n_fft, hop
        self.frames = check(lib.yoho_stft_frames(self.n_samples, n_fft, hop))
        self.seq_len = conv_out_len(self.frames)
        self.norm, self.compute = _NORM[norm], _COMPUTE[compute]
        # dtype of the encoder states handed back: float32 is what the reference returns; bfloat16 (rounded once, by the
        # final LayerNorm) halves the D2H bytes of the host-facing path
        self.states_dtype = _STATES[states_dtype]
        if self.states_dtype == torch.bfloat16 and self.compute != COMPUTE_BF16:
            raise ValueError("bf16 encoder states need compute='bf16'")
        self.fb = self.rt.filterbank(mel_filter_banks(sample_rate, n_fft, self.n_mels))
        params = params.get("params", params)
        params = params.get("audio_encoder", params)
        self.desc = EncDesc(self.n_mels, self.dims, self.n_heads, self.n_layers, self.seq_len)
        tensors = flatten_encoder_params(params, self.n_layers, sin_positional_encoding(self.seq_len, self.dims))
        self.blob = self.rt.packed_weights(params, self.desc, tensors)
        self._mel: dict[int, torch.Tensor] = {}
        self._stage: dict = {}
        self._streams = None
        self._graphs: dict = {}

    # ---- device-resident path -------------------------------------------------------------
    def _mel_buf(self, batch: int, slot: int = 0) -> torch.Tensor:
        buf = self._mel.get(slot)
        if buf is None or buf.shape[0] < batch:
            buf = torch.empty((batch, self.frames, self.n_mels), dtype=torch.float32, device=self.rt.device)
            self._mel[slot] = buf
        return buf[:batch]

    def logmel(self, pcm: torch.Tensor, out: torch.Tensor | None = None, ws_tag: str = "logmel") -> torch.Tensor:
        """(B, n_samples) float32/int16 CUDA tensor -> (B, frames, n_mels) float32, on the current stream."""
        rt, lib = self.rt, self.rt.lib
        B = pcm.shape[0]
        if pcm.shape[1] != self.n_samples or not pcm.is_cuda or not pcm.is_contiguous():
            raise ValueError(f"pcm must be a contiguous CUDA tensor (B, {self.n_samples}); got {tuple(pcm.shape)} on {pcm.device}")
        if out is None:
            out = torch.empty((B, self.frames, self.n_mels), dtype=torch.float32, device=rt.device)
        ws_bytes = lib.yoho_logmel_workspace_bytes(B, self.n_samples, self.n_fft, self.hop, self.n_mels)
        ws = rt.workspace(ws_tag, ws_bytes)
        fn = lib.yoho_logmel_s16 if pcm.dtype == torch.int16 else lib.yoho_logmel_f32
        check(fn(rt.ctx, self.fb, C.c_void_p(pcm.data_ptr()), B, self.n_samples, self.n_fft, self.hop,
                 C.c_void_p(out.data_ptr()), self.norm, C.c_void_p(rt.aligned_ptr(ws)), ws_bytes, rt.stream_ptr()))
        return out

    def encoder(self, mel: torch.Tensor, out: torch.Tensor | None = None, ws_tag: str = "encoder") -> torch.Tensor:
        """(B, frames, n_mels) float32 CUDA tensor -> (B, seq_len, dims) float32, on the current stream."""
        rt, lib = self.rt, self.rt.lib
        B = mel.shape[0]
        if out is None:
            out = torch.empty((B, self.seq_len, self.dims), dtype=self.states_dtype, device=rt.device)
        if out.dtype not in (torch.float32, torch.bfloat16) or not out.is_contiguous():
            raise ValueError("encoder states go to a contiguous float32 or bfloat16 tensor")
        ws_bytes = lib.yoho_encoder_workspace_bytes(C.byref(self.desc), B, self.frames, self.compute)
        ws = rt.workspace(ws_tag, ws_bytes)
        check(lib.yoho_encoder_fwd(rt.ctx, C.byref(self.desc), C.c_void_p(rt.aligned_ptr(self.blob)), rt.blob_bytes(self.blob),
                                   C.c_void_p(mel.data_ptr()), B, self.frames, C.c_void_p(out.data_ptr()),
                                   DTYPE_BF16 if out.dtype == torch.bfloat16 else DTYPE_F32, self.compute,
                                   C.c_void_p(rt.aligned_ptr(ws)), ws_bytes, rt.stream_ptr()))
        return out

    def encode(self, pcm: torch.Tensor, out: torch.Tensor | None = None, chunk_clips: int | None = None) -> torch.Tensor:
        """PCM resident in HBM -> encoder states resident in HBM (current stream, no host sync)."""
        B = pcm.shape[0]
        if out is None:
            out = torch.empty((B, self.seq_len, self.dims), dtype=self.states_dtype, device=self.rt.device)
        if B == 0:
            return out
        step = B if not chunk_clips else max(1, min(int(chunk_clips), B))
        with torch.cuda.device(self.rt.device):
            for b0 in range(0, B, step):
                b1 = min(B, b0 + step)
                mel = self.logmel(pcm[b0:b1], self._mel_buf(b1 - b0))
                self.encoder(mel, out[b0:b1])
        return out

    def encode_graphed(self, pcm: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
        """``encode`` replayed from a CUDA graph: the launch sequence of the library depends only on (shapes, pointers),
        so the ~60 launches of a small batch (BASELINE configs[0]: one clip, whisper.py:255-258) cost one graph launch
        instead of 60 launch latencies.  The graph is captured once per (pcm, out) buffer pair; both must stay alive and
        the caller refills ``pcm`` in place between calls."""
        B = pcm.shape[0]
        if out is None:
            out = torch.empty((B, self.seq_len, self.dims), dtype=self.states_dtype, device=self.rt.device)
        if B == 0:
            return out
        key = (pcm.data_ptr(), out.data_ptr(), B, pcm.dtype, out.dtype)
        entry = self._graphs.get(key)
        if entry is None:
            with torch.cuda.device(self.rt.device):
                self._reserve(B)
                side = torch.cuda.Stream(device=self.rt.device)
                side.wait_stream(torch.cuda.current_stream(self.rt.device))
                with torch.cuda.stream(side):
                    self.encode(pcm, out)                       # warm-up outside the capture (function attributes, workspaces)
                torch.cuda.current_stream(self.rt.device).wait_stream(side)
                torch.cuda.synchronize(self.rt.device)
                graph = torch.cuda.CUDAGraph()
                n0 = self.rt.lib.yoho_launch_count()
                with torch.cuda.graph(graph):
                    self.encode(pcm, out)
                self.graph_kernels = int(self.rt.lib.yoho_launch_count() - n0)   # kernel nodes replayed per call
            entry = self._graphs[key] = (graph, pcm, out)
        entry[0].replay()
        return out

    def _reserve(self, batch: int, lanes: int = 1) -> None:
        """Size the grow-only scratch buffers for ``batch`` clips up front, so that no chunk of a pipelined
        run reallocates one while kernels of the previous chunk still use it.  ``lanes`` > 1 reserves one set per
        compute stream of the host-facing path (tags / mel slots ``...{lane}``)."""
        lib = self.rt.lib
        lm = lib.yoho_logmel_workspace_bytes(batch, self.n_samples, self.n_fft, self.hop, self.n_mels)
        en = lib.yoho_encoder_workspace_bytes(C.byref(self.desc), batch, self.frames, self.compute)
        if lanes == 1:
            self._mel_buf(batch)
            self.rt.workspace("logmel", lm)
            self.rt.workspace("encoder", en)
        for lane in range(lanes if lanes > 1 else 0):
            self._mel_buf(batch, slot=100 + lane)
            self.rt.workspace(f"logmel{lane}", lm)
            self.rt.workspace(f"encoder{lane}", en)

    # ---- host-facing path -------------------------------------------------------------------
    def host_buffers(self, batch: int, pcm_dtype: torch.dtype = torch.float32) -> tuple[torch.Tensor, torch.Tensor]:
        """Pinned host buffers shaped for ``encode_host``: PCM (batch, n_samples) and states (batch, seq_len, dims)."""
        return (torch.empty((batch, self.n_samples), dtype=pcm_dtype, pin_memory=True),
                torch.empty((batch, self.seq_len, self.dims), dtype=self.states_dtype, pin_memory=True))

    def encode_host(self, pcm_host: torch.Tensor, out_host: torch.Tensor | None = None, chunk_clips: int = 64,
                    sync: bool = True, taper: bool = True) -> torch.Tensor:
        """Host PCM -> host encoder states.  ``pcm_host``: CPU tensor / ndarray (B, n_samples), float32
        or int16; pinned memory makes the H2D copies asynchronous (pageable memory still works, the
        copies then serialise with the host).  Returns ``out_host`` (pinned unless supplied)."""
        if isinstance(pcm_host, np.ndarray):
            pcm_host = torch.from_numpy(np.ascontiguousarray(pcm_host))
        B = pcm_host.shape[0]
        if out_host is None:
            out_host = torch.empty((B, self.seq_len, self.dims), dtype=self.states_dtype, pin_memory=True)
        if B == 0:                                              # the C ABI treats batch == 0 as a no-op; so does the pipeline
            return out_host
        dev = self.rt.device
        chunk = max(1, min(int(chunk_clips), (B + 1) // 2))     # at least two chunks, so that copies can hide behind kernels
        self.last_chunk_clips = chunk
        with torch.cuda.device(dev):
            if self._streams is None:
                self._streams = tuple(torch.cuda.Stream(device=dev) for _ in range(4))
            # one copy stream per direction and TWO compute streams: chunk i runs on compute stream i & 1 with its own
            # scratch, so the wave tails and launch gaps of one chunk's ~100 kernels are filled by the other chunk's
            # (chunked like this without copies: 94.2 ms on one stream, 91.1 ms on two, 89.4 ms as a whole batch)
            s_in, s_out, s_run0, s_run1 = self._streams
            s_runs = (s_run0, s_run1)
            key = (chunk, pcm_host.dtype, out_host.dtype)
            st = self._stage.get(key)
            if st is None:
                st = {"pcm": [torch.empty((chunk, self.n_samples), dtype=pcm_host.dtype, device=dev) for _ in range(2)],
                      "out": [torch.empty((chunk, self.seq_len, self.dims), dtype=out_host.dtype, device=dev) for _ in range(2)]}
                self._stage = {key: st}
            entry = torch.cuda.current_stream(dev)
            ready = torch.cuda.Event()
            ready.record(entry)
            for s in (s_in, s_out, s_run0, s_run1):
                s.wait_event(ready)
            self._reserve(chunk, lanes=2)
            ran: list = []
            drained: list = []
            for i, (b0, n) in enumerate(chunk_plan(B, chunk, taper)):
                b1, slot = b0 + n, i & 1
                with torch.cuda.stream(s_in):
                    if i >= 2:
                        s_in.wait_event(ran[i - 2])            # staging slot is free again
                    st["pcm"][slot][:n].copy_(pcm_host[b0:b1], non_blocking=True)
                    e_in = torch.cuda.Event()
                    e_in.record(s_in)
                s_run = s_runs[slot]
                with torch.cuda.stream(s_run):
                    s_run.wait_event(e_in)
                    if i >= 2:
                        s_run.wait_event(drained[i - 2])
                    mel = self.logmel(st["pcm"][slot][:n], self._mel_buf(n, slot=100 + slot), ws_tag=f"logmel{slot}")
                    self.encoder(mel, st["out"][slot][:n], ws_tag=f"encoder{slot}")
                    e_run = torch.cuda.Event()
                    e_run.record(s_run)
                    ran.append(e_run)
                with torch.cuda.stream(s_out):
                    s_out.wait_event(e_run)
                    out_host[b0:b1].copy_(st["out"][slot][:n], non_blocking=True)
                    e_out = torch.cuda.Event()
                    e_out.record(s_out)
                    drained.append(e_out)
            entry.wait_event(drained[-1])
            for e in ran[-2:]:
                entry.wait_event(e)
            if sync:
                drained[-1].synchronize()
        return out_host


class Transcriber:
    """PCM clips -> token ids: the reference's demo (whisper.py:249-285) for a batch, everything on the device.

    ``params`` is the reference's Whisper pytree ``{"audio_encoder": ..., "text_decoder": ...}`` (optionally under
    ``"params"``); ``model`` names the encoder shape, the decoder shape is read from the pytree.  The prompt layout is
    the caller's: ``prompt`` maps positions to forced token ids (the reference sets 0 -> <|startoftranscript|>,
    2 -> <|transcribe|>, 3 -> <|notimestamps|> and lets the model choose position 1, the language, :266-272)."""

    def __init__(self, model, params: dict, text_heads: int, text_seq_len: int = 448, n_samples: int = 480000, norm: str = "reference",
                 compute: str = "bf16", device: int | None = None):
        from .nn.whisper import TextDecoder
        params = params.get("params", params)
        self.encoder = ClipEncoder(model, params["audio_encoder"], n_samples=n_samples, norm=norm, compute=compute, device=device)
        self.dec_params = params["text_decoder"]
        vocab, dims = np.shape(self.dec_params["Embed_0"]["embedding"])
        n_layers = sum(1 for k in self.dec_params if k.startswith("DecoderBlock_"))
        self.decoder = TextDecoder(vocab_size=int(vocab), seq_len=text_seq_len, dims=int(dims), n_heads=text_heads, n_layers=n_layers)

    def transcribe(self, pcm, prompt: dict[int, int], max_new_tokens: int | None = None, eot: int = -1, use_graph: bool = False):
        """pcm (B, n_samples) float32 / int16 (host or CUDA) -> tokens (B, text_seq_len) int32 (same residency)."""
        rt = self.encoder.rt
        host = not (isinstance(pcm, torch.Tensor) and pcm.is_cuda)
        is16 = (pcm.dtype == torch.int16) if isinstance(pcm, torch.Tensor) else (np.asarray(pcm).dtype == np.int16)
        dev_pcm = rt.to_device(pcm, "pcm", torch.int16 if is16 else torch.float32)
        with torch.cuda.device(rt.device):
            states = self.encoder.encode(dev_pcm)
        L = self.decoder.seq_len
        tokens = torch.zeros((dev_pcm.shape[0], L), dtype=torch.int32, device=rt.device)
        forced = np.zeros(L, np.uint8)
        for pos, tok in prompt.items():
            tokens[:, pos] = int(tok)
            forced[pos] = 1
        n_steps = L - 1 if max_new_tokens is None else min(L - 1, int(max_new_tokens))
        out = self.decoder.generate({"params": self.dec_params}, states, tokens, forced=forced, n_steps=n_steps, eot=eot, use_graph=use_graph)
        return out.cpu().numpy() if host else out
