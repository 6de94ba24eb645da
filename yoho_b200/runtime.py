"""Per-device runtime: library context, filterbank / packed-weight caches, workspaces, host staging.

PyTorch is used only as plumbing here -- device allocations, pinned host buffers, the current CUDA
stream and (in bench.py) torch.distributed; every computation is a call into libyoho_b200.so.
"""
from __future__ import annotations

import ctypes as C
import threading

import numpy as np
import torch

from . import _lib
from ._lib import EncDesc, check


def _require_cuda():
    if not torch.cuda.is_available():
        raise RuntimeError("yoho_b200 needs a CUDA device (sm_100a); there is no CPU fallback")


class Runtime:
    """One per CUDA device; obtain with :func:`get_runtime`."""

    def __init__(self, device: int):
        _require_cuda()
        self.lib = _lib.load()
        self.device = torch.device("cuda", device)
        handle = C.c_void_p()
        with torch.cuda.device(self.device):       # the library restores the current device as well; belt and braces
            check(self.lib.yoho_ctx_create(device, C.byref(handle)))
        self.ctx = handle
        self._filterbanks: dict = {}
        self._workspaces: dict[str, torch.Tensor] = {}
        self._weights: dict = {}
        self._pinned: dict[str, torch.Tensor] = {}

    # ---- plumbing ---------------------------------------------------------------------
    def stream_ptr(self) -> C.c_void_p:
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def workspace(self, tag: str, nbytes: int) -> torch.Tensor:
        """Grow-only uint8 scratch buffer (1024-byte aligned by the caching allocator's 512 B
        granularity plus explicit over-allocation)."""
        buf = self._workspaces.get(tag)
        if buf is None or buf.numel() < nbytes + 1024:
            self._workspaces.pop(tag, None)
            buf = torch.empty(int(nbytes) + 1024, dtype=torch.uint8, device=self.device)
            self._workspaces[tag] = buf
        return buf

    @staticmethod
    def aligned_ptr(buf: torch.Tensor, align: int = 1024) -> int:
        p = buf.data_ptr()
        return (p + align - 1) // align * align

    @classmethod
    def blob_bytes(cls, blob: torch.Tensor) -> int:
        """Bytes usable behind ``aligned_ptr(blob)`` (what the ``*_fwd`` entry points check the descriptor's layout against)."""
        return int(blob.numel() - (cls.aligned_ptr(blob) - blob.data_ptr()))

    def pinned(self, tag: str, shape, dtype: torch.dtype) -> torch.Tensor:
        n = int(np.prod(shape))
        buf = self._pinned.get(tag)
        if buf is None or buf.dtype != dtype or buf.numel() < n:
            buf = torch.empty(n, dtype=dtype, pin_memory=True)
            self._pinned[tag] = buf
        return buf[:n].view(*shape)

    def to_device(self, array, tag: str, dtype: torch.dtype = torch.float32) -> torch.Tensor:
        """numpy / CPU tensor -> pinned staging -> async H2D on the current stream; CUDA tensors pass through."""
        if isinstance(array, torch.Tensor) and array.is_cuda:
            if array.device != self.device:
                raise ValueError(f"tensor on {array.device}, runtime on {self.device}")
            return array.to(dtype).contiguous()
        src = torch.as_tensor(np.ascontiguousarray(array)) if not isinstance(array, torch.Tensor) else array
        src = src.to(dtype)
        stage = self.pinned(tag, tuple(src.shape), dtype)
        stage.copy_(src)
        return stage.to(self.device, non_blocking=True)

    # ---- frontend constants -------------------------------------------------------------
    def filterbank(self, filters: np.ndarray):
        """Device filterbank handle for a dense (n_mels, n_bins) float32 matrix (cached by content)."""
        filters = np.ascontiguousarray(filters, np.float32)
        key = (filters.shape, filters.tobytes())
        fb = self._filterbanks.get(key)
        if fb is None:
            handle = C.c_void_p()
            with torch.cuda.device(self.device):
                check(self.lib.yoho_filterbank_create(self.ctx, filters.ctypes.data_as(C.c_void_p), filters.shape[0],
                                                      filters.shape[1], C.byref(handle)))
            fb = handle
            self._filterbanks[key] = fb
        return fb

    # ---- encoder weights ------------------------------------------------------------------
    def packed_weights(self, params: dict, desc, tensors: list, family: str = "encoder"):
        """Pack (once per params object) the float32 tensors into the device blob the kernels read.
        ``family``: "encoder" (yoho_encoder_*: Whisper AudioEncoder), "model" (yoho_model_*: YOHO's own encoder),
        "crosskv" (yoho_crosskv_*) or "decoder" (yoho_decoder_*: Whisper TextDecoder)."""
        # the blob layout depends on every desc field (the sinusoid / RoPE table is sized by seq_len and every later
        # offset follows it), so the descriptor is part of the key: one pytree used at two clip lengths gets two blobs
        key = (family, id(params), bytes(desc))
        hit = self._weights.get(key)
        if hit is not None and hit[0] is params:
            return hit[1]
        bytes_fn = getattr(self.lib, f"yoho_{family}_weights_bytes")
        pack_fn = getattr(self.lib, f"yoho_{family}_weights_pack")
        elems_fn = getattr(self.lib, f"yoho_{family}_tensor_elems")
        nbytes = bytes_fn(C.byref(desc))
        if nbytes == 0:
            raise _lib.YohoError(-1, self.lib.yoho_last_error().decode())
        # the C side copies tensor_elems(desc, i) floats from each pointer: a leaf of another shape would be read past its end
        for i, t in enumerate(tensors):
            want, have = int(elems_fn(C.byref(desc), i)), int(np.prod(np.shape(t)))
            if want != have:
                raise ValueError(f"{family} weight tensor {i} has {have} elements (shape {tuple(np.shape(t))}); "
                                 f"the descriptor {[getattr(desc, f) for f, _ in desc._fields_]} needs {want}")
        blob = torch.empty(nbytes + 1024, dtype=torch.uint8, device=self.device)
        keep, ptrs = [], (C.c_void_p * len(tensors))()
        for i, t in enumerate(tensors):
            if isinstance(t, torch.Tensor) and t.is_cuda:
                a = t.to(torch.float32).contiguous()
                ptrs[i] = a.data_ptr()
            else:
                a = np.ascontiguousarray(np.asarray(t), np.float32)
                ptrs[i] = a.ctypes.data
            keep.append(a)
        with torch.cuda.device(self.device):
            check(pack_fn(self.ctx, C.byref(desc), ptrs, len(tensors), C.c_void_p(self.aligned_ptr(blob)), nbytes,
                          self.stream_ptr()))
            torch.cuda.current_stream(self.device).synchronize()   # host sources may now be released
        self._weights[key] = (params, blob)
        return blob


_runtimes: dict[int, Runtime] = {}
_lock = threading.Lock()


def get_runtime(device: int | None = None) -> Runtime:
    _require_cuda()
    if device is None:
        device = torch.cuda.current_device()
    with _lock:
        rt = _runtimes.get(device)
        if rt is None:
            rt = _runtimes[device] = Runtime(device)
        return rt
