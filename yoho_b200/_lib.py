"""ctypes binding of ``libyoho_b200.so`` (include/yoho_b200.h).  No fallback: if the library is
missing the import fails loudly and tells the caller how to build it."""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import os

# YOHO_B200_LIB selects another build of the same library (A/B timing of kernel variants); default: the in-tree build
LIB_PATH = Path(os.environ.get("YOHO_B200_LIB") or (Path(__file__).resolve().parent / "lib" / "libyoho_b200.so"))

YOHO_OK = 0
NORM_NONE, NORM_REFERENCE, NORM_WHISPER = 0, 1, 2
COMPUTE_FP32, COMPUTE_BF16 = 0, 1
DTYPE_F32, DTYPE_BF16 = 0, 1
ABI_VERSION = 2


class YohoError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"libyoho_b200 error {code}: {message}")
        self.code = code


class EncDesc(C.Structure):
    _fields_ = [("n_mels", C.c_int32), ("dims", C.c_int32), ("n_heads", C.c_int32),
                ("n_layers", C.c_int32), ("seq_len", C.c_int32)]


class ModelDesc(C.Structure):
    _fields_ = [("n_mels", C.c_int32), ("dims", C.c_int32), ("q_heads", C.c_int32), ("kv_heads", C.c_int32),
                ("n_blocks", C.c_int32), ("seq_len", C.c_int32)]


class CrossKvDesc(C.Structure):
    _fields_ = [("d_in", C.c_int32), ("d_out", C.c_int32), ("n_layers", C.c_int32)]


class DecoderDesc(C.Structure):
    _fields_ = [("vocab", C.c_int32), ("seq_len", C.c_int32), ("dims", C.c_int32), ("n_heads", C.c_int32),
                ("n_layers", C.c_int32), ("audio_dims", C.c_int32), ("audio_len", C.c_int32)]


_p, _i, _sz = C.c_void_p, C.c_int, C.c_size_t

# name -> (restype, argtypes); mirrors include/yoho_b200.h one to one
SIGNATURES = {
    "yoho_abi_version": (_i, []),
    "yoho_last_error": (C.c_char_p, []),
    "yoho_xla_ffi_available": (_i, []),
    "yoho_ctx_create": (_i, [_i, C.POINTER(_p)]),
    "yoho_ctx_destroy": (_i, [_p]),
    "yoho_stft_frames": (_i, [_i, _i, _i]),
    "yoho_filterbank_create": (_i, [_p, _p, _i, _i, C.POINTER(_p)]),
    "yoho_filterbank_destroy": (_i, [_p]),
    "yoho_logmel_workspace_bytes": (_sz, [_i, _i, _i, _i, _i]),
    "yoho_logmel_f32": (_i, [_p, _p, _p, _i, _i, _i, _i, _p, _i, _p, _sz, _p]),
    "yoho_logmel_s16": (_i, [_p, _p, _p, _i, _i, _i, _i, _p, _i, _p, _sz, _p]),
    "yoho_logmel_static_max_tiles": (_sz, []),
    "yoho_logmel_f32_static": (_i, [_p, _p, _p, _i, _i, _i, _i, _p, _i, _p]),
    "yoho_chunk_count": (_i, [C.c_longlong, _i]),
    "yoho_logmel_chunked_f32": (_i, [_p, _p, _p, C.c_longlong, _i, _i, _i, _p, _i, _p, _sz, _p]),
    "yoho_logmel_chunked_s16": (_i, [_p, _p, _p, C.c_longlong, _i, _i, _i, _p, _i, _p, _sz, _p]),
    "yoho_normalize_workspace_bytes": (_sz, [_i, _i, _i]),
    "yoho_normalize_f32": (_i, [_p, _p, _i, _i, _i, _p, _i, _p, _sz, _p]),
    "yoho_encoder_num_tensors": (_i, [C.POINTER(EncDesc)]),
    "yoho_encoder_weights_bytes": (_sz, [C.POINTER(EncDesc)]),
    "yoho_encoder_weights_pack": (_i, [_p, C.POINTER(EncDesc), C.POINTER(_p), _i, _p, _sz, _p]),
    "yoho_encoder_out_len": (_i, [_i]),
    "yoho_encoder_workspace_bytes": (_sz, [C.POINTER(EncDesc), _i, _i, _i]),
    "yoho_encoder_tensor_elems": (C.c_longlong, [C.POINTER(EncDesc), _i]),
    "yoho_encoder_fwd": (_i, [_p, C.POINTER(EncDesc), _p, _sz, _p, _i, _i, _p, _i, _i, _p, _sz, _p]),
    "yoho_allgather_states": (_i, [_p, _p, _p, _p, _sz, _i, _p]),
    "yoho_crosskv_num_tensors": (_i, [C.POINTER(CrossKvDesc)]),
    "yoho_crosskv_tensor_elems": (C.c_longlong, [C.POINTER(CrossKvDesc), _i]),
    "yoho_crosskv_weights_bytes": (_sz, [C.POINTER(CrossKvDesc)]),
    "yoho_crosskv_weights_pack": (_i, [_p, C.POINTER(CrossKvDesc), C.POINTER(_p), _i, _p, _sz, _p]),
    "yoho_crosskv_workspace_bytes": (_sz, [C.POINTER(CrossKvDesc), _i, _i, _i]),
    "yoho_crosskv_fwd": (_i, [_p, C.POINTER(CrossKvDesc), _p, _p, _i, _i, _p, _i, _p, _sz, _p]),
    "yoho_decoder_num_tensors": (_i, [C.POINTER(DecoderDesc)]),
    "yoho_decoder_tensor_elems": (C.c_longlong, [C.POINTER(DecoderDesc), _i]),
    "yoho_decoder_weights_bytes": (_sz, [C.POINTER(DecoderDesc)]),
    "yoho_decoder_weights_pack": (_i, [_p, C.POINTER(DecoderDesc), C.POINTER(_p), _i, _p, _sz, _p]),
    "yoho_decoder_workspace_bytes": (_sz, [C.POINTER(DecoderDesc), _i, _i]),
    "yoho_decoder_fwd": (_i, [_p, C.POINTER(DecoderDesc), _p, _p, _p, _i, _i, _p, _p, _sz, _p]),
    "yoho_decoder_cache_bytes": (_sz, [C.POINTER(DecoderDesc), _i]),
    "yoho_decoder_step_workspace_bytes": (_sz, [C.POINTER(DecoderDesc), _i]),
    "yoho_decoder_begin": (_i, [_p, C.POINTER(DecoderDesc), _p, _p, _i, _p, _sz, _p]),
    "yoho_decoder_step": (_i, [_p, C.POINTER(DecoderDesc), _p, _p, _i, _p, _sz, _p, _p, _i, _p, _sz, _p]),
    "yoho_decoder_generate": (_i, [_p, C.POINTER(DecoderDesc), _p, _p, _i, _p, _sz, _p, _i, _i, _p, _sz, _p]),
    "yoho_model_num_tensors": (_i, [C.POINTER(ModelDesc)]),
    "yoho_model_tensor_elems": (C.c_longlong, [C.POINTER(ModelDesc), _i]),
    "yoho_model_weights_bytes": (_sz, [C.POINTER(ModelDesc)]),
    "yoho_model_weights_pack": (_i, [_p, C.POINTER(ModelDesc), C.POINTER(_p), _i, _p, _sz, _p]),
    "yoho_model_out_len": (_i, [_i]),
    "yoho_model_workspace_bytes": (_sz, [C.POINTER(ModelDesc), _i, _i, _i]),
    "yoho_model_encoder_fwd": (_i, [_p, C.POINTER(ModelDesc), _p, _p, _i, _i, _p, _i, _p, _sz, _p]),
    "yoho_launch_count": (C.c_longlong, []),
    "yoho_prof_enable": (_i, [_i]),
    "yoho_prof_num_classes": (_i, []),
    "yoho_prof_class_name": (C.c_char_p, [_i]),
    "yoho_prof_collect": (_i, [C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_longlong), _i]),
    "yoho_test_gemm_workspace_bytes": (_sz, [_i, _i, _i]),
    "yoho_test_gemm_bf16": (_i, [_p, _p, _p, _p, _p, _i, _i, _i, _i, _p, _sz, _p]),
    "yoho_test_attention_workspace_bytes": (_sz, [_i, _i, _i]),
    "yoho_test_skinny_gemm": (_i, [_p, _p, _p, _p, _p, _i, _i, _i, _i, _i, _p]),
    "yoho_test_attention_bf16": (_i, [_p, _p, _p, _i, _i, _i, _p, _sz, _p]),
    "yoho_test_attention_tiles": (_i, [_i]),
}

_lib = None


def load() -> C.CDLL:
    """Load (once) and type the shared library."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise ImportError(
            f"{LIB_PATH} is missing.  yoho_b200 has no CPU or PyTorch fallback: build the CUDA "
            "library first with `python -m yoho_b200.build` (needs nvcc, no GPU required).")
    lib = C.CDLL(str(LIB_PATH))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError here == header / library mismatch
        fn.restype = res
        fn.argtypes = args
    if lib.yoho_abi_version() != ABI_VERSION:
        raise ImportError(f"libyoho_b200 ABI {lib.yoho_abi_version()} != {ABI_VERSION}: rebuild with `python -m yoho_b200.build --force`")
    _lib = lib
    return lib


def check(code: int) -> int:
    """Raise YohoError for a negative status; pass non-negative results through."""
    if code < 0:
        raise YohoError(code, load().yoho_last_error().decode("utf-8", "replace"))
    return code
