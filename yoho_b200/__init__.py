"""yoho_b200 -- B200-native (sm_100a) implementation of YOHO's data-parallel hot path:
the log-mel frontend and the Whisper audio-encoder forward, behind the reference's own call
signatures.  All computation happens in ``lib/libyoho_b200.so`` (hand-written CUDA, C ABI in
``include/yoho_b200.h``); build it with ``python -m yoho_b200.build``.  No CPU fallback.
"""
__version__ = "0.1.0"

MODEL_SHAPES = {
    # name: (n_mels, dims, n_heads, n_layers); seq_len is 1500 for 30 s clips
    "tiny": (80, 384, 6, 4),
    "base": (80, 512, 8, 6),
    "small": (80, 768, 12, 12),
    "large-v3": (128, 1280, 20, 32),
}
