"""Execute the reference's OWN hot-path source files on numpy (TEST INFRASTRUCTURE ONLY).

jax, jaxlib and flax are not installed on the build or GPU image (no network), so the reference cannot run as shipped and the
oracle under ``oracle/`` was "unpinned by the reference" beyond ``mel_filter_banks``.  This package narrows that: it installs
small numpy stand-ins for exactly the ``jax`` / ``jax.numpy`` / ``jax.scipy.signal`` / ``flax.linen`` / ``ffmpeg`` names those
files import (``jnp_shim.py``, ``flax_shim.py``), after which

    yoho/src/preprocessing/audio.py      mel_spectogram, get_batched_spectogram, normalize_spectogram
    yoho/src/nn/whisper.py               SinPositionalEncoding, MultiHeadAttention, EncoderBlock, AudioEncoder, Whisper
    yoho/src/nn/model.py, layers.py      YOHO's own AudioEncoder, EncoderBlock, GroupedQueryAttention, RoPE, SwiGLU

are imported from ``/root/reference`` UNMODIFIED and executed.  What that pins is everything the reference's authors wrote:
the order and composition of operations, every scale factor and transpose, which projection has a bias, the einsum / einops
patterns, the parameter pytree's names (they come out of the reference's module structure through Flax's auto-naming rules,
restated in the shim).  What it cannot pin is the inside of the library primitives (stft, Dense, Conv, LayerNorm, gelu,
softmax, RMSNorm, max_pool): those are the shim's restatement of the documented jax 0.4.28 / flax 0.8.4 / scipy behaviour
(stft is SciPy's own); ``tests/test_refshim_primitives.py`` holds each of them to torch's implementation of the same operation.
``tests/golden/make_reference_golden.py`` runs the reference this way and commits its outputs as
``tests/golden/reference_shim_golden.npz``; ``tests/test_reference_shim.py`` checks the oracle (and, on the GPU, the CUDA
path) against them.  ``/root/reference`` does not exist on the GPU box: only the committed vectors travel.
"""
from __future__ import annotations

import sys
import types
from pathlib import Path

REFERENCE_ROOT = Path("/root/reference")


def reference_available() -> bool:
    return (REFERENCE_ROOT / "yoho" / "src" / "nn" / "whisper.py").exists()


def install() -> None:
    """Put the stand-in modules into ``sys.modules`` and the reference checkout on ``sys.path`` (idempotent).  Refuses to
    shadow a real jax / flax installation."""
    if getattr(install, "_done", False):
        return
    for real in ("jax", "flax"):
        if real in sys.modules and not getattr(sys.modules[real], "_yoho_refshim", False):
            raise RuntimeError(f"a real {real} is already imported; the shim is only for images without it")
    from .flax_shim import make_flax_modules
    from .jnp_shim import make_jax_modules
    mods = {**make_jax_modules(), **make_flax_modules()}
    ffmpeg = types.ModuleType("ffmpeg")             # audio.py imports it for load_audio / save_audio (out of scope, never called)
    mods["ffmpeg"] = ffmpeg
    for name, mod in mods.items():
        mod._yoho_refshim = True
        sys.modules[name] = mod
    if str(REFERENCE_ROOT) not in sys.path:
        sys.path.insert(0, str(REFERENCE_ROOT))
    install._done = True


def load_reference():
    """-> (audio module, whisper module, model module, layers module) of the reference, imported under the shim."""
    if not reference_available():
        raise FileNotFoundError(f"{REFERENCE_ROOT} is not mounted")
    install()
    import importlib
    return tuple(importlib.import_module(m) for m in ("yoho.src.preprocessing.audio", "yoho.src.nn.whisper", "yoho.src.nn.model", "yoho.src.nn.layers"))
