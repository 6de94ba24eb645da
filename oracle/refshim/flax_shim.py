"""Minimal stand-in for ``flax.linen`` (flax 0.8.4 semantics, as pinned by the reference's poetry.lock), enough to EXECUTE
the reference's own module code -- ``yoho/src/nn/whisper.py`` and ``yoho/src/nn/model.py`` / ``layers.py`` -- unmodified on
numpy arrays (TEST INFRASTRUCTURE ONLY; see oracle/refshim/__init__.py).

What it reproduces of Flax:
* modules are dataclasses built from their class annotations; ``name`` / ``parent`` are keyword-only extras;
* ``@nn.compact``: sub-modules created inside the method are auto-named ``<ClassName>_<i>``, i counting that class within
  the parent (``Conv_0, Conv_1, EncoderBlock_0 .. , LayerNorm_0``) -- these names ARE the parameter pytree's keys;
* ``setup()``: runs lazily before the first method call; a sub-module assigned to an attribute takes the attribute's name
  (``query_proj``, ``audio_encoder``);
* ``module.apply({"params": tree}, *args, method=Cls.fn)``: binds the tree to the top module; a child's parameters are
  ``parent_params[child.name]``, resolved when ``self.param`` is first used (parameter-free modules need no entry);
* layers, with Flax's defaults: ``Dense`` (kernel (in, out), ``x @ kernel + bias``), ``DenseGeneral``, ``Conv`` (kernel
  (k, in, out), cross-correlation over axis -2, int / 'SAME' padding, strides), ``LayerNorm`` (eps 1e-6, variance as
  E[x^2] - E[x]^2 clipped at 0, scale + bias), ``RMSNorm`` (eps 1e-6), ``gelu`` (tanh approximation), ``silu``,
  ``softmax`` (last axis), ``max_pool`` (VALID; the window's dims are the dims in front of the feature axis).
"""
from __future__ import annotations

import dataclasses
import functools
import math
import types
from typing import Any, Optional

import numpy as np

from .jnp_shim import JArray

_stack: list["Module"] = []          # modules whose setup() or method is executing (innermost last)


def _j(x):
    return x.view(JArray) if isinstance(x, np.ndarray) and not isinstance(x, JArray) else x


def _wrap_method(fn):
    @functools.wraps(fn)
    def wrapper(self, *args, **kwargs):
        self._ensure_setup()
        if fn.__name__ == "__call__" or getattr(fn, "_compact", False):
            object.__setattr__(self, "_counts", {})
        _stack.append(self)
        try:
            return fn(self, *args, **kwargs)
        finally:
            _stack.pop()
    return wrapper


@dataclasses.dataclass(eq=False, repr=False)
class Module:
    name: Optional[str] = dataclasses.field(default=None, kw_only=True)
    parent: Any = dataclasses.field(default=None, kw_only=True)

    def __init_subclass__(cls, **kw):
        super().__init_subclass__(**kw)
        for key, val in list(cls.__dict__.items()):
            if isinstance(val, types.FunctionType) and key not in ("setup", "__post_init__") and (not key.startswith("_") or key == "__call__"):
                setattr(cls, key, _wrap_method(val))
        dataclasses.dataclass(cls, eq=False, repr=False)

    def __post_init__(self):
        object.__setattr__(self, "_bound", None)
        object.__setattr__(self, "_setup_done", False)
        object.__setattr__(self, "_in_setup", False)
        object.__setattr__(self, "_counts", {})
        if _stack and self.parent is None:
            parent = _stack[-1]
            object.__setattr__(self, "parent", parent)
            if not parent._in_setup and self.name is None:       # inside a compact method: auto-name
                i = parent._counts.get(type(self).__name__, 0)
                parent._counts[type(self).__name__] = i + 1
                object.__setattr__(self, "name", f"{type(self).__name__}_{i}")

    def __setattr__(self, key, value):
        if isinstance(value, Module) and getattr(self, "_in_setup", False) and value.name is None:
            object.__setattr__(value, "name", key)                # setup(): the attribute name is the module's name
            object.__setattr__(value, "parent", self)
        object.__setattr__(self, key, value)

    def _ensure_setup(self):
        if self._setup_done:
            return
        object.__setattr__(self, "_setup_done", True)
        setup = getattr(self, "setup", None)
        if setup is not None:
            object.__setattr__(self, "_in_setup", True)
            _stack.append(self)
            try:
                setup()
            finally:
                _stack.pop()
                object.__setattr__(self, "_in_setup", False)

    def _params(self) -> dict:
        if self._bound is not None:
            return self._bound
        if self.parent is None:
            raise ValueError(f"{type(self).__name__}: no parameters bound (call through .apply)")
        tree = self.parent._params()
        if self.name not in tree:
            raise KeyError(f"parameter subtree {self.name!r} missing under {type(self.parent).__name__} (have {sorted(tree)})")
        return tree[self.name]

    def param(self, name, init_fn=None, *shape_args):
        tree = self._params()
        if name not in tree:
            raise KeyError(f"parameter {name!r} missing in {type(self).__name__} {self.name!r} (have {sorted(tree)})")
        return _j(np.asarray(tree[name]))

    def apply(self, variables, *args, method=None, **kwargs):
        object.__setattr__(self, "_bound", variables["params"] if "params" in variables else variables)
        try:
            if method is None:
                return self(*args, **kwargs)
            return method(self, *args, **kwargs)
        finally:
            object.__setattr__(self, "_bound", None)


def compact(fn):
    fn._compact = True
    return fn


# ---------------------------------------------------------------------------------------------------------- layers
class Dense(Module):
    features: int
    use_bias: bool = True

    def __call__(self, x):
        kernel = self.param("kernel")
        y = x @ kernel.astype(x.dtype)
        if self.use_bias:
            y = y + self.param("bias").astype(x.dtype)
        return _j(y)


class DenseGeneral(Module):
    features: Any
    use_bias: bool = True
    axis: int = -1

    def __call__(self, x):
        kernel = self.param("kernel")                                     # (in, *features)
        y = np.tensordot(np.asarray(x), np.asarray(kernel).astype(x.dtype), axes=1)
        if self.use_bias:
            y = y + np.asarray(self.param("bias")).astype(x.dtype)
        return _j(y)


class Conv(Module):
    features: int
    kernel_size: Any
    strides: Any = 1
    padding: Any = "SAME"
    dtype: Any = None
    use_bias: bool = True

    def __call__(self, x):
        k = self.kernel_size if isinstance(self.kernel_size, int) else self.kernel_size[0]
        stride = self.strides if isinstance(self.strides, int) else (self.strides[0] if self.strides else 1)
        length = x.shape[-2]
        if isinstance(self.padding, int):
            lo = hi = self.padding
        elif self.padding == "SAME":
            out_len = -(-length // stride)
            total = max((out_len - 1) * stride + k - length, 0)
            lo, hi = total // 2, total - total // 2
        elif self.padding == "VALID":
            lo = hi = 0
        else:
            lo, hi = self.padding[0]
        kernel = np.asarray(self.param("kernel")).astype(x.dtype)          # (k, in, out): cross-correlation
        xp = np.pad(np.asarray(x), [(0, 0)] * (x.ndim - 2) + [(lo, hi), (0, 0)])
        out_len = (length + lo + hi - k) // stride + 1
        y = np.zeros(x.shape[:-2] + (out_len, self.features), x.dtype)
        for tap in range(k):
            y += xp[..., tap: tap + stride * (out_len - 1) + 1: stride, :] @ kernel[tap]
        if self.use_bias:
            y = y + np.asarray(self.param("bias")).astype(x.dtype)
        return _j(y)


class LayerNorm(Module):
    epsilon: float = 1e-6

    def __call__(self, x):
        x = np.asarray(x)
        mean = x.mean(axis=-1, keepdims=True)
        mean2 = (x * x).mean(axis=-1, keepdims=True)
        var = np.maximum(0.0, mean2 - mean * mean)                         # use_fast_variance=True
        y = (x - mean) / np.sqrt(var + x.dtype.type(self.epsilon))
        return _j(y * np.asarray(self.param("scale")).astype(x.dtype) + np.asarray(self.param("bias")).astype(x.dtype))


class RMSNorm(Module):
    epsilon: float = 1e-6

    def __call__(self, x):
        x = np.asarray(x)
        ms = np.mean(np.square(x), axis=-1, keepdims=True)
        return _j(x / np.sqrt(ms + x.dtype.type(self.epsilon)) * np.asarray(self.param("scale")).astype(x.dtype))


class Embed(Module):
    num_embeddings: int
    features: int

    def setup(self):
        pass

    @property
    def embedding(self):
        return self.param("embedding")

    def __call__(self, ids):
        return _j(np.asarray(self.param("embedding"))[np.asarray(ids)])


def gelu(x, approximate: bool = True):
    x = np.asarray(x)
    if approximate:
        c = math.sqrt(2.0 / math.pi)
        return _j(0.5 * x * (1.0 + np.tanh(c * (x + 0.044715 * (x * x * x)))))
    from scipy.special import erf
    return _j(0.5 * x * (1.0 + erf(x / math.sqrt(2.0))))


def silu(x):
    x = np.asarray(x)
    return _j(x / (1.0 + np.exp(-x)))


def softmax(x, axis: int = -1):
    x = np.asarray(x)
    e = np.exp(x - x.max(axis=axis, keepdims=True))
    return _j(e / e.sum(axis=axis, keepdims=True))


def max_pool(inputs, window_shape, strides=None, padding="VALID"):
    """flax.linen.max_pool: the window dims are the dims in front of the trailing feature axis; dims before them are batch
    dims.  For a rank-3 (B, S, D) input and a two-element window -- the reference's call, model.py:20 -- there is no batch dim
    left, so the window (1, 2) runs over (B, S): a max over time pairs."""
    x = np.asarray(inputs)
    assert padding == "VALID"
    strides = tuple(strides) if strides is not None else (1,) * len(window_shape)
    nb = x.ndim - (len(window_shape) + 1)
    out = x
    for i, (w, s) in enumerate(zip(window_shape, strides)):
        ax = nb + i
        n_out = (out.shape[ax] - w) // s + 1
        parts = [np.take(out, np.arange(j, j + s * (n_out - 1) + 1, s), axis=ax) for j in range(w)]
        out = functools.reduce(np.maximum, parts)
    return _j(out)


def make_flax_modules() -> dict[str, types.ModuleType]:
    linen = types.ModuleType("flax.linen")
    for name, obj in dict(Module=Module, compact=compact, Dense=Dense, DenseGeneral=DenseGeneral, Conv=Conv, LayerNorm=LayerNorm,
                          RMSNorm=RMSNorm, Embed=Embed, gelu=gelu, silu=silu, softmax=softmax, max_pool=max_pool).items():
        setattr(linen, name, obj)
    initializers = types.ModuleType("flax.linen.initializers")
    initializers.glorot_uniform = lambda *a, **k: None
    linen.initializers = initializers
    flax = types.ModuleType("flax")
    flax.linen = linen
    return {"flax": flax, "flax.linen": linen, "flax.linen.initializers": initializers}
