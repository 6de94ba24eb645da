"""numpy stand-in for the few ``jax`` / ``jax.numpy`` / ``jax.scipy`` names the reference's hot path uses
(TEST INFRASTRUCTURE ONLY; see oracle/refshim/__init__.py).

What JAX does that plain numpy does not, and the reference relies on:

* x64 is disabled: ``jnp.arange(n)`` is int32, ``jnp.log(10000)`` is float32, and int32 (x) float32 -> float32 (numpy would
  promote to float64).  ``JArray`` is an ndarray subclass whose ufunc hook casts integer operands to the widest floating
  operand's dtype, which reproduces that for the expressions in ``SinPositionalEncoding`` (whisper.py:11-17) and ``RoPE``
  (layers.py:31-33).  Floating inputs keep their dtype, so feeding float64 data gives a float64 run of the reference's code
  (the tight comparison with the float64 oracle) and float32 data a dtype-faithful one.
* ``x.at[idx].set(v)`` (layers.py:39-41): functional update, returns a copy.
* ``jax.scipy.signal.stft`` mirrors ``scipy.signal.stft`` (same algorithm and defaults: JAX's docstring says so and
  tests/test_oracle_witness.py pins the oracle to it): delegated to SciPy.
"""
from __future__ import annotations

import types

import numpy as np
import scipy.signal


class _At:
    def __init__(self, arr):
        self._arr = arr

    def __getitem__(self, idx):
        arr = self._arr

        class _Ref:
            def set(self, value):
                out = np.array(arr, copy=True).view(JArray)
                out[idx] = value
                return out
        return _Ref()


class JArray(np.ndarray):
    """ndarray with JAX's int (x) float32 -> float32 promotion and the ``.at[...]`` functional update."""

    def __array_ufunc__(self, ufunc, method, *inputs, out=None, **kwargs):
        raw = [np.asarray(x) if isinstance(x, np.ndarray) else x for x in inputs]
        floats = [x.dtype for x in raw if isinstance(x, np.ndarray) and x.dtype.kind in "fc"]
        floats += [np.asarray(x).dtype for x in raw if isinstance(x, np.generic) and np.asarray(x).dtype.kind in "fc"]
        if floats:
            target = max(floats, key=lambda d: d.itemsize)
            if target.kind == "c":
                target = np.dtype(np.float64 if target.itemsize == 16 else np.float32)
            raw = [x.astype(target) if isinstance(x, np.ndarray) and x.dtype.kind in "iu" else x for x in raw]
        if out is not None:
            kwargs["out"] = tuple(np.asarray(o) if isinstance(o, np.ndarray) else o for o in out)
        res = getattr(ufunc, method)(*raw, **kwargs)
        if isinstance(res, tuple):
            return tuple(r.view(JArray) if isinstance(r, np.ndarray) else r for r in res)
        return res.view(JArray) if isinstance(res, np.ndarray) else res

    @property
    def at(self):
        return _At(self)


def _wrap(x):
    return x.view(JArray) if isinstance(x, np.ndarray) and not isinstance(x, JArray) else x


def _arange(*args, dtype=None, **kw):
    out = np.arange(*args, dtype=dtype, **kw)
    if dtype is None and out.dtype.kind == "i":
        out = out.astype(np.int32)                  # x64 disabled
    elif dtype is None and out.dtype == np.float64:
        out = out.astype(np.float32)
    return out.view(JArray)


def _log(x):
    if isinstance(x, (int, float)):                 # python scalar -> float32 (x64 disabled)
        return np.float32(np.log(np.float32(x)))
    return _wrap(np.log(x))


def _outer(a, b):
    a, b = np.asarray(a), np.asarray(b)
    if a.dtype.kind in "iu" and b.dtype.kind == "f":
        a = a.astype(b.dtype)
    return np.outer(a, b).view(JArray)


def _empty_like(x):
    return np.empty_like(np.asarray(x)).view(JArray)


def make_jax_modules() -> dict[str, types.ModuleType]:
    jnp = types.ModuleType("jax.numpy")

    def _getattr(name):                              # everything else is numpy's (abs, einsum, log10, maximum, mean, std, where, ...)
        return getattr(np, name)
    jnp.__getattr__ = _getattr
    jnp.arange, jnp.log, jnp.outer, jnp.empty_like = _arange, _log, _outer, _empty_like
    jnp.ndarray = np.ndarray
    jnp.e, jnp.inf, jnp.float32 = float(np.e), float("inf"), np.float32

    def _stft(x, fs=1.0, window="hann", nperseg=256, noverlap=None, **kw):
        # jax.scipy.signal.stft builds the window with SciPy in float64 and then casts it to the data's dtype, which is
        # float32 with x64 disabled: the window the reference multiplies by is the float32-ROUNDED periodic Hann
        win = scipy.signal.get_window(window, nperseg).astype(np.float32).astype(np.asarray(x).dtype)
        return scipy.signal.stft(x, fs=fs, window=win, nperseg=nperseg, noverlap=noverlap, **kw)

    signal = types.ModuleType("jax.scipy.signal")
    signal.stft = _stft
    jscipy = types.ModuleType("jax.scipy")
    jscipy.signal = signal
    jax = types.ModuleType("jax")
    jax.numpy, jax.scipy = jnp, jscipy
    return {"jax": jax, "jax.numpy": jnp, "jax.scipy": jscipy, "jax.scipy.signal": signal}
