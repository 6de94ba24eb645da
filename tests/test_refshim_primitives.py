"""The numpy stand-ins of ``oracle/refshim`` against a third party's implementation of the same primitive (torch, CPU).

``tests/test_reference_shim.py`` executes the reference's own source on these stand-ins; what that cannot pin is the INSIDE of
the library primitives the reference calls (flax.linen's Dense / Conv / LayerNorm / RMSNorm / Embed / gelu / silu / softmax /
max_pool, jax.scipy.signal.stft).  Here each stand-in is held to torch's implementation of the documented semantics, so a slip
in the restatement (a transposed kernel, a wrong epsilon, 'SAME' padding on the wrong side, a correlation written as a
convolution, the erf GELU instead of the tanh one) shows up independently of the oracle, which was written by the same hand.
What remains an assumption is the mapping itself, i.e. that flax 0.8.4 / jax 0.4.28 mean what their documentation says:
  Dense: y = x @ kernel + bias with kernel (in, out);  Conv: cross-correlation, kernel (width, in, out), channels last;
  LayerNorm: eps 1e-6, biased variance (fast variance E[x^2] - E[x]^2), scale and bias;  RMSNorm: eps 1e-6, scale only;
  nn.gelu: approximate=True (tanh);  max_pool: window over the dims in front of the feature axis, VALID.
CPU only, float64 (the stand-ins compute in the dtype of their input)."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle.refshim import flax_shim as fs
from oracle.refshim import jnp_shim as js

RNG = np.random.default_rng(20261105)


def _apply(module, params, *args):
    return np.asarray(module.apply({"params": params}, *args))


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a))


@pytest.mark.parametrize("use_bias", [True, False])
def test_dense(use_bias):
    x = RNG.standard_normal((3, 7, 20))
    params = {"kernel": RNG.standard_normal((20, 12))}
    if use_bias:
        params["bias"] = RNG.standard_normal(12)
    got = _apply(fs.Dense(features=12, use_bias=use_bias), params, x)
    want = F.linear(_t(x), _t(params["kernel"]).T, _t(params["bias"]) if use_bias else None).numpy()   # torch keeps (out, in)
    assert np.abs(got - want).max() <= 1e-12


def test_dense_general_splits_heads():
    x = RNG.standard_normal((2, 5, 16))
    params = {"kernel": RNG.standard_normal((16, 4, 8)), "bias": RNG.standard_normal((4, 8))}
    got = _apply(fs.DenseGeneral(features=(4, 8)), params, x)
    want = (F.linear(_t(x), _t(params["kernel"].reshape(16, 32)).T) + _t(params["bias"].reshape(32))).reshape(2, 5, 4, 8).numpy()
    assert got.shape == (2, 5, 4, 8) and np.abs(got - want).max() <= 1e-12


@pytest.mark.parametrize("length,stride,padding", [(50, 1, 1), (50, 2, 1), (51, 2, 1), (50, 1, "SAME"), (51, 2, "SAME"), (50, 2, "SAME"),
                                                   (50, 1, "VALID")])
def test_conv_is_a_cross_correlation_channels_last(length, stride, padding):
    """whisper.py:125,127 (padding=1, strides 1 and 2) and model.py:15-16 ('SAME' by default)."""
    x = RNG.standard_normal((2, length, 6))
    params = {"kernel": RNG.standard_normal((3, 6, 10)), "bias": RNG.standard_normal(10)}
    got = _apply(fs.Conv(features=10, kernel_size=(3,), strides=stride, padding=padding), params, x)
    xt = _t(x).transpose(1, 2)                               # torch: (batch, channels, length), weight (out, in, width)
    w = _t(params["kernel"]).permute(2, 1, 0)
    if padding == "SAME":                                    # XLA / lax 'SAME': out = ceil(L / s), the odd pad element goes to the end
        out_len = -(-length // stride)
        total = max((out_len - 1) * stride + 3 - length, 0)
        xt = F.pad(xt, (total // 2, total - total // 2))
        pad = 0
    else:
        pad = 0 if padding == "VALID" else padding
    want = F.conv1d(xt, w, _t(params["bias"]), stride=stride, padding=pad).transpose(1, 2).numpy()
    assert got.shape == want.shape and np.abs(got - want).max() <= 1e-12


def test_layer_norm_eps_and_biased_variance():
    x = RNG.standard_normal((4, 9, 32)) * 3.0 + 1.5
    params = {"scale": RNG.standard_normal(32), "bias": RNG.standard_normal(32)}
    got = _apply(fs.LayerNorm(), params, x)
    want = F.layer_norm(_t(x), (32,), _t(params["scale"]), _t(params["bias"]), eps=1e-6).numpy()
    assert np.abs(got - want).max() <= 1e-10
    # the epsilon is visible on a nearly constant row: 1e-5 (torch's default) instead of 1e-6 would change it by a factor ~3
    row = np.full((1, 32), 2.0)
    row[0, 0] += 1e-3
    a = _apply(fs.LayerNorm(), params, row)
    b = F.layer_norm(_t(row), (32,), _t(params["scale"]), _t(params["bias"]), eps=1e-5).numpy()
    assert np.abs(a - b).max() > 0.1


def test_rms_norm():
    x = RNG.standard_normal((3, 5, 24)) * 2.0
    params = {"scale": RNG.standard_normal(24)}
    got = _apply(fs.RMSNorm(), params, x)
    xt = _t(x)
    want = (xt * torch.rsqrt(xt.pow(2).mean(-1, keepdim=True) + 1e-6) * _t(params["scale"])).numpy()
    assert np.abs(got - want).max() <= 1e-12
    if hasattr(F, "rms_norm"):
        assert np.abs(got - F.rms_norm(xt, (24,), _t(params["scale"]), eps=1e-6).numpy()).max() <= 1e-12


def test_embed_and_attend_weight():
    ids = RNG.integers(0, 30, (2, 6))
    table = RNG.standard_normal((30, 8))
    emb = fs.Embed(num_embeddings=30, features=8)
    got = _apply(emb, {"embedding": table}, ids)
    assert np.array_equal(got, F.embedding(_t(ids), _t(table)).numpy())


def test_activations():
    x = np.linspace(-6.0, 6.0, 1001)
    assert np.abs(np.asarray(fs.gelu(x)) - F.gelu(_t(x), approximate="tanh").numpy()).max() <= 1e-12      # nn.gelu default: tanh
    assert np.abs(np.asarray(fs.gelu(x, approximate=False)) - F.gelu(_t(x)).numpy()).max() <= 1e-12
    assert np.abs(np.asarray(fs.gelu(x)) - F.gelu(_t(x)).numpy()).max() > 1e-4                            # the two differ visibly
    assert np.abs(np.asarray(fs.silu(x)) - F.silu(_t(x)).numpy()).max() <= 1e-12
    s = RNG.standard_normal((3, 4, 11)) * 5.0
    assert np.abs(np.asarray(fs.softmax(s)) - F.softmax(_t(s), dim=-1).numpy()).max() <= 1e-14
    s[0, 0, 3] = -np.inf                                                                                  # the causal mask of whisper.py:150
    assert np.abs(np.asarray(fs.softmax(s)) - F.softmax(_t(s), dim=-1).numpy()).max() <= 1e-14


def test_max_pool_of_a_rank3_input_pools_time_pairs():
    """model.py:20 calls nn.max_pool(x, (1, 2), strides=(1, 2)) on (B, S, D): the window dims are the dims in front of the
    feature axis, so with rank 3 the window (1, 2) lies over (batch, time): a maximum over time pairs, odd tail dropped."""
    x = RNG.standard_normal((3, 11, 5))
    got = np.asarray(fs.max_pool(x, (1, 2), strides=(1, 2)))
    want = F.max_pool1d(_t(x).transpose(1, 2), kernel_size=2, stride=2).transpose(1, 2).numpy()
    assert got.shape == (3, 5, 5) and np.array_equal(got, want)


@pytest.mark.parametrize("n", [1600, 16001])
def test_stft_stand_in_equals_torch_stft(n):
    """jax.scipy.signal.stft as audio.py:33 calls it (nperseg=n_fft, noverlap=n_fft-hop, boundary=None, padded=False; 'hann'
    periodic, rounded to float32 as JAX does with x64 off; scaling 'spectrum': divide by the window's sum)."""
    jnp = js.make_jax_modules()["jax.scipy.signal"]
    x = RNG.standard_normal((2, n))
    _, _, z = jnp.stft(x, nperseg=400, noverlap=240, boundary=None, padded=False)
    win = torch.hann_window(400, periodic=True, dtype=torch.float64).to(torch.float32).to(torch.float64)
    want = torch.stft(_t(x), n_fft=400, hop_length=160, window=win, center=False, onesided=True, return_complex=True) / win.sum()
    assert z.shape == tuple(want.shape)
    assert np.abs(np.asarray(z) - want.numpy()).max() <= 1e-12


def test_jax_dtype_rules_of_the_stand_in():
    """x64 disabled: arange is int32, int32 (x) float32 stays float32, jnp.log of a python scalar is float32 -- the rules the
    sinusoid table (whisper.py:11-17) and the RoPE angles (layers.py:31-33) are evaluated under."""
    jnp = js.make_jax_modules()["jax.numpy"]
    a = jnp.arange(10)
    assert a.dtype == np.int32
    f = (a * np.float32(0.5))
    assert np.asarray(f).dtype == np.float32
    assert isinstance(jnp.log(10000), np.float32)
    assert np.asarray(jnp.outer(a, jnp.arange(4) * np.float32(2.0))).dtype == np.float32
    # torch evaluates the same int32 x float32 product in float32 as well
    want = (torch.arange(10, dtype=torch.int32) * torch.tensor(0.5, dtype=torch.float32)).numpy()
    assert want.dtype == np.float32 and np.array_equal(np.asarray(f), want)
    # functional update
    b = jnp.arange(6).astype(np.float32).view(js.JArray)
    c = b.at[1::2].set(0.0)
    assert np.array_equal(np.asarray(c), [0, 0, 2, 0, 4, 0]) and np.array_equal(np.asarray(b), [0, 1, 2, 3, 4, 5])
