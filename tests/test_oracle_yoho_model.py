"""CPU checks of the oracle for YOHO's own audio encoder (oracle/yoho_model.py follows
yoho/src/nn/model.py:9-29 and yoho/src/nn/layers.py:7-116) and of the host-side constants the product builds."""
import numpy as np
import pytest

from oracle import yoho_model as ym


def test_grouped_query_attention_matches_the_reference_einops_patterns():
    """Re-evaluate layers.py:65-88 with the reference's literal einops patterns (einops is a reference dependency
    and is installed here) and compare with the oracle's explicit transposes."""
    from einops import einsum, rearrange
    rng = np.random.default_rng(0)
    B, S, D, qh, kvh = 2, 10, 256, 4, 2
    hd, groups = D // qh, qh // kvh
    p = ym.init_params(128, D, qh, kvh, 1, seed=1)["EncoderBlock_0"]["GroupedQueryAttention_0"]
    x = rng.standard_normal((B, S, D))
    q = np.einsum("bsi,ihd->bshd", x, p["DenseGeneral_0"]["kernel"].astype(np.float64))
    k = np.einsum("bsi,ihd->bshd", x, p["DenseGeneral_1"]["kernel"].astype(np.float64))
    v = np.einsum("bsi,ihd->bshd", x, p["DenseGeneral_2"]["kernel"].astype(np.float64))
    q = rearrange(q, "... s (h g) d -> ... g h s d", g=groups)          # layers.py:69
    k = rearrange(k, "... s h d -> ... h s d")
    v = rearrange(v, "... s h d -> ... h s d")
    thetas = (10.0 ** -np.arange(hd // 2, dtype=np.float32)).astype(np.float32)
    angles = np.outer(np.arange(1, S + 1).astype(np.float32), thetas).astype(np.float32).astype(np.float64)
    rot = np.exp(1j * angles)                                            # layers.py:33

    def rope(t):
        c = (t[..., ::2] + 1j * t[..., 1::2]) * rot                      # layers.py:35-36
        o = np.empty_like(t)
        o[..., ::2], o[..., 1::2] = c.real, c.imag
        return o
    q, k = rope(q), rope(k)
    scores = einsum(q, k, "... g h s d, ... h a d -> ... h s a")        # layers.py:77
    att = np.exp(scores / hd ** 0.5 - (scores / hd ** 0.5).max(-1, keepdims=True))
    att /= att.sum(-1, keepdims=True)
    out = einsum(att, v, "... h s a, ... h a d -> ... h s d")
    out = rearrange(out, "... h s d -> ... s (h d)") @ p["Dense_0"]["kernel"].astype(np.float64)
    got = ym.grouped_query_attention(x, p, qh, kvh)
    assert np.abs(got - out).max() < 1e-12


def test_query_group_sum_can_be_folded_into_the_weight():
    """What the CUDA path relies on: the scores are summed over g, RoPE is linear and head-independent,
    so attention only sees x @ sum_g Wq[:, h, g, :]."""
    rng = np.random.default_rng(1)
    B, S, D, qh, kvh = 1, 12, 256, 4, 2
    p = ym.init_params(128, D, qh, kvh, 1, seed=2)["EncoderBlock_0"]["GroupedQueryAttention_0"]
    x = rng.standard_normal((B, S, D))
    want = ym.grouped_query_attention(x, p, qh, kvh)
    wq = p["DenseGeneral_0"]["kernel"].astype(np.float64).reshape(D, kvh, qh // kvh, 64).sum(axis=2)
    folded = dict(p, DenseGeneral_0={"kernel": wq})
    got = ym.grouped_query_attention(x, folded, kvh, kvh)
    assert np.abs(got - want).max() < 1e-12


def test_max_pool_and_conv_shapes_for_the_reference_config():
    cfg = ym.YOHO_CONFIG
    p = ym.init_params(cfg["n_mel_bands"], cfg["dims"], cfg["n_audio_heads"], cfg["n_audio_heads"] // 2, 1, seed=0)
    mel = np.random.default_rng(0).standard_normal((1, 64, cfg["n_mel_bands"]))
    y, st = ym.audio_encoder(p, mel, cfg["n_audio_heads"], cfg["n_audio_heads"] // 2, return_stages=True)
    assert st["stem0"].shape == (1, 32, 384) and y.shape == (1, 32, 384)
    x = np.arange(2 * 6 * 3, dtype=np.float64).reshape(2, 6, 3)
    assert np.array_equal(ym.max_pool_time2(x), x[:, 1::2])            # increasing in time -> the odd rows win
    assert ym.max_pool_time2(np.zeros((1, 7, 2))).shape == (1, 3, 2)   # VALID: the odd tail row is dropped


def test_conv_same_padding_is_a_cross_correlation():
    p = {"kernel": np.zeros((3, 1, 1)), "bias": np.zeros(1)}
    p["kernel"][0, 0, 0] = 1.0                                        # tap 0 reads x[t-1]
    x = np.arange(1, 6, dtype=np.float64).reshape(1, 5, 1)
    assert np.array_equal(ym.conv1d_k3_same(x, p)[0, :, 0], [0, 1, 2, 3, 4])


def test_rms_norm_and_rope_constants():
    x = np.random.default_rng(3).standard_normal((4, 8))
    y = ym.rms_norm(x, {"scale": np.ones(8)})
    assert np.allclose(np.mean(y * y, -1), 1.0, atol=1e-5)
    from yoho_b200.nn.model import rope_table
    t = rope_table(512, 64)
    c, s = ym.rope_tables(512, 64)
    assert t.shape == (512, 2, 32) and t.dtype == np.float32
    assert np.abs(t[:, 0] - c).max() < 1e-7 and np.abs(t[:, 1] - s).max() < 1e-7
    assert abs(t[0, 0, 0] - np.cos(1.0)) < 1e-7 and abs(t[0, 1, 1] - np.sin(0.1)) < 1e-7   # positions start at 1


def test_product_and_oracle_share_the_parameter_generator():
    from yoho_b200.nn.model import flatten_yoho_encoder_params, init_yoho_encoder_params, rope_table
    a = ym.init_params(128, 128, 2, 1, 2, seed=5)
    b = init_yoho_encoder_params(128, 128, 2, 1, 2, seed=5)
    fa, fb = flatten_yoho_encoder_params(a, 2, rope_table(8)), flatten_yoho_encoder_params(b, 2, rope_table(8))
    assert len(fa) == 7 + 9 * 2 and all(np.array_equal(u, v) for u, v in zip(fa, fb))
    with pytest.raises(KeyError):
        flatten_yoho_encoder_params({"Conv_0": {}}, 0, rope_table(8))
