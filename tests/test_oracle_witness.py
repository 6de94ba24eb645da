"""Cross-check of the encoder oracle against an independent implementation available offline:
``transformers`` WhisperEncoder sub-modules on torch-CPU float64, driven by hand because
``WhisperEncoder.forward`` hard-codes erf-GELU in the stem and rejects inputs that are not 3000 frames.
Mapping: Flax Dense.kernel -> Linear.weight = kernel.T; Conv.kernel (k,in,out) -> Conv1d.weight (out,in,k);
every LayerNorm.eps = 1e-6; activation gelu_pytorch_tanh; sinusoid table added by hand."""
import warnings

import numpy as np
import pytest

from oracle import encoder as enc


def test_encoder_oracle_matches_hf_whisper_submodules():
    torch = pytest.importorskip("torch")
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        tr = pytest.importorskip("transformers")
        from transformers.models.whisper.modeling_whisper import WhisperEncoder
    M, D, H, L, F = 80, 128, 2, 2, 199
    T = enc.conv_out_len(F)
    p = enc.init_params(M, D, L, seed=42)
    mel = np.random.default_rng(0).standard_normal((2, F, M)).astype(np.float32)
    want = enc.audio_encoder(p, mel, H, np.float64)

    cfg = tr.WhisperConfig(num_mel_bins=M, d_model=D, encoder_layers=L, encoder_attention_heads=H, encoder_ffn_dim=4 * D,
                           max_source_positions=T, activation_function="gelu_pytorch_tanh")
    hf = WhisperEncoder(cfg).double().eval()
    t = lambda a: torch.from_numpy(np.asarray(a, np.float64))
    with torch.no_grad():
        hf.conv1.weight.copy_(t(p["Conv_0"]["kernel"]).permute(2, 1, 0)); hf.conv1.bias.copy_(t(p["Conv_0"]["bias"]))
        hf.conv2.weight.copy_(t(p["Conv_1"]["kernel"]).permute(2, 1, 0)); hf.conv2.bias.copy_(t(p["Conv_1"]["bias"]))
        for i, layer in enumerate(hf.layers):
            b = p[f"EncoderBlock_{i}"]; a = b["MultiHeadAttention_0"]
            for ln, src in ((layer.self_attn_layer_norm, b["LayerNorm_0"]), (layer.final_layer_norm, b["LayerNorm_1"])):
                ln.weight.copy_(t(src["scale"])); ln.bias.copy_(t(src["bias"])); ln.eps = 1e-6
            layer.self_attn.q_proj.weight.copy_(t(a["query_proj"]["kernel"]).T); layer.self_attn.q_proj.bias.copy_(t(a["query_proj"]["bias"]))
            layer.self_attn.k_proj.weight.copy_(t(a["key_proj"]["kernel"]).T)
            layer.self_attn.v_proj.weight.copy_(t(a["value_proj"]["kernel"]).T); layer.self_attn.v_proj.bias.copy_(t(a["value_proj"]["bias"]))
            layer.self_attn.out_proj.weight.copy_(t(a["out_proj"]["kernel"]).T); layer.self_attn.out_proj.bias.copy_(t(a["out_proj"]["bias"]))
            layer.fc1.weight.copy_(t(b["Dense_0"]["kernel"]).T); layer.fc1.bias.copy_(t(b["Dense_0"]["bias"]))
            layer.fc2.weight.copy_(t(b["Dense_1"]["kernel"]).T); layer.fc2.bias.copy_(t(b["Dense_1"]["bias"]))
        hf.layer_norm.weight.copy_(t(p["LayerNorm_0"]["scale"])); hf.layer_norm.bias.copy_(t(p["LayerNorm_0"]["bias"])); hf.layer_norm.eps = 1e-6
        gelu = lambda v: torch.nn.functional.gelu(v, approximate="tanh")
        x = t(mel).transpose(1, 2)
        x = gelu(hf.conv2(gelu(hf.conv1(x)))).transpose(1, 2) + t(enc.sinusoid_table(T, D))
        for layer in hf.layers:
            out = layer(x, None)
            x = out[0] if isinstance(out, tuple) else out
        got = hf.layer_norm(x).numpy()
    assert np.abs(got - want).max() < 1e-10


def test_decoder_oracle_matches_hf_whisper_decoder():
    """oracle.decoder.text_decoder against transformers' WhisperDecoder (float64, torch CPU) with the same weights:
    learned positions, tied embedding, masked self attention, cross attention over the encoder states, tanh GELU."""
    torch = pytest.importorskip("torch")
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        tr = pytest.importorskip("transformers")
        from transformers.models.whisper.modeling_whisper import WhisperDecoder
    from oracle import decoder as dec
    V, L_MAX, D, H, L, T, S, B = 97, 24, 128, 2, 2, 50, 11, 2
    p = dec.init_params(V, L_MAX, D, L, seed=7)
    rng = np.random.default_rng(1)
    tokens = rng.integers(0, V, (B, S))
    states = rng.standard_normal((B, T, D)).astype(np.float32)
    want = dec.text_decoder(p, tokens, states, H, np.float64)

    cfg = tr.WhisperConfig(vocab_size=V, d_model=D, decoder_layers=L, decoder_attention_heads=H, decoder_ffn_dim=4 * D,
                           max_target_positions=L_MAX, activation_function="gelu_pytorch_tanh", pad_token_id=0,
                           scale_embedding=False, dropout=0.0, attention_dropout=0.0, activation_dropout=0.0)
    cfg._attn_implementation = "eager"
    hf = WhisperDecoder(cfg).double().eval()
    t = lambda a: torch.from_numpy(np.asarray(a, np.float64))

    def put_attn(mod, a):
        mod.q_proj.weight.copy_(t(a["query_proj"]["kernel"]).T); mod.q_proj.bias.copy_(t(a["query_proj"]["bias"]))
        mod.k_proj.weight.copy_(t(a["key_proj"]["kernel"]).T)
        mod.v_proj.weight.copy_(t(a["value_proj"]["kernel"]).T); mod.v_proj.bias.copy_(t(a["value_proj"]["bias"]))
        mod.out_proj.weight.copy_(t(a["out_proj"]["kernel"]).T); mod.out_proj.bias.copy_(t(a["out_proj"]["bias"]))

    with torch.no_grad():
        hf.embed_tokens.weight.copy_(t(p["Embed_0"]["embedding"]))
        hf.embed_positions.weight.copy_(t(p["positional_embedding"]))
        for i, layer in enumerate(hf.layers):
            b = p[f"DecoderBlock_{i}"]
            for ln, src in ((layer.self_attn_layer_norm, b["LayerNorm_0"]), (layer.encoder_attn_layer_norm, b["LayerNorm_1"]),
                            (layer.final_layer_norm, b["LayerNorm_2"])):
                ln.weight.copy_(t(src["scale"])); ln.bias.copy_(t(src["bias"])); ln.eps = 1e-6
            put_attn(layer.self_attn, b["MultiHeadAttention_0"])
            put_attn(layer.encoder_attn, b["MultiHeadAttention_1"])
            layer.fc1.weight.copy_(t(b["Dense_0"]["kernel"]).T); layer.fc1.bias.copy_(t(b["Dense_0"]["bias"]))
            layer.fc2.weight.copy_(t(b["Dense_1"]["kernel"]).T); layer.fc2.bias.copy_(t(b["Dense_1"]["bias"]))
        hf.layer_norm.weight.copy_(t(p["LayerNorm_0"]["scale"])); hf.layer_norm.bias.copy_(t(p["LayerNorm_0"]["bias"])); hf.layer_norm.eps = 1e-6
        out = hf(input_ids=torch.from_numpy(tokens), encoder_hidden_states=t(states), use_cache=False)
        hidden = out.last_hidden_state
        got = (hidden @ hf.embed_tokens.weight.T).numpy()
    assert np.abs(got - want).max() < 1e-9 * max(1.0, np.abs(want).max())
    # the reference's greedy loop re-runs the full decoder per token; causality makes the unwritten tail irrelevant
    forced = np.zeros(L_MAX, bool); forced[0] = forced[2] = True
    buf = np.zeros((B, L_MAX), np.int64); buf[:, 0] = 3; buf[:, 2] = 5
    seq = dec.greedy_decode(p, states, H, buf, forced, 6)
    full = dec.text_decoder(p, seq[:, :7], states, H)
    for pos in (1, 3, 4, 5, 6):
        assert np.array_equal(seq[:, pos], full[:, pos - 1].argmax(-1))
    assert np.array_equal(seq[:, 2], buf[:, 2])
