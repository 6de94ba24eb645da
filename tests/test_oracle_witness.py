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
