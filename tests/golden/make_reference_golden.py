"""Generate tests/golden/reference_shim_golden.npz by EXECUTING THE REFERENCE'S OWN SOURCE FILES (imported unmodified from
/root/reference) on the numpy stand-ins of oracle/refshim -- see that package's docstring for what this does and does not pin.

    python tests/golden/make_reference_golden.py          # needs /root/reference (this container; not the GPU box)

Inputs are the ones the other golden files already hold (frontend_golden.npz's audio, encoder_golden.npz's mel and seed), so
every consumer -- the oracle, the CUDA path -- can be compared on identical data.  All reference runs are float64 runs of the
reference's code (float64 inputs; the window, the sinusoid table and the RoPE angles keep the float32 roundings JAX applies
with x64 disabled)."""
from __future__ import annotations

import hashlib
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))

from oracle import decoder as od  # noqa: E402
from oracle import encoder as eo  # noqa: E402
from oracle import refshim  # noqa: E402
from oracle import yoho_model as ym  # noqa: E402

GOLDEN = ROOT / "tests" / "golden"
YOHO_SHAPE = dict(n_mels=64, dims=256, q_heads=4, kv_heads=2, blocks=2, frames=50, seed=7)   # head dim 64: what the CUDA path supports
TABLE_SHAPES = [(1500, 384), (1500, 512), (1500, 768), (1500, 1280)]
DECODER_SHAPE = dict(vocab=97, text_seq_len=24, dims=128, heads=2, layers=2, audio_frames=50, batch=3, tokens=11, seed=3)


def as_float64(tree):
    """Token ids carry no floating dtype, so a float64 run of the reference's decoder needs float64 leaves (float32 leaves
    make it a float32 run: measured 7e-6 from the float64 oracle, pure rounding)."""
    return {k: as_float64(v) for k, v in tree.items()} if isinstance(tree, dict) else np.asarray(tree, np.float64)


def run_reference() -> dict:
    audio, whisper, model, _layers = refshim.load_reference()
    out: dict = {}
    # ---- frontend: audio.py:32-37 (mel_spectogram), :40-50 (get_batched_spectogram), :53-58 (normalize_spectogram)
    fg = np.load(GOLDEN / "frontend_golden.npz")
    pcm = fg["audio"][:2].astype(np.float64)
    out["pcm_clips"] = np.array([0, 1])
    for m in (80, 128):
        power = np.asarray(audio.mel_spectogram(pcm, 400, 160, 16000, m))
        out[f"m{m}_power"] = power
        out[f"m{m}_reference"] = np.asarray(audio.normalize_spectogram(power))
    from yoho.src.config import YOHOConfig
    cfg = YOHOConfig(sample_rate=16000, n_mel_bands=128, n_fft=400, stft_hop=160, max_audio_len=1024, max_text_len=256,
                     dims=384, n_audio_heads=6, n_audio_blocks=4, n_text_heads=6, n_text_blocks=4)
    assert np.array_equal(np.asarray(audio.get_batched_spectogram(cfg)(pcm)), out["m128_power"])
    # ---- encoder: whisper.py:117-135 through Whisper.encode_audio (:194-195) and model.apply(..., method=...) (:258)
    eg = np.load(GOLDEN / "encoder_golden.npz")
    n_mels, dims, heads, layers, seed = (int(eg[k]) for k in ("n_mels", "dims", "n_heads", "n_layers", "seed"))
    params = eo.init_params(n_mels, dims, layers, seed=seed)
    mel = eg["mel"].astype(np.float64)
    seq_len = eo.conv_out_len(mel.shape[1])
    net = whisper.Whisper(audio_dims=dims, audio_seq_len=seq_len, audio_heads=heads, audio_layers=layers, text_vocab_size=16,
                          text_seq_len=8, text_dims=dims, text_heads=heads, text_layers=1)
    out["encoder_out"] = np.asarray(net.apply({"params": {"audio_encoder": params}}, mel, method=whisper.Whisper.encode_audio))
    bare = whisper.AudioEncoder(seq_len=seq_len, dims=dims, n_heads=heads, n_layers=layers)
    assert np.array_equal(np.asarray(bare.apply({"params": params}, mel)), out["encoder_out"])
    # ---- sinusoid tables (whisper.py:11-17), float32 as JAX evaluates them: kept as digests (the full tables are MBs)
    for (length, dim) in TABLE_SHAPES:
        table = np.asarray(whisper.SinPositionalEncoding(length, dim).sin_positional_encoding())
        assert table.dtype == np.float32 and table.shape == (length, dim)
        out[f"table_sha256_{length}x{dim}"] = np.frombuffer(hashlib.sha256(np.ascontiguousarray(table).tobytes()).digest(), np.uint8)
    out["table_100x128"] = np.asarray(whisper.SinPositionalEncoding(100, 128).sin_positional_encoding())
    # ---- YOHO's own encoder: model.py:9-29 with layers.py:7-116
    s = YOHO_SHAPE
    yparams = ym.init_params(s["n_mels"], s["dims"], s["q_heads"], s["kv_heads"], s["blocks"], seed=s["seed"])
    ymel = np.random.default_rng(3).standard_normal((2, s["frames"], s["n_mels"]))
    yenc = model.AudioEncoder(q_heads=s["q_heads"], kv_heads=s["kv_heads"], dims=s["dims"], seq_len=s["frames"], blocks=s["blocks"])
    out["yoho_mel"] = ymel.astype(np.float32)
    out["yoho_out"] = np.asarray(yenc.apply({"params": yparams}, ymel.astype(np.float32).astype(np.float64)))
    # ---- text decoder: whisper.py:138-165 through Whisper.decode_text (:197-198) and the whole model (Whisper.__call__, :200-203)
    d = DECODER_SHAPE
    dparams = od.init_params(d["vocab"], d["text_seq_len"], d["dims"], d["layers"], seed=d["seed"])
    rng = np.random.default_rng(d["seed"] + 1)
    states = rng.standard_normal((d["batch"], d["audio_frames"], d["dims"])).astype(np.float32)
    tokens = rng.integers(0, d["vocab"], (d["batch"], d["tokens"]))
    aparams = eo.init_params(80, d["dims"], 1, seed=d["seed"])
    dmel = rng.standard_normal((d["batch"], 2 * d["audio_frames"] - 1, 80)).astype(np.float32)
    net = whisper.Whisper(audio_dims=d["dims"], audio_seq_len=d["audio_frames"], audio_heads=d["heads"], audio_layers=1,
                          text_vocab_size=d["vocab"], text_seq_len=d["text_seq_len"], text_dims=d["dims"], text_heads=d["heads"],
                          text_layers=d["layers"])
    variables = {"params": {"audio_encoder": as_float64(aparams), "text_decoder": as_float64(dparams)}}
    out["decoder_states"], out["decoder_tokens"], out["decoder_mel"] = states, tokens.astype(np.int32), dmel
    out["decoder_logits"] = np.asarray(net.apply(variables, states.astype(np.float64), tokens, method=whisper.Whisper.decode_text))
    out["whisper_logits"] = np.asarray(net.apply(variables, dmel.astype(np.float64), tokens))
    return {k: np.asarray(v) for k, v in out.items()}


if __name__ == "__main__":
    res = run_reference()
    np.savez_compressed(GOLDEN / "reference_shim_golden.npz", **res)
    size = (GOLDEN / "reference_shim_golden.npz").stat().st_size
    print(f"wrote reference_shim_golden.npz ({size / 1e6:.2f} MB): " + ", ".join(f"{k}{tuple(v.shape)}" for k, v in res.items()))
