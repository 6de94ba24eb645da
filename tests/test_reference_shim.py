"""The oracle -- and, on the GPU, the CUDA path -- against outputs of THE REFERENCE'S OWN SOURCE FILES.

jax / flax cannot be installed here, so ``oracle/refshim`` supplies numpy stand-ins for the dozen library names the reference's
hot path imports, and ``yoho/src/preprocessing/audio.py``, ``yoho/src/nn/whisper.py``, ``yoho/src/nn/model.py`` and
``layers.py`` are imported from /root/reference UNMODIFIED and executed.  That pins what the reference's authors wrote (the
composition and order of operations, scale factors, transposes, einsum / einops patterns, bias-free projections, the
parameter pytree's names); the inside of the library primitives stays the shim's restatement (docstring of oracle/refshim).

* live tests (this container only: /root/reference is not on the GPU box) run the reference code and hold the oracle AND the
  committed vectors to it;
* ``tests/golden/reference_shim_golden.npz`` (made by tests/golden/make_reference_golden.py) travels: the oracle is held to
  it everywhere, the CUDA path in the ``gpu`` tests below, with the north_star tolerances (log-mel 1e-4 absolute; encoder 1e-4
  on the float32 path, 2e-2 max-relative on the bf16 path)."""
import hashlib

import numpy as np
import pytest

from oracle import decoder as od
from oracle import encoder as eo
from oracle import frontend as fo
from oracle import refshim
from oracle import yoho_model as ym

needs_reference = pytest.mark.skipif(not refshim.reference_available(), reason="/root/reference is not mounted (GPU box)")


@pytest.fixture(scope="module")
def ref_golden(golden_dir):
    return np.load(golden_dir / "reference_shim_golden.npz")


def _encoder_case(golden_dir):
    eg = np.load(golden_dir / "encoder_golden.npz")
    n_mels, dims, heads, layers, seed = (int(eg[k]) for k in ("n_mels", "dims", "n_heads", "n_layers", "seed"))
    return eg["mel"], eo.init_params(n_mels, dims, layers, seed=seed), heads, layers, dims


def _yoho_case():
    import importlib.util
    from pathlib import Path
    spec = importlib.util.spec_from_file_location("make_reference_golden", Path(__file__).parent / "golden" / "make_reference_golden.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    s = mod.YOHO_SHAPE
    return s, ym.init_params(s["n_mels"], s["dims"], s["q_heads"], s["kv_heads"], s["blocks"], seed=s["seed"]), mod


# ------------------------------------------------------------------ live: the reference's code, executed here
@needs_reference
def test_committed_vectors_are_what_the_reference_produces(ref_golden):
    _, _, mod = _yoho_case()
    live = mod.run_reference()
    assert set(live) == set(ref_golden.files)
    for key, val in live.items():
        assert np.array_equal(val, ref_golden[key]), key


@needs_reference
def test_reference_module_structure_names_the_pytree():
    """The parameter names the product's flattener and the oracle expect come out of the reference's own module code
    through Flax's naming rules: a tree with one leaf renamed is refused with the offending name."""
    _, whisper, _, _ = refshim.load_reference()
    params = eo.init_params(80, 128, 1, seed=0)
    net = whisper.AudioEncoder(seq_len=100, dims=128, n_heads=2, n_layers=1)
    mel = np.zeros((1, 199, 80))
    assert np.asarray(net.apply({"params": params}, mel)).shape == (1, 100, 128)
    for path in (("Conv_1",), ("EncoderBlock_0", "MultiHeadAttention_0", "key_proj"), ("EncoderBlock_0", "LayerNorm_1"), ("LayerNorm_0",)):
        broken = {k: (dict(v) if isinstance(v, dict) else v) for k, v in params.items()}
        node = broken
        for key in path[:-1]:
            node[key] = dict(node[key])
            node = node[key]
        node["renamed"] = node.pop(path[-1])
        with pytest.raises(KeyError, match=path[-1]):
            net.apply({"params": broken}, mel)
    # key_proj has no bias in the reference (whisper.py:33): a tree WITH one is simply not read, one WITHOUT query bias fails
    blk = params["EncoderBlock_0"]["MultiHeadAttention_0"]
    assert "bias" not in blk["key_proj"]
    broken = dict(params); broken["EncoderBlock_0"] = dict(params["EncoderBlock_0"])
    broken["EncoderBlock_0"]["MultiHeadAttention_0"] = dict(blk, query_proj={"kernel": blk["query_proj"]["kernel"]})
    with pytest.raises(KeyError, match="bias"):
        net.apply({"params": broken}, mel)


@needs_reference
@pytest.mark.parametrize("n_samples,n_mels", [(16000, 80), (16001, 128), (164080, 128)])
def test_live_frontend_float32_run(n_samples, n_mels):
    """A dtype-faithful (float32) run of the reference's frontend code stays within the float32 envelope of the float64
    oracle; a float64 run agrees to round-off."""
    audio, _, _, _ = refshim.load_reference()
    from yoho_b200.synth import synth_batch
    x = synth_batch(2, n_samples)
    want = fo.mel_spectogram(x.astype(np.float64), 400, 160, 16000, n_mels)
    got64 = np.asarray(audio.mel_spectogram(x.astype(np.float64), 400, 160, 16000, n_mels))
    assert np.abs(got64 - want).max() <= 1e-13 * want.max()
    assert np.abs(np.asarray(audio.normalize_spectogram(got64)) - fo.normalize_spectogram(want)).max() <= 1e-9
    got32 = np.asarray(audio.mel_spectogram(x, 400, 160, 16000, n_mels))
    assert got32.dtype == np.float32
    strong = want > want.max() * 1e-6
    assert np.abs(got32[strong] / want[strong] - 1.0).max() <= 2e-3


# ------------------------------------------------------------------ the oracle against the committed vectors (runs everywhere)
def test_oracle_frontend_equals_reference(golden_dir, ref_golden):
    pcm = np.load(golden_dir / "frontend_golden.npz")["audio"][ref_golden["pcm_clips"]].astype(np.float64)
    for m in (80, 128):
        want = ref_golden[f"m{m}_power"]
        got = fo.mel_spectogram(pcm, 400, 160, 16000, m)
        assert np.abs(got - want).max() <= 1e-13 * want.max()
        assert np.abs(fo.normalize_spectogram(got) - ref_golden[f"m{m}_reference"]).max() <= 1e-9
        assert np.abs(fo.log_mel(pcm, n_mels=m, norm_mode=fo.NORM_REFERENCE) - ref_golden[f"m{m}_reference"]).max() <= 1e-9


def test_oracle_encoder_equals_reference(golden_dir, ref_golden):
    mel, params, heads, _, _ = _encoder_case(golden_dir)
    got = eo.audio_encoder(params, mel.astype(np.float64), heads, np.float64)
    assert np.abs(got - ref_golden["encoder_out"]).max() <= 1e-11
    # the float32 golden the GPU tests have used since round 1 is the same function
    assert np.abs(np.load(golden_dir / "encoder_golden.npz")["out"] - ref_golden["encoder_out"]).max() <= 1e-5


def test_sinusoid_table_is_bit_equal_to_the_reference(ref_golden):
    from yoho_b200.nn import sin_positional_encoding
    assert np.array_equal(eo.sinusoid_table(100, 128), ref_golden["table_100x128"])
    assert np.array_equal(sin_positional_encoding(100, 128), ref_golden["table_100x128"])
    for key in ref_golden.files:
        if key.startswith("table_sha256_"):
            length, dim = (int(v) for v in key.rsplit("_", 1)[1].split("x"))
            for table in (eo.sinusoid_table(length, dim), sin_positional_encoding(length, dim)):
                digest = np.frombuffer(hashlib.sha256(np.ascontiguousarray(table).tobytes()).digest(), np.uint8)
                assert np.array_equal(digest, ref_golden[key]), key


def test_oracle_yoho_encoder_equals_reference(ref_golden):
    s, params, _ = _yoho_case()
    got = ym.audio_encoder(params, ref_golden["yoho_mel"].astype(np.float64), s["q_heads"], s["kv_heads"], np.float64)
    want = ref_golden["yoho_out"]
    # the reference forms the RoPE rotation e^{i angle} in complex64 (x64 disabled); the oracle takes cos / sin of the same
    # float32 angles in float64: they differ by float32 round-off of the rotation, 1e-7 relative
    assert np.abs(got - want).max() <= 5e-6 * np.abs(want).max()


def _decoder_case(mod):
    d = mod.DECODER_SHAPE
    return d, od.init_params(d["vocab"], d["text_seq_len"], d["dims"], d["layers"], seed=d["seed"]), eo.init_params(80, d["dims"], 1, seed=d["seed"])


def test_oracle_decoder_equals_reference(ref_golden):
    """TextDecoder (whisper.py:138-165) through Whisper.decode_text, and the whole model (Whisper.__call__, :200-203)."""
    _, _, mod = _yoho_case()
    d, dparams, aparams = _decoder_case(mod)
    tokens = ref_golden["decoder_tokens"]
    got = od.text_decoder(dparams, tokens, ref_golden["decoder_states"], d["heads"], np.float64)
    assert np.abs(got - ref_golden["decoder_logits"]).max() <= 1e-11
    states = eo.audio_encoder(aparams, ref_golden["decoder_mel"].astype(np.float64), d["heads"], np.float64)
    got = od.text_decoder(dparams, tokens, states, d["heads"], np.float64)
    assert np.abs(got - ref_golden["whisper_logits"]).max() <= 1e-10


@needs_reference
def test_live_decoder_float32_run(ref_golden):
    """Token ids carry no floating dtype: with the float32 leaves a checkpoint holds, the reference's decoder code runs in
    float32 and lands within float32 round-off of the float64 vectors."""
    _, whisper, _, _ = refshim.load_reference()
    _, _, mod = _yoho_case()
    d, dparams, _ = _decoder_case(mod)
    net = whisper.Whisper(audio_dims=d["dims"], audio_seq_len=d["audio_frames"], audio_heads=d["heads"], audio_layers=1,
                          text_vocab_size=d["vocab"], text_seq_len=d["text_seq_len"], text_dims=d["dims"], text_heads=d["heads"],
                          text_layers=d["layers"])
    got = np.asarray(net.apply({"params": {"text_decoder": dparams}}, ref_golden["decoder_states"], ref_golden["decoder_tokens"],
                               method=whisper.Whisper.decode_text))
    want = ref_golden["decoder_logits"]
    assert got.dtype == np.float32
    assert 0 < np.abs(got - want).max() <= 1e-5 * np.abs(want).max()


# ------------------------------------------------------------------ the CUDA path against the reference's outputs
@pytest.fixture(scope="module")
def gpu(built_library):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    return True


@pytest.mark.gpu
@pytest.mark.parametrize("n_mels", [80, 128])
def test_gpu_frontend_vs_reference(gpu, golden_dir, ref_golden, n_mels):
    from yoho_b200 import preprocessing as pp
    pcm = np.load(golden_dir / "frontend_golden.npz")["audio"][ref_golden["pcm_clips"]]
    power = pp.mel_spectogram(pcm, 400, 160, 16000, n_mels)
    want = ref_golden[f"m{n_mels}_power"]
    assert np.all(np.abs(power - want) <= 2.3e-4 * np.abs(want) + 1e-12)            # 1e-4 in log10
    assert np.abs(pp.normalize_spectogram(power) - ref_golden[f"m{n_mels}_reference"]).max() <= 1e-4
    assert np.abs(pp.log_mel_spectogram(pcm, n_mels=n_mels, norm="reference") - ref_golden[f"m{n_mels}_reference"]).max() <= 1e-4


@pytest.mark.gpu
def test_gpu_encoder_vs_reference(gpu, golden_dir, ref_golden):
    from yoho_b200.nn import Whisper
    mel, params, heads, layers, dims = _encoder_case(golden_dir)
    want = ref_golden["encoder_out"]
    net = Whisper(audio_dims=dims, audio_seq_len=want.shape[1], audio_heads=heads, audio_layers=layers)
    variables = {"params": {"audio_encoder": params}}
    out32 = net.apply(variables, mel, method=Whisper.encode_audio, compute="fp32")
    assert np.abs(out32 - want).max() <= 1e-4
    out16 = net.apply(variables, mel, method=Whisper.encode_audio, compute="bf16")
    assert np.abs(out16 - want).max() / np.abs(want).max() <= 2e-2


@pytest.mark.gpu
def test_gpu_yoho_encoder_vs_reference(gpu, ref_golden):
    from yoho_b200.nn.model import AudioEncoder
    s, params, _ = _yoho_case()
    want = ref_golden["yoho_out"]
    enc = AudioEncoder(q_heads=s["q_heads"], kv_heads=s["kv_heads"], dims=s["dims"], seq_len=s["frames"], blocks=s["blocks"])
    out32 = enc.apply({"params": params}, ref_golden["yoho_mel"], compute="fp32")
    assert np.abs(out32 - want).max() <= 1e-4 * max(1.0, float(np.abs(want).max()))
    out16 = enc.apply({"params": params}, ref_golden["yoho_mel"], compute="bf16")
    assert np.abs(out16 - want).max() / np.abs(want).max() <= 2e-2


@pytest.mark.gpu
def test_gpu_decoder_vs_reference(gpu, ref_golden):
    """The float32 decoder kernels against the reference's decode_text output; then the whole model
    (encoder float32 -> decoder) against Whisper.__call__'s."""
    from yoho_b200.nn import Whisper
    _, _, mod = _yoho_case()
    d, dparams, aparams = _decoder_case(mod)
    net = Whisper(audio_dims=d["dims"], audio_seq_len=d["audio_frames"], audio_heads=d["heads"], audio_layers=1,
                  text_vocab_size=d["vocab"], text_seq_len=d["text_seq_len"], text_dims=d["dims"], text_heads=d["heads"],
                  text_layers=d["layers"])
    variables = {"params": {"audio_encoder": aparams, "text_decoder": dparams}}
    tokens = ref_golden["decoder_tokens"].astype(np.int64)
    want = ref_golden["decoder_logits"]
    got = np.asarray(net.apply(variables, ref_golden["decoder_states"], tokens, method=Whisper.decode_text))
    assert np.abs(got - want).max() <= 1e-4 * max(1.0, np.abs(want).max())
    want = ref_golden["whisper_logits"]
    got = np.asarray(net.apply(variables, ref_golden["decoder_mel"], tokens, compute="fp32"))
    assert np.abs(got - want).max() <= 1e-4 * max(1.0, np.abs(want).max())
