"""GPU parity tests of YOHO's own audio encoder: product path (Flax-style apply ->
ctypes -> C ABI -> sm_100a kernels) against the float64 oracle (oracle/yoho_model.py).
Tolerances as for the Whisper path: fp32 path 1e-4 max-abs (relative to max|oracle| when that exceeds 1);
bf16/tcgen05 path 2e-2 max-relative."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL_FP32 = 1e-4
TOL_BF16 = 2e-2


@pytest.fixture(scope="module")
def rt(built_library):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from yoho_b200.runtime import get_runtime
    return get_runtime()


def _case(n_mels, dims, qh, kvh, blocks, batch, frames, seed):
    from oracle import yoho_model as ym
    params = ym.init_params(n_mels, dims, qh, kvh, blocks, seed=seed)
    mel = np.random.default_rng(seed + 1).standard_normal((batch, frames, n_mels)).astype(np.float32)
    want = ym.audio_encoder(params, mel, qh, kvh)
    return params, mel, want


@pytest.mark.parametrize("n_mels,dims,qh,kvh,blocks,batch,frames", [
    (128, 128, 2, 1, 2, 2, 64),       # small
    (128, 384, 6, 3, 4, 2, 1024),     # the reference config (train/config.toml:7-18)
    (64, 256, 4, 4, 1, 3, 200),       # no grouping, ragged tiles
    (128, 384, 6, 3, 1, 1, 301),      # odd frame count: max_pool drops the tail row
])
def test_yoho_encoder_fp32_and_bf16(rt, n_mels, dims, qh, kvh, blocks, batch, frames):
    from yoho_b200.nn.model import AudioEncoder
    params, mel, want = _case(n_mels, dims, qh, kvh, blocks, batch, frames, seed=frames)
    enc = AudioEncoder(qh, kvh, dims, max(frames, 1024), blocks)
    scale = max(1.0, float(np.abs(want).max()))
    got32 = enc.apply({"params": params}, mel, compute="fp32")
    assert got32.shape == want.shape == (batch, frames // 2, dims)
    assert np.isfinite(got32).all()
    assert np.abs(got32 - want).max() <= TOL_FP32 * scale
    got16 = enc.apply({"params": params}, mel, compute="bf16")
    assert np.isfinite(got16).all()
    assert np.abs(got16 - want).max() / np.abs(want).max() <= TOL_BF16


def test_model_apply_signature_and_device_tensors(rt):
    """Model(config, vocab).apply({"params": {"encoder": ...}}, spectogram, method=Model.encode_audio) as at
    transcription_pretrain_no_voiceprints.py:198-203, with a CUDA tensor in and out."""
    import torch
    from yoho_b200.config import YOHOConfig
    from yoho_b200.nn.model import Model
    cfg = YOHOConfig(sample_rate=16000, n_mel_bands=128, n_fft=400, stft_hop=160, max_audio_len=256, max_text_len=16, dims=128,
                     n_audio_heads=2, n_audio_blocks=1, n_text_heads=2, n_text_blocks=1)
    params, mel, want = _case(128, 128, 2, 1, 1, 2, 256, seed=11)
    model = Model(cfg, 100)
    out = model.apply({"params": {"encoder": params}}, torch.from_numpy(mel).cuda(), method=Model.encode_audio, compute="fp32")
    assert out.is_cuda and tuple(out.shape) == (2, 128, 128)
    assert np.abs(out.cpu().numpy() - want).max() <= TOL_FP32 * max(1.0, float(np.abs(want).max()))
    with pytest.raises(NotImplementedError):
        model.apply({"params": {"encoder": params}}, mel, method=Model.decode_text)


def test_yoho_encoder_rejects_bad_shapes(rt):
    from yoho_b200._lib import YohoError
    from yoho_b200.nn.model import AudioEncoder
    params, mel, _ = _case(128, 128, 2, 1, 1, 1, 64, seed=3)
    with pytest.raises(ValueError):
        AudioEncoder(2, 1, 128, 32, 1).apply({"params": params}, mel)            # more frames than seq_len
    with pytest.raises(ValueError):
        AudioEncoder(2, 1, 256, 64, 1).apply({"params": params}, mel)            # Conv_0 kernel / dims mismatch
    with pytest.raises(YohoError):
        AudioEncoder(4, 2, 128, 64, 1).apply({"params": params}, mel)            # head_dim != 64
