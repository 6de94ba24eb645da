"""GPU parity tests of the Whisper audio encoder: product path (Flax-style apply -> ctypes -> C ABI ->
sm_100a kernels) against the float64 oracle / golden vectors.
Tolerances (BASELINE.json north_star): fp32 path 1e-4; bf16 path 2e-2 max-relative
(max|ours - oracle| / max|oracle|)."""
import ctypes as C

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


def _bf16(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a, np.float32)).to(torch.bfloat16).to(torch.float32).numpy()


def _max_rel(got, want):
    return float(np.abs(got - want).max() / np.abs(want).max())


# ------------------------------------------------------------------ tcgen05 GEMM in isolation
@pytest.mark.parametrize("m,n,k,epi", [(128, 128, 64, 0), (256, 256, 128, 0), (1000, 384, 192, 0), (300, 768, 768, 1),
                                       (4099, 1152, 384, 0), (129, 160, 72, 1), (20000, 2304, 768, 0), (513, 512, 3072, 0)])
def test_tcgen05_gemm(rt, m, n, k, epi):
    import torch
    rng = np.random.default_rng(m + n + k)
    a = rng.standard_normal((m, k)).astype(np.float32)
    w = (rng.standard_normal((k, n)) / np.sqrt(k)).astype(np.float32)
    bias = rng.standard_normal(n).astype(np.float32)
    want = _bf16(a).astype(np.float64) @ _bf16(w).astype(np.float64) + bias
    if epi == 1:
        want = 0.5 * want * (1 + np.tanh(np.sqrt(2 / np.pi) * (want + 0.044715 * want ** 3)))
    ad, wd, bd = (torch.from_numpy(v).cuda() for v in (a, w, bias))
    out = torch.full((m, n), float("nan"), device="cuda")
    nbytes = rt.lib.yoho_test_gemm_workspace_bytes(m, n, k)
    ws = rt.workspace("test", nbytes)
    from yoho_b200._lib import check
    check(rt.lib.yoho_test_gemm_bf16(rt.ctx, C.c_void_p(ad.data_ptr()), C.c_void_p(wd.data_ptr()), C.c_void_p(bd.data_ptr()),
                                     C.c_void_p(out.data_ptr()), m, n, k, epi, C.c_void_p(rt.aligned_ptr(ws)), nbytes, rt.stream_ptr()))
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    assert np.isfinite(got).all()
    assert np.abs(got - want).max() <= 2e-3 * max(1.0, np.abs(want).max())       # fp32 accumulate; tanh.approx in the GELU case


# ------------------------------------------------------------------ tcgen05 attention in isolation
def _attention_case(rt, qkv, tiles):
    """softmax(q k^T / 8) v through the C ABI with the chosen kernel (0 = automatic dispatch) vs float64 numpy."""
    import torch
    from yoho_b200._lib import check
    batch, t, _, heads, _ = qkv.shape
    r = _bf16(qkv).astype(np.float64)
    q, k, v = (r[:, :, i].transpose(0, 2, 1, 3) for i in range(3))       # (B, H, T, 64)
    s = q @ k.transpose(0, 1, 3, 2) / 8.0
    p = np.exp(s - s.max(-1, keepdims=True))
    p /= p.sum(-1, keepdims=True)
    want = (p @ v).transpose(0, 2, 1, 3).reshape(batch, t, heads * 64)
    qd = torch.from_numpy(qkv).cuda()
    out = torch.full((batch, t, heads * 64), float("nan"), device="cuda")
    nbytes = rt.lib.yoho_test_attention_workspace_bytes(batch, t, heads)
    ws = rt.workspace("test", nbytes)
    prev = rt.lib.yoho_test_attention_tiles(tiles)
    try:
        check(rt.lib.yoho_test_attention_bf16(rt.ctx, C.c_void_p(qd.data_ptr()), C.c_void_p(out.data_ptr()), batch, t, heads,
                                              C.c_void_p(rt.aligned_ptr(ws)), nbytes, rt.stream_ptr()))
        torch.cuda.synchronize()
    finally:
        rt.lib.yoho_test_attention_tiles(prev)
    return out.cpu().numpy(), want


@pytest.mark.parametrize("tiles", [0, 2, 3])
@pytest.mark.parametrize("batch,t,heads", [(1, 128, 1), (2, 100, 2), (1, 1500, 2), (3, 300, 6), (2, 257, 1), (2, 384, 2), (1, 385, 1),
                                           (1, 513, 3), (1, 64, 1), (1, 1, 1), (1, 1500, 20),
                                           (40, 130, 4), (50, 60, 4), (9, 400, 20)])   # several work items per CTA, down to one 64-key step each
def test_tcgen05_attention(rt, batch, t, heads, tiles):
    rng = np.random.default_rng(batch * 1000 + t + heads)
    qkv = rng.standard_normal((batch, t, 3, heads, 64)).astype(np.float32)
    qkv[:, :, 0] *= 2.0                                                  # sharper softmax
    got, want = _attention_case(rt, qkv, tiles)
    assert np.isfinite(got).all()
    assert np.abs(got - want).max() <= 1.5e-2 * np.abs(want).max()       # P and the output are rounded to bf16


@pytest.mark.parametrize("tiles", [2, 3])
def test_tcgen05_attention_growing_maximum(rt, tiles):
    """Scores that grow along the first half of the keys and fall along the second: the running maximum rises by more than
    2^8 from one 64-key step to the next, so the lazy rescaling of O and of the row sum runs in every step up to the
    middle; after it the exp2 arguments fall below -126 (the clamp of the polynomial exp2); the last step is ragged
    (1500 = 23 * 64 + 28)."""
    rng = np.random.default_rng(7)
    batch, t, heads = 2, 1500, 2
    qkv = rng.standard_normal((batch, t, 3, heads, 64)).astype(np.float32)
    ramp = 10.0 - np.abs(np.linspace(-9.8, 9.8, t, dtype=np.float32))          # 0.2 -> 10 at the middle key -> 0.2
    qkv[:, :, 0] = np.abs(qkv[:, :, 0]) * 1.5                                  # queries: every component positive
    qkv[:, :, 1] = qkv[:, :, 1] * 0.3 + ramp[None, :, None, None]              # keys: score / 8 ~ 9.6 * ramp (138 log2 units at the peak)
    got, want = _attention_case(rt, qkv, tiles)
    assert np.isfinite(got).all()
    assert np.abs(got - want).max() <= 1.5e-2 * np.abs(want).max()


def test_tcgen05_attention_kernels_agree(rt):
    """Same inputs through the two-tile and the three-tile kernel: both round P to bf16 after the same scaling, so they may
    differ only by the summation order, the lazy-rescaling points and the polynomial exp2 (7.5e-5 relative)."""
    rng = np.random.default_rng(11)
    qkv = rng.standard_normal((2, 700, 3, 3, 64)).astype(np.float32)
    a, want = _attention_case(rt, qkv, 2)
    b, _ = _attention_case(rt, qkv, 3)
    assert np.abs(a - b).max() <= 8e-3 * np.abs(want).max()


# ------------------------------------------------------------------ whole encoder
def _golden(golden_dir):
    from yoho_b200.nn import AudioEncoder, init_encoder_params
    g = np.load(golden_dir / "encoder_golden.npz")
    params = init_encoder_params(int(g["n_mels"]), int(g["dims"]), int(g["n_layers"]), seed=int(g["seed"]))
    model = AudioEncoder(seq_len=100, dims=int(g["dims"]), n_heads=int(g["n_heads"]), n_layers=int(g["n_layers"]))
    return g, params, model


def test_encoder_fp32_golden(rt, golden_dir):
    g, params, model = _golden(golden_dir)
    out = model.apply({"params": params}, g["mel"], compute="fp32")
    assert isinstance(out, np.ndarray) and out.shape == (2, 100, 128) and out.dtype == np.float32
    assert np.abs(out - g["out"]).max() <= TOL_FP32


def test_encoder_bf16_golden(rt, golden_dir):
    g, params, model = _golden(golden_dir)
    out = model.apply({"params": params}, g["mel"], compute="bf16")
    assert np.isfinite(out).all()
    assert _max_rel(out, g["out"]) <= TOL_BF16


@pytest.mark.parametrize("frames", [2999, 3000])
def test_whisper_tiny_full_clip(rt, frames):
    """BASELINE.json configs[0]: Whisper-tiny, batch 1, one 30 s clip, through Whisper.apply(method=encode_audio)."""
    from oracle import encoder as eo
    from oracle import frontend as fe
    from yoho_b200.nn import Whisper, init_encoder_params
    from yoho_b200.synth import synth_clip
    n_mels, dims, heads, layers = 80, 384, 6, 4
    n_samples = 480000 if frames == 2999 else 480240
    mel = fe.log_mel(synth_clip(0, n_samples)[None], norm_mode=1).astype(np.float32)
    assert mel.shape == (1, frames, n_mels)
    params = init_encoder_params(n_mels, dims, layers, seed=42)
    want = eo.audio_encoder(params, mel, heads, np.float64)
    model = Whisper(audio_dims=dims, audio_seq_len=1500, audio_heads=heads, audio_layers=layers)
    variables = {"params": {"audio_encoder": params}}
    out32 = model.apply(variables, mel, method=Whisper.encode_audio, compute="fp32")
    assert out32.shape == (1, 1500, dims)
    assert np.abs(out32 - want).max() <= TOL_FP32
    out16 = model.apply(variables, mel, method=Whisper.encode_audio, compute="bf16")
    assert _max_rel(out16, want) <= TOL_BF16


def test_pipeline_pcm_to_states_batch(rt):
    """log-mel kernels feeding the encoder on the device, batch of 4 base-size clips; bf16 vs fp32 path, and
    batch independence (a property that holds at any size)."""
    import torch
    from yoho_b200.nn import AudioEncoder, init_encoder_params
    from yoho_b200.preprocessing import log_mel_spectogram
    from yoho_b200.synth import synth_batch
    n_mels, dims, heads, layers = 80, 512, 8, 6
    pcm = torch.from_numpy(synth_batch(4, 480000)).cuda()
    mel = log_mel_spectogram(pcm, norm="reference")
    params = init_encoder_params(n_mels, dims, layers, seed=3)
    model = AudioEncoder(1500, dims, heads, layers)
    ref = model.apply({"params": params}, mel, compute="fp32")
    out = model.apply({"params": params}, mel, compute="bf16")
    assert out.is_cuda and out.shape == (4, 1500, dims)
    assert float((out - ref).abs().max() / ref.abs().max()) <= TOL_BF16
    alone = model.apply({"params": params}, mel[2:3].contiguous(), compute="bf16")
    assert torch.equal(alone[0], out[2])


def test_shape_mismatch_errors(rt):
    from yoho_b200.nn import AudioEncoder, init_encoder_params
    params = init_encoder_params(80, 128, 1, seed=0)
    with pytest.raises(ValueError, match="positions"):
        AudioEncoder(1500, 128, 2, 1).apply({"params": params}, np.zeros((1, 199, 80), np.float32))
    with pytest.raises(ValueError, match="Conv_0"):
        AudioEncoder(100, 128, 2, 1).apply({"params": params}, np.zeros((1, 199, 128), np.float32))


# ------------------------------------------------------------------ cross-attention K/V pre-projection
@pytest.mark.parametrize("batch,t,d_in,d_out,layers", [(2, 1500, 384, 384, 4), (3, 100, 512, 256, 2), (1, 1500, 768, 768, 1)])
def test_cross_attention_kv(rt, batch, t, d_in, d_out, layers):
    from oracle import encoder as eo
    from yoho_b200.nn import cross_attention_kv
    params = eo.init_cross_kv_params(d_in, d_out, layers, seed=d_in + t)
    states = np.random.default_rng(t).standard_normal((batch, t, d_in)).astype(np.float32)
    want = eo.cross_attention_kv(params, states, layers)
    got32 = cross_attention_kv(params, states, layers, compute="fp32")
    assert got32.shape == (layers, batch, t, 2 * d_out)
    assert np.abs(got32 - want).max() <= TOL_FP32 * max(1.0, float(np.abs(want).max()))
    got16 = cross_attention_kv(params, states, layers, compute="bf16")
    assert np.isfinite(got16).all() and _max_rel(got16, want) <= TOL_BF16
    # key_proj has no bias (whisper.py:33): zero states give k == 0 and v == value_proj.bias exactly
    zero = cross_attention_kv(params, np.zeros((1, 8, d_in), np.float32), layers, compute="bf16")
    assert np.array_equal(zero[..., :d_out], np.zeros_like(zero[..., :d_out]))
    bias = np.stack([params[f"DecoderBlock_{i}"]["MultiHeadAttention_1"]["value_proj"]["bias"] for i in range(layers)])
    assert np.array_equal(zero[..., d_out:], np.broadcast_to(bias[:, None, None, :], zero[..., d_out:].shape))


# ------------------------------------------------------------------ pipelined host path
@pytest.mark.parametrize("pcm_dtype", ["float32", "int16"])
def test_encode_host_matches_device_path(rt, pcm_dtype):
    """ClipEncoder.encode_host (three streams, double-buffered staging, tapered chunk plan) returns exactly what
    the single-stream device path computes for the same clips, and the log-mel stage matches the oracle."""
    import torch
    from oracle import frontend as ofe
    from yoho_b200.nn import init_encoder_params
    from yoho_b200.pipeline import ClipEncoder, chunk_plan
    from yoho_b200.synth import synth_batch
    shape = (80, 128, 2, 2)
    n_samples, B, chunk = 32000, 19, 4
    enc = ClipEncoder(shape, init_encoder_params(80, 128, 2, seed=5), n_samples=n_samples)
    pcm = synth_batch(B, n_samples, unique=B)
    if pcm_dtype == "int16":
        pcm = np.round(pcm * 32767.0).astype(np.int16)
    assert [n for _, n in chunk_plan(B, chunk)] == [1, 3, 4, 4, 3, 2, 1, 1]
    pcm_host, out_host = enc.host_buffers(B, torch.int16 if pcm_dtype == "int16" else torch.float32)
    pcm_host.copy_(torch.from_numpy(pcm))
    out_host.fill_(float("nan"))
    got = enc.encode_host(pcm_host, out_host, chunk_clips=chunk).numpy().copy()
    plain = enc.encode_host(pcm_host, None, chunk_clips=chunk, taper=False).numpy().copy()
    dev = enc.encode(torch.from_numpy(pcm).cuda()).cpu().numpy()
    assert np.isfinite(got).all()
    # same kernels on the same clips: per-clip results do not depend on how the batch was cut
    assert np.array_equal(got, plain)
    assert _max_rel(got, dev) <= 1e-6
    mel = enc.logmel(torch.from_numpy(pcm[:2]).cuda()).cpu().numpy()
    f32 = pcm[:2].astype(np.float32) / 32768.0 if pcm_dtype == "int16" else pcm[:2]
    want = ofe.normalize_spectogram(ofe.mel_spectogram(f32.astype(np.float64), 400, 160, 16000, 80))
    assert np.abs(mel - want).max() <= 1e-4


def test_encoder_properties_at_bench_size(rt):
    """BASELINE configs[2] per-GPU shard (whisper-small, 256 x 30 s clips), where the float64 oracle is out of reach:
    size-independent properties of the path.  A clip's states do not depend on its position in the batch or on its
    neighbours (clips 128..255 repeat clips 0..127; one clip alone == its slot in the batch), two runs are bit-identical
    (no atomics, no split-k on the data path), every value is finite, and the tcgen05 path agrees with the float32 SIMT
    path (itself checked against the oracle at the sizes above) within the bf16 tolerance."""
    import torch
    from yoho_b200 import MODEL_SHAPES
    from yoho_b200.nn import init_encoder_params
    from yoho_b200.pipeline import ClipEncoder
    from yoho_b200.synth import synth_batch
    n_mels, dims, heads, layers = MODEL_SHAPES["small"]
    params = init_encoder_params(n_mels, dims, layers, seed=42)
    enc = ClipEncoder("small", params)
    half = synth_batch(128, 480000, unique=128)
    pcm = torch.from_numpy(np.concatenate([half, half])).cuda()
    out = enc.encode(pcm)
    assert out.shape == (256, 1500, dims) and bool(torch.isfinite(out).all())
    assert torch.equal(out[:128], out[128:])
    again = enc.encode(pcm)
    assert torch.equal(out, again)
    alone = enc.encode(pcm[37:38].contiguous())
    assert torch.equal(alone[0], out[37])
    chunked = enc.encode(pcm, chunk_clips=48)                  # 48-clip kernel sequences: same tiles per clip
    assert torch.equal(chunked, out)
    enc32 = ClipEncoder("small", params, compute="fp32")
    ref = enc32.encode(pcm[:2].contiguous())
    assert float((out[:2] - ref).abs().max() / ref.abs().max()) <= TOL_BF16
    # distinct clips give distinct states (the batch is not collapsing onto one clip)
    assert float((out[0] - out[1]).abs().max()) > 1e-3


# ------------------------------------------------------------------ binding hygiene
def test_one_pytree_at_two_clip_lengths(rt):
    """The packed blob embeds the sinusoid table (seq_len x D) in front of the layer weights: the same params object used
    at two clip lengths must get two blobs (the cache key holds the descriptor), and each result must match the oracle."""
    from oracle import encoder as eo
    from yoho_b200.nn import AudioEncoder, init_encoder_params
    params = init_encoder_params(80, 128, 2, seed=9)
    rng = np.random.default_rng(0)
    for frames in (199, 61, 199):
        mel = rng.standard_normal((2, frames, 80)).astype(np.float32)
        T = (frames + 2 - 3) // 2 + 1
        got = AudioEncoder(T, 128, 2, 2).apply({"params": params}, mel, compute="fp32")
        want = eo.audio_encoder(params, mel, 2, np.float64)
        assert np.abs(got - want).max() <= TOL_FP32


def test_wrong_leaf_shape_is_refused(rt):
    """A leaf whose element count differs from what the descriptor implies is refused before the library reads it."""
    from yoho_b200.nn import AudioEncoder, init_encoder_params
    params = init_encoder_params(80, 128, 2, seed=9)
    params["EncoderBlock_1"]["Dense_0"]["kernel"] = params["EncoderBlock_1"]["Dense_0"]["kernel"][:, :-1]
    with pytest.raises(ValueError, match="elements"):
        AudioEncoder(100, 128, 2, 2).apply({"params": params}, np.zeros((1, 199, 80), np.float32))


def test_blob_of_another_descriptor_is_refused(rt):
    import torch
    from yoho_b200._lib import EncDesc, YohoError, check
    small, big = EncDesc(80, 128, 2, 1, 100), EncDesc(80, 128, 2, 2, 100)
    nbytes = rt.lib.yoho_encoder_weights_bytes(C.byref(small))
    blob = torch.zeros(nbytes + 1024, dtype=torch.uint8, device="cuda")
    mel = torch.zeros((1, 199, 80), device="cuda")
    out = torch.zeros((1, 100, 128), device="cuda")
    ws_bytes = rt.lib.yoho_encoder_workspace_bytes(C.byref(big), 1, 199, 1)
    ws = rt.workspace("test", ws_bytes)
    with pytest.raises(YohoError, match="blob"):
        check(rt.lib.yoho_encoder_fwd(rt.ctx, C.byref(big), C.c_void_p(rt.aligned_ptr(blob)), rt.blob_bytes(blob), C.c_void_p(mel.data_ptr()),
                                      1, 199, C.c_void_p(out.data_ptr()), 0, 1, C.c_void_p(rt.aligned_ptr(ws)), ws_bytes, rt.stream_ptr()))


def test_empty_batch_is_a_no_op(rt):
    import torch
    from yoho_b200.nn import init_encoder_params
    from yoho_b200.pipeline import ClipEncoder
    from yoho_b200.preprocessing import log_mel_spectogram_chunked
    enc = ClipEncoder((80, 128, 2, 2), init_encoder_params(80, 128, 2, seed=5), n_samples=32000)
    out = enc.encode_host(torch.empty((0, 32000), dtype=torch.float32))
    assert tuple(out.shape) == (0, enc.seq_len, 128)
    assert tuple(enc.encode(torch.empty((0, 32000), device="cuda")).shape) == (0, enc.seq_len, 128)
    assert log_mel_spectogram_chunked(np.zeros(0, np.float32), clip_samples=32000).shape == (0, 199, 80)


def test_ctx_create_keeps_the_current_device(rt):
    """yoho_ctx_create / yoho_filterbank_create allocate on their own device and put the caller's device back."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two devices")
    from yoho_b200.runtime import get_runtime
    torch.cuda.set_device(0)
    get_runtime(1)
    assert torch.cuda.current_device() == 0


def test_graph_replay_matches_plain_launches(rt):
    """BASELINE configs[0] geometry (whisper-tiny, one clip): the CUDA-graph replay of the launch sequence returns exactly
    what the plain launches do, also after the input buffer was refilled in place."""
    import torch
    from yoho_b200.nn import init_encoder_params
    from yoho_b200.pipeline import ClipEncoder
    from yoho_b200.synth import synth_batch
    enc = ClipEncoder("tiny", init_encoder_params(80, 384, 4, seed=42))
    pcm = torch.from_numpy(synth_batch(1, 480000)).cuda()
    plain = enc.encode(pcm).clone()
    out = torch.empty_like(plain)
    assert torch.equal(enc.encode_graphed(pcm, out), plain)
    pcm.copy_(torch.from_numpy(synth_batch(1, 480000, first_index=5)).cuda())
    plain2 = enc.encode(pcm).clone()
    assert not torch.equal(plain2, plain)
    assert torch.equal(enc.encode_graphed(pcm, out), plain2)


def test_allgather_states_c_abi(rt):
    """yoho_allgather_states resolves ncclAllGather from the libnccl already in the process (torch's) and runs it on the
    caller's communicator: here a one-rank communicator built through ctypes, for which the gather is the identity.  The
    multi-rank gather is measured by bench.py under torchrun (``allgather`` key) through torch.distributed."""
    import torch
    from yoho_b200._lib import check
    nccl = C.CDLL("libnccl.so.2")
    comm = C.c_void_p()
    devs = (C.c_int * 1)(torch.cuda.current_device())
    assert nccl.ncclCommInitAll(C.byref(comm), 1, devs) == 0
    try:
        for dtype, code in ((torch.float32, 0), (torch.bfloat16, 1)):
            states = torch.randn((3, 50, 64), device="cuda").to(dtype)
            out = torch.zeros_like(states)
            check(rt.lib.yoho_allgather_states(rt.ctx, comm, C.c_void_p(states.data_ptr()), C.c_void_p(out.data_ptr()),
                                               states.numel(), code, rt.stream_ptr()))
            torch.cuda.synchronize()
            assert torch.equal(out, states)
    finally:
        nccl.ncclCommDestroy.argtypes = [C.c_void_p]
        nccl.ncclCommDestroy(comm)


@pytest.mark.parametrize("switch", ["YOHO_LN_FUSE", "YOHO_PDL"])
def test_ab_switches_still_match_the_oracle(rt, switch):
    """The documented A/B switches, each read once per process: YOHO_LN_FUSE=0 selects the stand-alone
    LayerNorm kernels and plain weights instead of folding LayerNorm_0 / LayerNorm_1 into the GEMMs around them; YOHO_PDL=0
    launches every kernel without programmatic stream serialisation (the `griddepcontrol` instructions become no-ops).  Run
    in a child process on the small all-kernels script, which checks every result against the oracle."""
    import os
    import subprocess
    import sys
    from pathlib import Path
    root = Path(__file__).resolve().parents[1]
    env = dict(os.environ, **{switch: "0"})
    r = subprocess.run([sys.executable, str(root / "tools" / "sanitize_case.py")], env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "[sanitize] done" in r.stdout


@pytest.mark.parametrize("n_mels,dims,heads,layers,frames,batch", [(80, 128, 2, 2, 41, 3), (80, 128, 2, 3, 199, 5), (128, 192, 3, 2, 777, 2),
                                                                   (96, 256, 4, 1, 1023, 1), (80, 320, 5, 2, 130, 7), (64, 64, 1, 2, 3, 2),
                                                                   (80, 448, 7, 1, 257, 3)])
def test_ragged_encoder_shapes_vs_oracle(rt, n_mels, dims, heads, layers, frames, batch):
    """Shapes no BASELINE config has: row counts that are not multiples of the 128-row GEMM tile (ragged last tile of the
    LayerNorm-producer epilogues: residual rows fetched and stored by TMA past the end), sequences shorter than one 64-key
    attention step, odd head counts, widths that are not multiples of 256, mel counts 64 / 96 / 128 in the Conv_0 taps.
    Both compute paths against the float64 oracle (whisper.py:117-135)."""
    from oracle import encoder as eo
    from yoho_b200.nn import AudioEncoder, init_encoder_params
    params = init_encoder_params(n_mels, dims, layers, seed=frames)
    rng = np.random.default_rng(frames + batch)
    mel = rng.standard_normal((batch, frames, n_mels)).astype(np.float32)
    T = eo.conv_out_len(frames)
    want = eo.audio_encoder(params, mel.astype(np.float64), heads, np.float64)
    model = AudioEncoder(T, dims, heads, layers)
    out32 = model.apply({"params": params}, mel, compute="fp32")
    assert out32.shape == want.shape == (batch, T, dims)
    assert np.abs(out32 - want).max() <= 1e-4 * max(1.0, float(np.abs(want).max()))
    out16 = model.apply({"params": params}, mel, compute="bf16")
    assert np.isfinite(out16).all()
    assert np.abs(out16 - want).max() / np.abs(want).max() <= 2e-2
