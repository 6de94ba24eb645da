"""GPU parity tests of the Whisper text decoder: product path (Flax-style apply / generate ->
ctypes -> C ABI -> sm_100a kernels) against the float64 oracle (oracle/decoder.py, itself cross-checked against
transformers' WhisperDecoder).  float32 kernels: 1e-4 tolerance relative to the logit scale."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TOL = 1e-4


@pytest.fixture(scope="module")
def rt(built_library):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from yoho_b200.runtime import get_runtime
    return get_runtime()


def _case(V, L_MAX, D, layers, T, B, seed, audio_dims=None):
    from oracle import decoder as od
    from yoho_b200.nn import init_decoder_params
    p = init_decoder_params(V, L_MAX, D, layers, seed=seed, audio_dims=audio_dims)
    q = od.init_params(V, L_MAX, D, layers, seed=seed, audio_dims=audio_dims)
    assert all(np.array_equal(p["Embed_0"]["embedding"], q["Embed_0"]["embedding"]) for _ in (0,))
    rng = np.random.default_rng(seed + 1)
    states = rng.standard_normal((B, T, audio_dims or D)).astype(np.float32)
    return p, states, rng


@pytest.mark.parametrize("V,L_MAX,D,H,layers,T,B,S,audio_dims", [(97, 24, 128, 2, 2, 50, 3, 11, None), (1000, 40, 384, 6, 2, 150, 2, 40, 256),
                                                                 (51865, 448, 512, 8, 1, 100, 1, 9, None)])
def test_decoder_forward_matches_oracle(rt, V, L_MAX, D, H, layers, T, B, S, audio_dims):
    """Whisper.decode_text / TextDecoder.__call__ (whisper.py:147-165, :197-198): teacher-forced logits."""
    from oracle import decoder as od
    from yoho_b200.nn import TextDecoder, Whisper
    p, states, rng = _case(V, L_MAX, D, layers, T, B, 3, audio_dims)
    tokens = rng.integers(0, V, (B, S))
    want = od.text_decoder(p, tokens, states, H, np.float64)
    dec = TextDecoder(V, L_MAX, D, H, layers)
    got = dec.apply({"params": p}, tokens, states)
    assert got.shape == (B, S, V) and np.isfinite(got).all()
    assert np.abs(got - want).max() <= TOL * max(1.0, np.abs(want).max())
    if audio_dims is None:
        model = Whisper(audio_dims=D, audio_seq_len=T, audio_heads=H, audio_layers=1, text_vocab_size=V, text_seq_len=L_MAX,
                        text_dims=D, text_heads=H, text_layers=layers)
        via = model.apply({"params": {"text_decoder": p}}, states, tokens, method=Whisper.decode_text)
        assert np.array_equal(via, got)


@pytest.mark.parametrize("B,use_graph", [(3, False), (3, True), (70, False)])
def test_generate_with_kv_cache_matches_full_rerun(rt, B, use_graph):
    """The cached token-by-token path reproduces the reference's loop (whisper.py:260-285: full decoder re-run per token,
    argmax at pos - 1, positions 0 / 2 / 3 pre-filled): same tokens as the oracle's greedy loop and, step by step, the
    logits of a teacher-forced run over the final sequence."""
    from oracle import decoder as od
    from yoho_b200.nn import TextDecoder
    V, L_MAX, D, H, layers, T, n_steps = 211, 24, 128, 2, 2, 50, 12
    p, states, rng = _case(V, L_MAX, D, layers, T, B, 11)
    buf = np.zeros((B, L_MAX), np.int64)
    buf[:, 0], buf[:, 2], buf[:, 3] = 5, 7, 9
    forced = np.zeros(L_MAX, bool)
    forced[[0, 2, 3]] = True
    dec = TextDecoder(V, L_MAX, D, H, layers)
    if use_graph:
        seq = dec.generate({"params": p}, states, buf, forced=forced, n_steps=n_steps, use_graph=True)
        plain = dec.generate({"params": p}, states, buf, forced=forced, n_steps=n_steps)
        assert np.array_equal(seq, plain)
        return
    seq, logits = dec.generate({"params": p}, states, buf, forced=forced, n_steps=n_steps, return_logits=True)
    fast = dec.generate({"params": p}, states, buf, forced=forced, n_steps=n_steps)
    assert np.array_equal(seq, fast)
    assert np.array_equal(seq[:, [0, 2, 3]], buf[:, [0, 2, 3]]) and (seq[:, n_steps + 1:] == 0).all()
    nb = min(B, 4)
    full = od.text_decoder(p, seq[:nb, :n_steps], states[:nb], H, np.float64)           # (nb, n_steps, V)
    got = np.transpose(logits[:, :nb], (1, 0, 2))                                        # step i = position i
    assert np.abs(got - full).max() <= TOL * max(1.0, np.abs(full).max())
    want = od.greedy_decode(p, states[:nb], H, buf[:nb], forced, n_steps)
    assert np.array_equal(seq[:nb], want[:nb])
    # every clip is independent of its batch neighbours (slices of 64 rows inside the skinny GEMMs)
    if B > 64:
        solo = dec.generate({"params": p}, states[66:67], buf[66:67], forced=forced, n_steps=n_steps)
        assert np.array_equal(solo[0], seq[66])


def test_eot_rows_stay_finished(rt):
    from yoho_b200.nn import TextDecoder
    V, L_MAX, D, H, layers, T, B = 50, 16, 64, 1, 1, 20, 4
    p, states, _ = _case(V, L_MAX, D, layers, T, B, 21)
    buf = np.zeros((B, L_MAX), np.int64)
    buf[:, 0] = 3
    dec = TextDecoder(V, L_MAX, D, H, layers)
    free = dec.generate({"params": p}, states, buf, n_steps=10)
    eot = int(free[0, 4])                       # pretend the token row 0 emits at position 4 is end-of-text
    seq = dec.generate({"params": p}, states, buf, n_steps=10, eot=eot)
    for b in range(B):
        hits = np.nonzero(seq[b, 1:11] == eot)[0]
        if len(hits):
            assert (seq[b, 1 + hits[0]:11] == eot).all()
            assert np.array_equal(seq[b, :1 + hits[0] + 1], free[b, :1 + hits[0] + 1])
    assert (seq[0, 4:11] == eot).all()


@pytest.mark.parametrize("b", [1, 3, 4, 5, 8, 9, 16, 17, 33, 64, 70])
@pytest.mark.parametrize("k,n,gelu,acc", [(128, 96, 0, 0), (512, 1536, 0, 1), (2048, 512, 1, 0), (320, 9600, 0, 0), (1280, 51865, 0, 0)])
def test_skinny_gemm(rt, b, k, n, gelu, acc):
    """The decoder's weight-streaming GEMM in isolation: every row-tile variant, with the cluster k split (n / 32 < 296
    column blocks) and without it (vocabulary-sized n), ragged n and k, GELU and accumulate epilogues."""
    import ctypes as C
    import torch
    from yoho_b200._lib import check
    rng = np.random.default_rng(b * 1000 + k + n)
    x = rng.standard_normal((b, k)).astype(np.float32)
    w = (rng.standard_normal((k, n)) / np.sqrt(k)).astype(np.float32)
    bias = rng.standard_normal(n).astype(np.float32)
    base = rng.standard_normal((b, n)).astype(np.float32)
    want = x.astype(np.float64) @ w.astype(np.float64) + bias
    if gelu:
        want = 0.5 * want * (1 + np.tanh(np.sqrt(2 / np.pi) * (want + 0.044715 * want ** 3)))
    if acc:
        want = want + base
    xd, wd, bd = (torch.from_numpy(v).cuda() for v in (x, w, bias))
    out = torch.from_numpy(base).cuda() if acc else torch.full((b, n), float("nan"), device="cuda")
    check(rt.lib.yoho_test_skinny_gemm(rt.ctx, C.c_void_p(xd.data_ptr()), C.c_void_p(wd.data_ptr()), C.c_void_p(bd.data_ptr()),
                                       C.c_void_p(out.data_ptr()), b, k, n, gelu, acc, rt.stream_ptr()))
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    assert np.isfinite(got).all()
    assert np.abs(got - want).max() <= 2e-5 * max(1.0, np.abs(want).max())


def test_transcriber_end_to_end(rt):
    """PCM -> log-mel -> encoder -> KV-cached greedy decode through yoho_b200.pipeline.Transcriber (fp32 compute, so that
    the token choice is comparable) equals the oracle chain: frontend -> audio_encoder -> full decoder re-run per token."""
    from oracle import decoder as od, encoder as oe, frontend as ofe
    from yoho_b200.nn import init_decoder_params, init_encoder_params
    from yoho_b200.pipeline import Transcriber
    from yoho_b200.synth import synth_batch
    n_samples, B, D, H, V, L_MAX, n_new = 16000, 2, 128, 2, 300, 16, 8
    params = {"audio_encoder": init_encoder_params(80, D, 2, seed=2), "text_decoder": init_decoder_params(V, L_MAX, D, 2, seed=3)}
    pcm = synth_batch(B, n_samples, unique=B)
    tr = Transcriber((80, D, H, 2), params, text_heads=H, text_seq_len=L_MAX, n_samples=n_samples, compute="fp32")
    prompt = {0: 7, 2: 11, 3: 13}
    got = tr.transcribe(pcm, prompt, max_new_tokens=n_new)
    assert got.shape == (B, L_MAX)
    mel = ofe.normalize_spectogram(ofe.mel_spectogram(pcm.astype(np.float64), 400, 160, 16000, 80))
    states = oe.audio_encoder(params["audio_encoder"], mel, H, np.float64)
    buf = np.zeros((B, L_MAX), np.int64)
    forced = np.zeros(L_MAX, bool)
    for pos, tok in prompt.items():
        buf[:, pos], forced[pos] = tok, True
    want = od.greedy_decode(params["text_decoder"], states, H, buf, forced, n_new)
    assert np.array_equal(got, want)
    on_dev = tr.transcribe(__import__("torch").from_numpy(pcm).cuda(), prompt, max_new_tokens=n_new, use_graph=True)
    assert np.array_equal(on_dev.cpu().numpy(), want)
