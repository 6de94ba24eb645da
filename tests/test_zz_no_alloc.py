"""GPU test of the no-allocation log-mel entry point ``yoho_logmel_f32_static`` (no counterpart in the reference).
Collected last on purpose (``pytest -x`` stops at the first failure).

The entry point keeps its per-tile statistics scratch in a file-scope ``__device__`` array of the library instead of a
caller-provided workspace.  It is called through the C ABI and compared with the existing ``yoho_logmel_f32`` on the
same inputs: two calls in a row with different batches (the second call finds the first call's statistics in the
shared scratch and must not be influenced by them), every normaliser, within the log-mel tolerance of the other tests
(1e-4 absolute; the two entry points launch the same kernels, so the results are in fact identical).  No torch allocation
happens between the two calls under test."""
import ctypes as C

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


def _fb(rt, n_mels):
    from yoho_b200.preprocessing.mel_filterbanks import mel_filter_banks
    return rt.filterbank(mel_filter_banks(16000, 400, n_mels))


@pytest.mark.parametrize("norm", [1, 2, 0])
def test_static_scratch_matches_workspace_entry_point(rt, norm):
    import torch
    from yoho_b200._lib import check
    from yoho_b200.synth import synth_batch
    lib, n_samples, n_mels = rt.lib, 48000, 80
    frames = check(lib.yoho_stft_frames(n_samples, 400, 160))
    fb = _fb(rt, n_mels)
    batches = [torch.from_numpy(synth_batch(5, n_samples, first_index=0)).cuda(),
               torch.from_numpy(synth_batch(3, n_samples, first_index=40)).cuda()]
    # everything is allocated before the calls under test
    want = [torch.empty((b.shape[0], frames, n_mels), device="cuda") for b in batches]
    got = [torch.full((b.shape[0], frames, n_mels), float("nan"), device="cuda") for b in batches]
    ws_bytes = lib.yoho_logmel_workspace_bytes(5, n_samples, 400, 160, n_mels)
    ws = rt.workspace("logmel", ws_bytes)
    stream = rt.stream_ptr()
    for pcm, out in zip(batches, want):
        check(lib.yoho_logmel_f32(rt.ctx, fb, C.c_void_p(pcm.data_ptr()), pcm.shape[0], n_samples, 400, 160, C.c_void_p(out.data_ptr()),
                                  norm, C.c_void_p(rt.aligned_ptr(ws)), ws_bytes, stream))
    torch.cuda.synchronize()
    mem0 = torch.cuda.memory_allocated()
    for pcm, out in zip(batches, got):                          # two calls in a row sharing the built-in scratch
        check(lib.yoho_logmel_f32_static(rt.ctx, fb, C.c_void_p(pcm.data_ptr()), pcm.shape[0], n_samples, 400, 160,
                                         C.c_void_p(out.data_ptr()), norm, stream))
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == mem0
    for g, w in zip(got, want):
        g, w = g.cpu().numpy(), w.cpu().numpy()
        assert np.isfinite(g).all()
        if norm == 0:                                           # mel power has no fixed scale: 1e-4 in log10 <=> 2.3e-4 relative
            assert np.all(np.abs(g - w) <= 2.3e-4 * np.abs(w) + 1e-12)
        else:
            assert np.abs(g - w).max() <= TOL
    # and against the oracle, like the other frontend tests
    if norm == 1:
        from oracle import frontend as fe
        ref = fe.log_mel(batches[1].cpu().numpy(), norm_mode=1)
        power = fe.mel_spectogram(batches[1].cpu().numpy().astype(np.float64), 400, 160, 16000, n_mels)
        well_posed = power >= 1e-6 * power.max(axis=-1, keepdims=True)
        assert np.abs(got[1].cpu().numpy() - ref)[well_posed].max() <= TOL


def test_static_scratch_limit_is_enforced(rt):
    """Sizes beyond the built-in scratch are refused with a clear error before anything is launched."""
    import torch
    from yoho_b200._lib import YohoError, check
    lib = rt.lib
    limit = lib.yoho_logmel_static_max_tiles()
    assert limit == 32768
    fb = _fb(rt, 80)
    tiny = torch.zeros(1024, device="cuda")
    frames = check(lib.yoho_stft_frames(480000, 400, 160))
    tiles = (frames + 31) // 32
    too_many = limit // tiles + 1                               # 349 clips of 30 s
    with pytest.raises(YohoError, match="statistics slots"):
        check(lib.yoho_logmel_f32_static(rt.ctx, fb, C.c_void_p(tiny.data_ptr()), too_many, 480000, 400, 160,
                                         C.c_void_p(tiny.data_ptr()), 1, rt.stream_ptr()))
    torch.cuda.synchronize()
