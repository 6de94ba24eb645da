"""CPU tests of the host layer: the C-ABI library loads and exports every declared symbol, the shape
helpers agree with the oracle, the reference-signature mirror validates arguments like the reference, and
the product path refuses to run without the CUDA device (no CPU fallback).  No compute calls here."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

from oracle import encoder as enc_oracle
from oracle import frontend as fe_oracle

ROOT = Path(__file__).resolve().parents[1]


def test_library_exports_every_declared_symbol(built_library):
    header = (ROOT / "include" / "yoho_b200.h").read_text()
    declared = set(re.findall(r"\b(yoho_[a-z0-9_]+)\s*\(", header))
    declared -= {"yoho_status", "yoho_norm", "yoho_compute", "yoho_dtype"}
    assert len(declared) >= 20
    lib = C.CDLL(str(built_library))
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in include/yoho_b200.h but not exported"
    from yoho_b200 import _lib
    assert set(_lib.SIGNATURES) == declared                      # the ctypes table covers the whole header


def test_shape_helpers_match_oracle(built_library):
    from yoho_b200 import _lib
    lib = _lib.load()
    for n in (400, 401, 16400, 164080, 480000, 480240, 481234):
        assert lib.yoho_stft_frames(n, 400, 160) == fe_oracle.stft_geometry(n, 400, 160)[1]
    assert lib.yoho_stft_frames(399, 400, 160) < 0 and b"n_samples" in lib.yoho_last_error()
    for f in (1, 2, 199, 1024, 2999, 3000):
        assert lib.yoho_encoder_out_len(f) == enc_oracle.conv_out_len(f)
    d = _lib.EncDesc(80, 768, 12, 12, 1500)
    assert lib.yoho_encoder_num_tensors(C.byref(d)) == 7 + 15 * 12
    n_params = 87_000_000
    assert 6 * n_params < lib.yoho_encoder_weights_bytes(C.byref(d)) < 7 * n_params     # fp32 + bf16 copies
    bad = _lib.EncDesc(80, 500, 8, 6, 1500)                                              # d_head != 64
    assert lib.yoho_encoder_weights_bytes(C.byref(bad)) == 0 and b"d_head" in lib.yoho_last_error()
    assert lib.yoho_encoder_workspace_bytes(C.byref(d), 256, 2999, 1) > 256 * 1500 * 768 * 4
    assert lib.yoho_logmel_workspace_bytes(256, 480000, 400, 160, 80) >= 256 * 94 * 8


def test_flatten_order_matches_header():
    from yoho_b200.nn import flatten_encoder_params, init_encoder_params, sin_positional_encoding
    p = init_encoder_params(80, 128, 3, seed=1)
    flat = flatten_encoder_params(p, 3, sin_positional_encoding(100, 128))
    assert len(flat) == 7 + 15 * 3
    assert flat[0].shape == (3, 80, 128) and flat[2].shape == (3, 128, 128) and flat[4].shape == (100, 128)
    blk = flat[7:22]
    assert [a.shape for a in blk] == [(128,), (128,), (128, 128), (128,), (128, 128), (128, 128), (128,), (128, 128), (128,),
                                      (128,), (128,), (128, 512), (512,), (512, 128), (128,)]
    assert blk[4] is p["EncoderBlock_0"]["MultiHeadAttention_0"]["key_proj"]["kernel"]          # key_proj has no bias
    o = enc_oracle.init_params(80, 128, 3, seed=1)                                             # same generator as the oracle
    assert np.array_equal(o["EncoderBlock_2"]["Dense_1"]["kernel"], p["EncoderBlock_2"]["Dense_1"]["kernel"])
    with pytest.raises(KeyError):
        flatten_encoder_params({k: v for k, v in p.items() if k != "Conv_1"}, 3, flat[4])
    # the element counts a binding checks before packing (yoho_encoder_tensor_elems) are those of the flattened leaves
    from yoho_b200 import _lib
    lib, d = _lib.load(), _lib.EncDesc(80, 128, 2, 3, 100)
    assert [lib.yoho_encoder_tensor_elems(C.byref(d), i) for i in range(len(flat))] == [int(np.prod(a.shape)) for a in flat]
    assert lib.yoho_encoder_tensor_elems(C.byref(d), len(flat)) < 0


def test_allgather_states_argument_checks(built_library):
    """No GPU, no communicator: the optional gather entry point refuses null arguments and unknown dtypes."""
    from yoho_b200 import _lib
    lib = _lib.load()
    assert lib.yoho_allgather_states(None, None, None, None, 16, 0, None) == -1
    assert b"null" in lib.yoho_last_error()


def test_product_filterbank_and_table_are_bit_exact(golden_dir):
    from yoho_b200.nn import sin_positional_encoding
    from yoho_b200.preprocessing import mel_filter_banks
    for m in (80, 128):
        assert np.array_equal(mel_filter_banks(16000, 400, m), np.load(golden_dir / f"mel_filter_banks_{m}.npy"))
    assert np.array_equal(sin_positional_encoding(1500, 384), enc_oracle.sinusoid_table(1500, 384))


def test_config_mirror():
    from yoho_b200.config import YOHOConfig
    cfg = YOHOConfig(sample_rate=16000, n_mel_bands=128, n_fft=400, stft_hop=160, max_audio_len=1024, max_text_len=256,
                     dims=384, n_audio_heads=6, n_audio_blocks=4, n_text_heads=6, n_text_blocks=4)
    assert cfg.n_samples == 164080 and abs(cfg.max_input_seconds - 10.255) < 1e-9     # reference train/config.toml shape


def test_argument_validation_like_the_reference():
    from yoho_b200.nn import AudioEncoder, Whisper
    from yoho_b200.preprocessing import mel_spectogram
    with pytest.raises(ValueError):                      # the reference einsum is batched-only (audio.py:36)
        mel_spectogram(np.zeros(480000, np.float32), 400, 160, 16000, 80)
    with pytest.raises(ValueError):
        AudioEncoder(1500, 384, 6, 4).apply({"params": {}}, np.zeros((2999, 80), np.float32))
    with pytest.raises(ValueError):                      # built without text_* fields: no decoder to run
        Whisper(384, 1500, 6, 4).apply({"params": {}}, None, None, method=Whisper.decode_text)


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    from yoho_b200.preprocessing import mel_spectogram
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        mel_spectogram(np.zeros((1, 480000), np.float32), 400, 160, 16000, 80)


def test_product_does_not_import_oracle():
    for path in (ROOT / "yoho_b200").rglob("*.py"):
        text = path.read_text()
        assert not re.search(r"^\s*(from|import)\s+oracle\b", text, re.M), f"{path} imports the oracle"


def test_shard_range_partitions():
    from yoho_b200.dist import shard_range
    for n, w in ((2048, 8), (120, 8), (7, 4), (3, 8)):
        spans = [shard_range(n, r, w) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
        sizes = [b - a for a, b in spans]
        assert max(sizes) - min(sizes) <= 1


def test_chunk_plan_partitions_the_batch():
    """The host pipeline's chunk plan (yoho_b200.pipeline.chunk_plan) covers every clip once, in order, never
    exceeds the staging size, and tapers only when the batch holds at least two chunks."""
    from yoho_b200.pipeline import chunk_plan
    for batch in (1, 7, 32, 127, 128, 130, 256, 1000):
        for chunk in (1, 2, 5, 32, 64):
            for taper in (False, True):
                plan = chunk_plan(batch, chunk, taper)
                assert plan[0][0] == 0 and sum(n for _, n in plan) == batch
                assert all(0 < n <= chunk for _, n in plan)
                assert all(plan[i][0] + plan[i][1] == plan[i + 1][0] for i in range(len(plan) - 1))
    assert [n for _, n in chunk_plan(256, 32)] == [8, 24, 32, 32, 32, 32, 32, 32, 16, 8, 8]
    assert [n for _, n in chunk_plan(64, 32)] == [8, 24, 16, 8, 8]
    assert [n for _, n in chunk_plan(48, 32, taper=True)] == [32, 16]


def test_bench_clock_sampler_reports_only_the_timed_region():
    """bench.py starts nvidia-smi during the warm-up and keeps the samples stamped inside the timed region; a region shorter
    than the sampling period falls back to the last samples before it and says so."""
    import datetime
    import importlib.util
    from pathlib import Path
    spec = importlib.util.spec_from_file_location("bench_mod", Path(__file__).resolve().parents[1] / "bench.py")
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    rows = ["0, 1965, 1965, 310.0, 0x0, Not Active, Not Active, Not Active, Not Active, 2026/11/05 10:00:00.100",
            "0, 1400, 1965, 990.0, 0x4, Not Active, Not Active, Not Active, Active, 2026/11/05 10:00:01.100",
            "0, 1320, 1965, 995.0, 0x4, Not Active, Not Active, Not Active, Active, 2026/11/05 10:00:02.100",
            "0, 1340, 1965, 991.0, 0x4, Not Active, Not Active, Not Active, Active, 2026/11/05 10:00:02.300"]
    out = "\n".join(rows)
    mark = datetime.datetime(2026, 11, 5, 10, 0, 2, 0)
    c = bench.ClockSampler.summarise(out, mark)
    assert c["samples"] == 2 and c["sm_mhz"] == 1330.0 and c["reasons"] == ["sw_power_cap"] and c["window"] == "timed region"
    assert bench.ClockSampler.summarise(out, None)["samples"] == 4
    late = bench.ClockSampler.summarise(out, datetime.datetime(2026, 11, 5, 10, 0, 3, 0))
    assert late["samples"] == 4 and late["window"].startswith("end of warm-up")
    hot = bench.ClockSampler.summarise(rows[1].replace("Not Active, Active, 2026", "Active, Active, 2026"), None)
    assert "sw_thermal_slowdown" in hot["reasons"]
    assert bench.ClockSampler.summarise("", None)["reasons"] == ["no samples"]
