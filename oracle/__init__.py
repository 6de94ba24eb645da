"""CPU oracle for the YOHO log-mel + Whisper-encoder hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is imported by the product
package ``yoho_b200``; only ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may use it, and there
only as the checker or the timed CPU baseline, never as the thing shipped.

Parity status: the reference (JirkaKlimes/YOHO) ships zero tests and its hot
path imports jax/flax, which exist neither in the build container nor on the
GPU box, so the reference cannot run AS SHIPPED and there are no golden vectors
of its own.  What pins this oracle:

* ``mel_filter_banks`` — the reference's own pure-NumPy function
  (yoho/src/preprocessing/mel_filterbanks.py:21-42) is importable; its output
  for (16000, 400, 80) and (16000, 400, 128) is committed bit-for-bit under
  ``tests/golden/`` by ``tests/golden/make_golden.py`` and every filterbank in
  this repo is compared against it with ``==``.
* THE REFERENCE'S OWN SOURCE FILES, executed (``oracle/refshim``, round 2):
  ``audio.py``, ``nn/whisper.py``, ``nn/model.py`` and ``nn/layers.py`` are
  imported unmodified from /root/reference on numpy stand-ins for the dozen
  jax / flax names they use, and their outputs are committed as
  ``tests/golden/reference_shim_golden.npz``.  The oracle agrees with them to
  1e-13 (log-mel), 0.0 / 1e-11 (Whisper encoder, float64), bit-for-bit (sinusoid
  table, float32), 1e-7 relative (YOHO's encoder: complex64 RoPE factors) and
  0.0 (text decoder and the whole Whisper.__call__, float64).
  This pins what the reference's authors wrote -- composition, order, scales,
  transposes, einsum patterns, pytree names -- but NOT the inside of the
  library primitives (stft, Dense, Conv, LayerNorm, gelu, softmax, RMSNorm,
  max_pool), which the stand-ins restate from the documented jax 0.4.28 /
  flax 0.8.4 behaviour.  "Parity unpinned" therefore still applies to those
  primitives, and to them only.
* independent witnesses for those primitives: ``scipy.signal.stft`` (the
  routine ``jax.scipy.signal.stft`` mirrors), ``torch.stft``, ``transformers``
  ``WhisperEncoder`` / ``WhisperDecoder`` sub-modules driven by hand, torch
  ``conv1d`` / ``scaled_dot_product_attention`` (``tests/test_oracle_witness.py``,
  ``tests/test_oracle_witness2.py``).
"""
