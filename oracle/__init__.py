"""CPU oracle for the YOHO log-mel + Whisper-encoder hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is imported by the product
package ``yoho_b200``; only ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may use it, and there
only as the checker or the timed CPU baseline, never as the thing shipped.

Parity status: **unpinned by the reference** for everything except the mel
filterbank.  The reference (JirkaKlimes/YOHO) ships zero tests and its hot path
imports jax/flax, which exist neither in the build container nor on the GPU
box, so the reference itself cannot be executed.  What we can and do pin:

* ``mel_filter_banks`` — the reference's own pure-NumPy function
  (yoho/src/preprocessing/mel_filterbanks.py:21-42) is importable; its output
  for (16000, 400, 80) and (16000, 400, 128) is committed bit-for-bit under
  ``tests/golden/`` by ``tests/golden/make_golden.py`` and every filterbank in
  this repo is compared against it with ``==``.
* the STFT restatement is checked against ``scipy.signal.stft`` (the routine
  ``jax.scipy.signal.stft`` mirrors) and the encoder restatement against
  ``transformers`` ``WhisperEncoder`` sub-modules driven by hand (see
  ``tests/test_oracle_witness.py``).
"""
