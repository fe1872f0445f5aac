"""CPU oracle for the TMDiff sampling hot path.

TEST INFRASTRUCTURE ONLY.  This package is a torch-fp32 CPU restatement of the
reference algorithm (WavBEST denoiser + DPM-Solver++ sampler).  It is the
checker for the CUDA path and the timed CPU baseline; nothing under
``tmdiff_b200/`` may import it.  Only ``tests/``, ``__graft_entry__.smoke()``
and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs use it.

Parity pin: every function here is checked against the *unmodified* reference
imported from /root/reference by ``tests/golden/make_golden.py`` (run in the
build container, where the reference is mounted); the resulting vectors are
committed under ``tests/golden/`` and re-checked by ``tests/test_oracle.py``.
The reference itself ships no tests or golden vectors.
"""
