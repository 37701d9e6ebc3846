"""CPU oracle for the CFG sampling hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product:
only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import it, and there only as the checker or as the
timed CPU baseline.  The shipped path (``diffusion-model_b200/``) never imports
this package and raises if the CUDA library is missing.

Parity status: the reference ships no tests, golden vectors, fixtures or
weights, so parity is *unpinned by the reference's own
artefacts*.  It is pinned instead against outputs of the reference itself run
in the build container (torch 2.11.0 CPU fp32): ``tests/golden/make_golden.py``
imports ``/root/reference`` unmodified, runs it on seeded inputs and commits the
results under ``tests/golden/``; ``tests/test_oracle_golden.py`` checks this
restatement against those vectors.
"""
