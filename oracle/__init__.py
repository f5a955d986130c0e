"""CPU oracle for the b200rl hot path — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs
of ``bench.py`` may import anything from this package. The product (``areal_vllm_b200``) never
does: it has no CPU path and fails loudly when its CUDA library is missing.

What is here is a restatement of the reference's algorithms (each function cites the reference
file:line it follows), pinned against the reference itself: ``oracle/gen_golden.py`` imports the
real reference code from /root/reference in the build container, runs it on seeded inputs and
stores inputs + outputs under ``tests/golden/``; ``tests/test_oracle_golden.py`` checks the
restatement against those vectors on every run (no GPU, no /root/reference needed).
"""
