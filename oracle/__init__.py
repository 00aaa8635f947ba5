"""TEST INFRASTRUCTURE ONLY — CPU oracle for the FastSpeech2 + HiFi-GAN + GMM-prior hot path.

Nothing under ``oracle/`` is part of the product.  Only ``tests/``, ``__graft_entry__.smoke()``
and the ``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import it, and only as
the checker or as the timed CPU baseline.  The shipped path (package
``mid-attribute-speaker-generation_b200``) never imports this package and has no CPU fallback.

Parity status: the reference ships NO tests, golden vectors or fixtures, so the
reference's own tests pin nothing ("parity unpinned" by reference tests).  ``oracle/restate.py``
is instead pinned against *outputs of the reference itself* run in the build container
(``oracle/ref_import.py`` imports /root/reference read-only; ``oracle/gen_golden.py`` writes
``tests/golden/*.npz``; ``tests/test_oracle_golden.py`` checks the restatement against them and,
when /root/reference is present, against the live reference modules).
"""
