"""CPU oracle for the speaker-embedding + scoring hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product:
only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline``
/ ``--impl reference`` legs may import it, and there only as the checker or as
the timed CPU baseline.  The product path (``speakerverification_b200``) never
imports this package and fails loudly when its CUDA library is missing.

Parity pinning: the reference (hiimmuc/SpeakerVerification) ships no tests,
golden vectors or fixtures for this path, so the oracle is
pinned against outputs of the reference's own Python code, imported from
/root/reference in the build container by ``oracle/make_golden.py`` and
committed as fixtures under ``tests/golden/``.
"""
