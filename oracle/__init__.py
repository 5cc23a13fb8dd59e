"""TEST INFRASTRUCTURE ONLY — the parity oracle for the tutoring-inference hot path.

Nothing under ``oracle/`` is part of the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / ``--impl reference``
legs may import it, and there only as the checker or the timed CPU baseline.

Parity pin: the reference ships no tests or golden vectors for this path
(SURVEY.md §4, §8c), so by the reference's own material parity is unpinned.
The pin we create: the reference's own modules (``tutoring_server.py``,
``lms_server.py``) imported unmodified in the build container under three
shims (``oracle/make_golden.py``), run in fp32 on CPU with fixed seeds on
``transformers 5.5.0`` / ``torch 2.11.0`` (the arithmetic's real home, a
third-party dependency the reference does not vendor or pin), with the
outputs committed under ``tests/golden/``.
"""
