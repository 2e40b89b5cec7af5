"""TEST INFRASTRUCTURE ONLY — CPU oracle for the ObjCAViT / AdaBins depth head.

Nothing under ``oracle/`` is product code. Only ``tests/``, ``__graft_entry__.smoke()``
and the ``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import it, and
there only as the checker (or as the timed CPU baseline), never as the thing shipped.

* ``oracle.head``      — CPU restatement (plain torch fp32 functional ops + explicit loops for the
                          index-mapping pieces) of the reference head, each function citing the
                          reference ``file:line`` it follows.
* ``oracle.ref_shim``  — imports the *unmodified* reference modules from ``/root/reference`` (only in
                          the build container; the GPU box has no ``/root/reference``).
* ``oracle.make_golden`` — runs the real reference on seeded inputs and writes ``tests/golden/``.
* ``oracle.synth``     — the seeded synthetic inputs / weights of SURVEY.md §8(d).

Parity status: PINNED for a1–a7 (restatement == real reference modules run in the build container,
see ``tests/golden/*.pt`` and ``oracle/make_golden.py``); the chamfer part of a8 is UNPINNED
(pytorch3d 0.7.0 is not vendored by the reference and not installed; restated from its documented
defaults and pinned only against a brute-force O(P·T) formula).
"""
