"""CPU oracle for the video-flow-ml optical-flow hot path.

TEST INFRASTRUCTURE ONLY.  Nothing under ``oracle/`` is part of the product:
only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` may import it, and only as the
checker / the timed CPU baseline.  The product package
(``video_flow_ml_b200``) never imports this package and has no CPU fallback.

The modules here restate, on the CPU (torch fp32 ops + numpy), the algorithms
of the reference's hot path.  Every function cites the reference file:line it
follows (paths relative to the reference checkout).  Parity of these
restatements is pinned by ``tests/golden/*.npz``: outputs of the *unmodified*
reference modules, executed in the build container by
``tests/golden/make_golden.py`` (import shims for the absent ``timm``/``yacs``
packages only), on seeded inputs and seeded synthetic weights.

Unpinned boundary: the Twins-SVT-L encoder lives in third-party
``timm==0.4.12`` (not in the reference tree, not installable offline).
``oracle/twins.py`` restates it from the vendored class definitions in
``VideoFlow/core/Networks/twins_ft.py``; no reference test or fixture pins its
outputs, so the encoder's parity is "unpinned" and encoder outputs are
treated as the *inputs* of the graded hot path.
"""
