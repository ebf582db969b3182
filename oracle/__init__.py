"""oracle/ -- TEST INFRASTRUCTURE ONLY (not product code).

A CPU restatement (torch-CPU fp32 / numpy) of the FI-ODE hot path named in
BASELINE.json:north_star.  Every function cites the reference file:line it
follows (paths relative to the reference checkout).  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl
reference`` legs may import this package -- as the checker or as the timed CPU
baseline, never as the thing shipped.  The product (``fiode_b200``) never
imports it and has no CPU fallback.

Parity pinning
--------------
The reference has no first-party tests or golden vectors for this path
(SURVEY.md section 4).  The oracle is therefore pinned against *outputs of the
reference itself*, produced by importing the reference's own modules in the
build container (``oracle/ref_harness.py`` -- needs the reference checkout,
which does not exist on the GPU box) and frozen under ``tests/golden/`` by
``tests/golden/make_golden.py``.  The one piece that cannot be pinned that way
is the ODE solver arithmetic: torchdiffeq==0.2.2 (reference ``env.yml:251``) is
an un-vendored third-party dependency absent from the reference tree and from
this image.  ``oracle/dopri5.py`` restates its published algorithm; it is
anchored on the reference's call sites (``models.py:236-241``,
``libs/core/core/dynamics/system_dynamics.py:48-52``), on the Dormand-Prince
order conditions, and on scipy's independent RK45 at tight tolerance.
=> solver parity: "parity unpinned" (restated, not diffed against an install).
"""
