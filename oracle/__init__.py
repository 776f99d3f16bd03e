"""CPU oracle for the PTv3 hot path.

TEST INFRASTRUCTURE ONLY.  This package is a CPU restatement (numpy for the
integer work, torch-CPU fp32 for the floating point work) of the algorithms on
the reference hot path.  Only `tests/`, `__graft_entry__.smoke()` and the
`cpu_baseline` / `--impl reference` legs of `bench.py` may import it, and there
only as the checker or the timed CPU baseline -- never as the product path.
The product (`ptv3_b200`) never imports `oracle`.

Parity status: PINNED.  Every function below is checked against outputs of the
reference's own Python (imported from /root/reference in the build container,
see tests/golden/make_golden.py) frozen under tests/golden/*.npz, and against
the known-answer vectors listed in SURVEY.md section 8c.  Where the arithmetic
lives in third-party packages that are not vendored in the reference
(spconv-cu120==2.3.6, torch-scatter==2.1.2, flash-attn==2.6.3) the published
semantics are restated and that part of the parity is "unpinned by the
reference" (the reference has no tests at those boundaries): see DESIGN.md.
"""
