"""CPU oracle for the run-analysis hot path.  TEST INFRASTRUCTURE ONLY.

This package restates, on the CPU, the arithmetic that the reference
(Siryoos/kserve-vllm-mini: analyze.py, energy/collector.py, cost_estimator.py,
cost_calculator.py) performs on a run directory.  It exists so the CUDA path can
be checked; it is never the thing that ships.

Who may import it: `tests/`, `__graft_entry__.smoke()` and the `cpu_baseline` /
`--impl reference` legs of `bench.py`.  Nothing under `kserve-vllm-mini_b200/` or
the root-level drop-in scripts may import it (tests/test_boundary.py greps for
exactly that).

Modules
  py_port   pure-Python restatement in the reference's own evaluation order
            (list.sort, sequential `+=`, statistics.mean).  Columnar inputs
            instead of the reference's list-of-dicts, otherwise step for step.
  np_port   numpy-vectorised restatement for sizes py_port cannot reach; proven
            bit-equal to py_port on order statistics / counts in tests/test_oracle.py.

Parity pinning: PINNED for every function that exists in the reference - py_port is
checked against the reference's own known-answer tests (tests/test_energy.py,
.github/workflows/lint-test.yml:112-136) and against golden fixtures produced by
running the unmodified reference in the build container
(tests/golden/make_golden.py -> tests/golden/*.json).
UNPINNED (no reference function exists): the 1-second
window series (`window_series`) and the per-interval energy join (`interval_energy`);
their oracle is the numpy restatement of the spec in DESIGN.md, and the per-interval
join is anchored on the reference's `trapezoidal_wh` for the one-interval case.
"""
