"""TEST INFRASTRUCTURE ONLY — CPU restatement of the ExoGround/TAN alignment hot path.

Nothing in the product package (`exoground_b200/`) may import this package.  The
only legitimate importers are `tests/`, `__graft_entry__.smoke()` and the
`cpu_baseline` / `--impl reference` legs of `bench.py`, and there only as the
checker or the timed CPU baseline, never as the thing shipped.

Parity status: the reference ships no golden vectors or known-answer tests for
this path.  The oracle is therefore pinned against outputs
of the reference modules themselves, run in the build container by
`oracle/make_golden.py` (which imports /root/reference) and committed as
fixtures under `tests/golden/`.  The one documented known-answer in the
reference (`circulant` docstring, train/loss.py:19-20) is checked as well.
"""
