"""CPU oracle for the DINCAE hot path -- TEST INFRASTRUCTURE ONLY.

This package restates, on the CPU, the algorithm of the reference
(gher-uliege/DINCAE.jl v2.0.6) for the path BASELINE.json's north_star names.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import it.  The product (dincae.jl_b200/) never does.

Pinning status
--------------
* interpnd / interp_adjn (src/points.jl:6-77): PINNED against the golden
  vectors of test/test_interp_adjoint.jl:20-47 and the adjoint identity
  test/test_interp_adjoint.jl:49-65 (tests/test_oracle_golden.py).
* SumSkip / CatSkip / normalize2: PINNED against test/test_utils.jl:4-13.
* Conv / MeanPool / Upsample / Adam / ClipGrad / glorot (Flux, NNlib,
  Optimisers.jl -- third-party, un-vendored, Manifest.toml git-ignored so no
  exact pin exists; compat Flux 0.15/0.16):  PARITY UNPINNED.  No test of the
  reference asserts a numeric value that flows through them and Julia is not
  installable here (no runtime, no network).  Their published semantics are
  restated (true convolution = flipped kernel, mean-pool divides by the full
  window incl. padding, bilinear align_corners=true, Adam eps outside the
  sqrt) and cross-checked between two independent restatements (torch ops in
  oracle/model.py versus plain loops in oracle/c/oracle.c).
"""
