"""CPU oracle for the cheetah trajectory-optimisation hot path.

TEST INFRASTRUCTURE -- NOT PRODUCT CODE.  Only `tests/`, `__graft_entry__.smoke()` and the
`cpu_baseline` / `--impl reference` legs of `bench.py` may import anything from this package,
and there only as the checker / the timed CPU baseline.  The product path
(`cheetah_pose_estimation_b200`) never imports it and fails loudly without its CUDA library.

What it is: a numpy FP64 restatement of the reference's arithmetic for the path named in
BASELINE.json (skeleton FK -> fisheye/pinhole projection -> likelihood-weighted redescending
reprojection cost + constant-acceleration motion cost -> Jacobians -> LM normal equations).
Each function cites the reference file:line it follows (paths relative to /root/reference).

Parity status: PINNED for FK, reduced pose, projection, robust loss, constraint functions and
cost assembly against (a) the reference's own functions imported in the build container
(oracle/make_golden.py, fixtures in tests/golden/) and (b) the shipped data/test_set golden
outputs.  UNPINNED for the optimiser itself: the reference solves with Pyomo + IPOPT/MA97, neither
of which is installed on this box; converged-cost parity is anchored on the shipped `q`/`obj_cost`
only up to the lost DLC-likelihood mask (SURVEY.md section 8c "What is NOT pinned").
"""
