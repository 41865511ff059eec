"""Minimal stand-in for `pyomo` so the reference's *arithmetic* functions can be imported in
this container (pyomo / ipopt are not installed and there is no network).  TEST INFRASTRUCTURE
ONLY: used by oracle/make_golden.py to generate tests/golden/ fixtures from /root/reference.
Nothing here builds or solves a model; the math functions are plain `math.*`."""
