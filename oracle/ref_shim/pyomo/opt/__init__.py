class SolverFactory:
    def __init__(self, *a, **k):
        raise RuntimeError("pyomo stub: no solver available (ipopt is not installed on this box)")


class SolverStatus:
    ok = "ok"


class TerminationCondition:
    optimal = "optimal"
