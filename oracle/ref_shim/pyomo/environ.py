"""Stub of pyomo.environ (see package docstring).  Import lines served:
acinoset_misc.py:11,19  acinoset_opt.py:11-13  cheetah.py:6-7  links.py:8  utils.py:4-5  system.py:4-11."""
from math import sin, cos, atan, sqrt, exp, log, tan, asin, acos  # noqa: F401


class _Dummy:
    Skip = None

    def __init__(self, *a, **k):
        self.args, self.kwargs = a, k

    def __call__(self, *a, **k):
        return self

    def __getattr__(self, name):
        raise AttributeError(name)


class ConcreteModel(_Dummy):
    pass


class Var(_Dummy):
    pass


class Set(_Dummy):
    pass


class Param(_Dummy):
    pass


class RangeSet(_Dummy):
    pass


class Constraint(_Dummy):
    Skip = None


class ConstraintList(_Dummy):
    pass


class Objective(_Dummy):
    pass


class Suffix(_Dummy):
    IMPORT = 1
    EXPORT = 2
    IMPORT_EXPORT = 3


class SolverStatus:
    ok = "ok"


class TerminationCondition:
    optimal = "optimal"


def value(x):
    return float(x)


Reals = NonNegativeReals = Binary = Any = None
minimize = 1
maximize = -1
