from math import log, exp, sin, cos, atan, sqrt  # noqa: F401
