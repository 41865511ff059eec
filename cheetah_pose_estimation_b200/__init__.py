"""B200-native replacement of the cheetah trajectory-optimisation hot path (FK -> projection -> robust
reprojection + motion cost -> Jacobians -> block-banded LM solve).  The arithmetic lives in the CUDA library
`libcpe_b200.so` (C ABI: include/cpe.h); there is no CPU fallback."""
from . import skeleton  # noqa: F401

__all__ = ["skeleton", "engine", "synthetic"]
