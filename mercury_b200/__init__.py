"""mercury_b200 -- B200-native importance-sampling hot path of AIoT-MLSys-Lab/Mercury.

Python/PyTorch surface over a C-ABI CUDA library (include/mercury_b200.h).  The kernels are
hand-written for sm_100a; there is no Triton, no multi-backend dispatch and no CPU fallback.
"""
from . import _lib  # noqa: F401  (does not load the .so until an op is called)
from . import ops  # noqa: F401

__all__ = ["ops"]
__version__ = "0.1.0"
