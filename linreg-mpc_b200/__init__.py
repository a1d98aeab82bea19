"""linreg-mpc phase-1 secure aggregation on B200 (sm_100a).

``csrc/``  CUDA kernels + the C-ABI (include/p1gpu.h) -> ``libp1gpu.so``
``host/``  C mirror of the reference's run_party / run_trusted_initializer on that ABI
``ffi``    ctypes bindings of the C-ABI
``phase1`` host-side orchestration of a whole phase-1 run (all parties of one box)

There is no CPU fallback: loading fails loudly when libp1gpu.so is missing, and every
compute call fails when no B200 is visible.
"""
from . import ffi  # noqa: F401
from .ffi import P1Lib, build  # noqa: F401
from . import phase1  # noqa: F401
