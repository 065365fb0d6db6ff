"""B200-native explicit time-step core for DynEarthSol (3D tets, FP64, sm_100a).

The product is `csrc/libdes_b200.so` (C-ABI: include/des_gpu.h); `core` is the ctypes
host-side mirror.  There is no CPU fallback."""
from . import core  # noqa: F401
from .core import DesGpu, DesParam, DesError, make_param  # noqa: F401
