"""cuda_samples_b200 -- B200-native (sm_100a) data-parallel primitives behind the cuda-samples entry points.

histogram64 / histogram256, reduce<T>, scanExclusiveShort / scanExclusiveLarge, mergeSort (key-value):
hand-written CUDA kernels in csrc/, exported through a C ABI (include/*.h) as libb200prim.so.  This
package is the thin host-side mirror of that ABI.  No CPU fallback: importing fails if the library
has not been built (python cuda_samples_b200/build.py).
"""
from ._lib import B200Error, LIB_PATH, lib  # noqa: F401
from .api import *  # noqa: F401,F403
from . import api  # noqa: F401
