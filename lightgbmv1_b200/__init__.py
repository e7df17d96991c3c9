"""lightgbmv1_b200 -- B200-native (sm_100a) histogram tree learner behind the LightGBM C API.

The product is the shared library ``lib_lightgbm_b200.so`` (C ABI declared in ``include/``); this package is a
thin ctypes mirror of that ABI for tests and benchmarks.  There is no CPU fallback.
"""
from . import _native  # noqa: F401  (fails loudly when the CUDA library has not been built)
from ._native import CudaLearner, LgbCuError, make_feature_meta, make_tree_params  # noqa: F401
