"""disco4est_b200 -- B200-native matrix-free SIPG apply + Chebyshev smoother.

The product is the C-ABI shared library `libd4est_b200.so` (CUDA, sm_100a;
sources in disco4est_b200/csrc, interface in include/d4est_b200.h) and the
reference-side C adapter (disco4est_b200/adapter).  This Python package is the
test/bench harness over that ABI.
"""
from .abi import MeshDesc, ChebyParams, make_desc, PENALTY  # noqa: F401
from .runtime import Mesh, load_library, LIB_PATH, EXPORTS  # noqa: F401
