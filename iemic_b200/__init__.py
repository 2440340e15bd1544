"""iemic_b200 -- B200-native hot path of I-EMIC (THCM assembly, CSR SpMV, GMRES orthogonalisation).

The product is the C-ABI shared library built from iemic_b200/csrc (see include/iemic_b200.h); this package only
holds the build recipe and a thin ctypes mirror of the reference's Ocean interface for tests and benchmarks.
"""
from .capi import Ocean, Config, B200Error, PAR, PAR_NAMES, MGS, CGS2, CGS2F, PREC_NONE, PREC_BLOCKJACOBI, PREC_CALLBACK, lib, LIB_PATH, EXPORTED  # noqa: F401
from . import grids  # noqa: F401
