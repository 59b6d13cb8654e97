"""epde_b200 -- B200-native (sm_100a) fitness-evaluation engine for EPDE's equation search.

Scope: the evaluator hot path only.  The numeric work lives in
``libepde_b200.so`` (csrc/epde_b200.cu, C ABI in include/epde_b200.h); this package is the thin
Python host mirroring the reference's cache / operator API.  Importing the package does not need a
GPU; creating an :class:`Engine` does, and fails loudly without one.
"""
from ._lib import EpdeB200Error, library_path, load as load_library  # noqa: F401
from .engine import (Batch, BatchResult, Encoding, Engine, encode, KIND_COS, KIND_INVERSE,  # noqa: F401
                     KIND_POWER, KIND_SIN, KS, MAX_FACTORS, MAX_TERMS)

__version__ = '0.1.0'
