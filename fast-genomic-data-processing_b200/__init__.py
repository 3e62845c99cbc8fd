"""B200-native PairHMM read-likelihood engine (drop-in for deepmutect/Mutect2Cpp's VectorLoglessPairHMM).

The product is ``libpairhmm_cuda.so`` (csrc/, C ABI in include/pairhmm_cuda.h). This package is the thin
Python host binding used by the tests and bench.py; it mirrors the reference's PairHMM interface
(M2/utils/pairhmm/PairHMM.h:13-69). There is NO CPU fallback: without the built library or without an
sm_100 GPU every compute call raises.
"""
from .pairhmm import (CudaPairHMM, Engine, PairHMMError, lib_path, load_library, device_count,  # noqa: F401
                      EXPORTED_SYMBOLS, prefix_plan, device_plan)
from . import synth  # noqa: F401
