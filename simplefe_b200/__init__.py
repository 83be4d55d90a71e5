"""simplefe_b200 -- B200 (sm_100a) implementation of simpleFE's libdsp sample-stream hot path.

Product surface:
  * ``libsfegpu.so``  -- C ABI (include/sfe_gpu.h), hand-written CUDA kernels, no CPU fallback
  * ``simplefe_b200.pydsp``  -- drop-in for the reference's SWIG module (libdsp/test/pydsp.i)
  * ``simplefe_b200.chain``  -- batched many-stream chain (additive API)
  * ``include/libdsp/*.h``   -- source-compatible C++ classes forwarding to the C ABI
"""
from ._lib import lib, last_error, SfeError  # noqa: F401

__all__ = ["lib", "last_error", "SfeError"]
