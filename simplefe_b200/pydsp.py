"""pydsp -- drop-in for the reference's SWIG Python module (libdsp/test/pydsp.i:1-22).

The SWIG/numpy typemaps there (pydsp.i:16-18) collapse ``(float *taps, int n_taps)`` and
``(float *in, int n_in)`` into one float32 array argument each and turn ``(float *out, int out_len)``
into an int argument plus a returned array, so the reference scripts call

    r = pydsp.resample(taps, upsample, blksize)          # libdsp/test/test_resample.py:17
    n_out, y = r.process(x_block, out_len, rate)         # libdsp/test/test_resample.py:24
    d = pydsp.decimate(taps, upsample, blksize)          # libdsp/test/test_decimate.py:17

Same shapes here, served by the CUDA kernels behind the C ABI.  ``blkconv`` is not wrapped by the
reference's pydsp.i; it is provided as an additive extra with the C++ method names
(libdsp/blkconv.h:35-48).
"""
from __future__ import annotations

import numpy as np

from ._lib import SfeError, check, last_error, lib, _f32p


def _f32(a) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(a, dtype=np.float32))


class _polyphase:
    _create = ""

    def __init__(self, taps, upsample: int, blksize: int):
        t = _f32(taps)
        self._l = lib()
        self._h = getattr(self._l, self._create)(t.ctypes.data_as(_f32p), len(t), int(upsample), int(blksize))
        if not self._h:
            raise SfeError(f"{self._create}: {last_error()}")

    def process(self, x, out_len: int, rate: float):
        x = _f32(x)
        y = np.zeros(int(out_len), dtype=np.float32)
        n = self._l.sfe_polyphase_process(self._h, x.ctypes.data_as(_f32p), len(x), y.ctypes.data_as(_f32p),
                                          int(out_len), float(rate))
        check(n, "process")
        return n, y

    def set_exact(self, exact: bool = True):
        """Additive: reproduce the reference's operation order bit for bit (slower)."""
        check(self._l.sfe_polyphase_set_exact(self._h, int(bool(exact))), "set_exact")
        return self

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            self._l.sfe_polyphase_destroy(h)


class resample(_polyphase):
    """libdsp/resample.h:33-47"""
    _create = "sfe_resample_create"


class decimate(_polyphase):
    """libdsp/decimate.h:33-46"""
    _create = "sfe_decimate_create"


class blkconv:
    """libdsp/blkconv.h:35-48 (additive: not in the reference's pydsp.i)."""

    def __init__(self, taps, fft_len: int):
        t = _f32(taps)
        self._l = lib()
        self._h = self._l.sfe_blkconv_create(t.ctypes.data_as(_f32p), len(t), int(fft_len))
        if not self._h:
            raise SfeError(f"sfe_blkconv_create: {last_error()}")
        self._buf = np.ctypeslib.as_array(self._l.sfe_blkconv_buf(self._h), shape=(int(fft_len) + 2,))

    def get_blksize(self) -> int:
        return self._l.sfe_blkconv_blksize(self._h)

    def get_process_buf(self) -> np.ndarray:
        return self._buf

    def process(self) -> None:
        check(self._l.sfe_blkconv_process(self._h), "blkconv.process")

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            self._buf = None
            self._l.sfe_blkconv_destroy(h)
