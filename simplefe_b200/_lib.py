"""ctypes binding of libsfegpu.so.  Loading fails loudly: there is no fallback implementation."""
from __future__ import annotations

import ctypes as C
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "libsfegpu.so")


class SfeError(RuntimeError):
    pass


class StageDesc(C.Structure):
    _fields_ = [("klass", C.c_int), ("taps", C.POINTER(C.c_float)), ("n_taps", C.c_int), ("up", C.c_int),
                ("blk", C.c_int), ("rate", C.c_float)]


class WalkStateC(C.Structure):
    """sfe_walk_state (include/sfe_gpu.h section 7): the scalars a decimate/resample object carries between calls"""
    _fields_ = [("pos", C.c_int), ("mu", C.c_float), ("held", C.c_int)]


class ChainDesc(C.Structure):
    _fields_ = [("in_fmt", C.c_int), ("n_streams", C.c_int), ("bc_taps", C.POINTER(C.c_float)),
                ("bc_ntaps", C.c_int), ("bc_fft", C.c_int), ("a", StageDesc), ("b", StageDesc), ("flags", C.c_int)]


RING_CONV = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_int)
RING_LEN = C.CFUNCTYPE(C.c_int, C.c_int)

# every symbol include/sfe_gpu.h declares: (restype, argtypes)
_f32p = C.POINTER(C.c_float)
_vp, _ci, _cl, _cf = C.c_void_p, C.c_int, C.c_long, C.c_float
SYMBOLS = {
    "sfe_version": (C.c_char_p, []),
    "sfe_last_error": (C.c_char_p, []),
    "sfe_device_count": (_ci, []),
    "sfe_set_device": (_ci, [_ci]),
    "sfe_kernel_launches": (C.c_int64, []),
    "sfe_pinned_alloc": (_vp, [C.c_size_t]),
    "sfe_pinned_free": (None, [_vp]),
    "sfe_dev_alloc": (_vp, [C.c_size_t]),
    "sfe_dev_free": (None, [_vp]),
    "sfe_memcpy_h2d": (_ci, [_vp, _vp, C.c_size_t, _vp]),
    "sfe_memcpy_d2h": (_ci, [_vp, _vp, C.c_size_t, _vp]),
    "sfe_stream_sync": (_ci, [_vp]),
    "sfe_blkconv_create": (_vp, [_f32p, _ci, _ci]),
    "sfe_blkconv_blksize": (_ci, [_vp]),
    "sfe_blkconv_buf": (_f32p, [_vp]),
    "sfe_blkconv_process": (_ci, [_vp]),
    "sfe_blkconv_destroy": (None, [_vp]),
    "sfe_decimate_create": (_vp, [_f32p, _ci, _ci, _ci]),
    "sfe_resample_create": (_vp, [_f32p, _ci, _ci, _ci]),
    "sfe_polyphase_process": (_ci, [_vp, _f32p, _ci, _f32p, _ci, _cf]),
    "sfe_polyphase_destroy": (None, [_vp]),
    "sfe_polyphase_set_exact": (_ci, [_vp, _ci]),
    "sfe_chain_create": (_vp, [C.POINTER(ChainDesc)]),
    "sfe_chain_destroy": (None, [_vp]),
    "sfe_chain_reset": (_ci, [_vp]),
    "sfe_chain_max_out": (_cl, [_vp, _cl]),
    "sfe_chain_process_dev": (_ci, [_vp, _vp, _cl, _cl, _vp, _cl, C.POINTER(_cl), _vp]),
    "sfe_chain_process_host": (_ci, [_vp, _vp, _cl, _cl, _vp, _cl, C.POINTER(_cl)]),
    "sfe_chain_plan": (C.c_char_p, [_vp]),
    "sfe_chain_n_ops": (_ci, [_vp]),
    "sfe_chain_op_name": (C.c_char_p, [_vp, _ci]),
    "sfe_chain_profile_enable": (_ci, [_vp, _ci]),
    "sfe_chain_profile_read": (_ci, [_vp, _ci, C.POINTER(C.c_float), C.POINTER(_ci)]),
    "sfe_box_create": (_vp, [C.POINTER(ChainDesc), _ci]),
    "sfe_box_destroy": (None, [_vp]),
    "sfe_box_n_devices": (_ci, [_vp]),
    "sfe_box_stream_range": (_ci, [_vp, _ci, C.POINTER(_ci), C.POINTER(_ci)]),
    "sfe_box_reset": (_ci, [_vp]),
    "sfe_box_max_out": (_cl, [_vp, _cl]),
    "sfe_box_plan": (C.c_char_p, [_vp]),
    "sfe_box_process_host": (_ci, [_vp, _vp, _cl, _cl, _vp, _cl, C.POINTER(_cl)]),
    "sfe_box_process_dev": (_ci, [_vp, C.POINTER(_vp), _cl, _cl, C.POINTER(_vp), _cl, C.POINTER(_cl), C.POINTER(_vp)]),
    "sfe_box_sync": (_ci, [_vp]),
    "sfe_pack_dac10_dev": (_ci, [_vp, _cl, _vp, _vp]),
    "sfe_polyphase_process_noalloc": (_ci, [_ci, _vp, _ci, _ci, C.POINTER(WalkStateC), _vp, _ci, _vp, _ci, _vp, _ci, _cf, _ci, _vp]),
    "sfe_bpsk_tx_create": (_vp, [_f32p, _ci, _ci, _cf, _ci]),
    "sfe_bpsk_tx_destroy": (None, [_vp]),
    "sfe_bpsk_tx_reset": (_ci, [_vp]),
    "sfe_bpsk_tx_process_dev": (_ci, [_vp, _vp, _cl, _cl, _vp, _cl, _ci, _vp]),
    "sfe_rx_feed_create": (_vp, [C.POINTER(ChainDesc), _ci, _ci]),
    "sfe_rx_feed_destroy": (None, [_vp]),
    "sfe_rx_feed_callback": (_ci, [_vp, _ci, _vp]),
    "sfe_rx_feed_pump": (_cl, [_vp, _vp, _cl, _ci]),
    "sfe_rx_feed_stat": (_cl, [_vp, _ci]),
    "sfe_tx_feed_create": (_vp, [_ci, _ci]),
    "sfe_tx_feed_destroy": (None, [_vp]),
    "sfe_tx_feed_write": (_cl, [_vp, _f32p, _cl, _ci]),
    "sfe_tx_feed_callback": (_ci, [_vp, _ci, _vp]),
    "sfe_tx_feed_stat": (_cl, [_vp, _ci]),
    "sfe_ring_create": (_vp, [_ci, _ci]),
    "sfe_ring_alloc": (_ci, [_vp, _ci]),
    "sfe_ring_destroy": (None, [_vp]),
    "sfe_ring_space": (_ci, [_vp]),
    "sfe_ring_count": (_ci, [_vp]),
    "sfe_ring_write": (_ci, [_vp, _vp, _ci]),
    "sfe_ring_read": (_ci, [_vp, _vp, C.c_uint, RING_CONV, RING_LEN]),
    "sfe_ring_read_dev": (_ci, [_vp, _vp, _ci, _vp]),
}

_lib = None


def lib():
    """The loaded library.  Raises if libsfegpu.so has not been built (python -m simplefe_b200.build)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise SfeError(f"{LIB_PATH} is missing: build it with `python -m simplefe_b200.build` "
                           "(the GPU library is the only implementation; there is no CPU fallback)")
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(l, name)  # AttributeError if the ABI and the header diverge
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def last_error() -> str:
    return lib().sfe_last_error().decode()


def check(rc: int, what: str) -> int:
    if rc < 0:
        raise SfeError(f"{what} failed ({rc}): {last_error()}")
    return rc
