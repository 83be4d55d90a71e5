"""oracle.py -- TEST INFRASTRUCTURE ONLY: ctypes front end of the CPU checkers.

Loads, in order of preference, ``oracle/_ref/libsfe_ref.so`` (the reference's own libdsp sources
compiled verbatim, kind "reference") or ``oracle/libsfe_oracle.so`` (the restatement in
sfe_oracle.c, kind "port").  Only tests/, ``__graft_entry__.smoke()`` and bench.py's
``cpu_baseline`` / ``--impl reference`` legs may import this module; nothing under
``simplefe_b200/`` does.

Classes mirror the reference's C++ classes (libdsp/blkconv.h:35-48, decimate.h:33-46,
resample.h:33-47) with the pydsp SWIG call shapes (libdsp/test/pydsp.i:16-18):
``n_out, y = obj.process(x_block, out_len, rate)``.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

_f32p = C.POINTER(C.c_float)


class _Stage(C.Structure):
    _fields_ = [("klass", C.c_int), ("taps", _f32p), ("n_taps", C.c_int), ("up", C.c_int),
                ("blk", C.c_int), ("rate", C.c_float)]


class _ChainCfg(C.Structure):
    _fields_ = [("in_fmt", C.c_int), ("bc_taps", _f32p), ("bc_ntaps", C.c_int), ("bc_fft", C.c_int),
                ("a", _Stage), ("b", _Stage)]


def _bind(lib):
    vp, ci, cf = C.c_void_p, C.c_int, C.c_float
    lib.sfeo_kind.restype = C.c_char_p
    lib.sfeo_build_flags.restype = C.c_char_p
    lib.sfeo_blkconv_new.restype = vp
    lib.sfeo_blkconv_new.argtypes = [_f32p, ci, ci]
    lib.sfeo_blkconv_blksize.argtypes = [vp]
    lib.sfeo_blkconv_buf.restype = _f32p
    lib.sfeo_blkconv_buf.argtypes = [vp]
    lib.sfeo_blkconv_process.argtypes = [vp]
    lib.sfeo_blkconv_free.argtypes = [vp]
    for k in ("decimate", "resample"):
        getattr(lib, f"sfeo_{k}_new").restype = vp
        getattr(lib, f"sfeo_{k}_new").argtypes = [_f32p, ci, ci, ci]
        getattr(lib, f"sfeo_{k}_process").argtypes = [vp, _f32p, ci, _f32p, ci, cf]
        getattr(lib, f"sfeo_{k}_free").argtypes = [vp]
    lib.sfeo_f32_to_dac10.argtypes = [_f32p, ci, C.c_void_p]
    if hasattr(lib, "sfeo_ref_tx_packets"):
        lib.sfeo_ref_tx_packets.restype = C.c_long
        lib.sfeo_ref_tx_packets.argtypes = [_f32p, ci, ci, ci, C.c_void_p, C.c_long]
    lib.sfeo_i16_to_f32.argtypes = [C.c_void_p, _f32p, C.c_size_t]
    lib.sfeo_u8_to_f32.argtypes = [C.c_void_p, _f32p, C.c_size_t]
    lib.sfeo_chain_run.argtypes = [C.POINTER(_ChainCfg), C.c_void_p, ci, C.c_long, _f32p, C.c_long,
                                   C.POINTER(ci), ci]
    lib.sfeo_ring_script.argtypes = [ci, C.POINTER(ci), ci, _f32p, ci]
    return lib


def build(force: bool = False) -> None:
    """Compile the checkers (port always; _ref only where /root/reference exists)."""
    args = ["make", "-C", _HERE, "all"] + (["-B"] if force else [])
    subprocess.run(args, check=True, stdout=subprocess.DEVNULL)


_cache: dict = {}


def load(kind: str = "auto", fast: bool = False):
    """kind: "reference" | "port" | "auto" (reference if present).  fast: the -O3 timing build."""
    key = (kind, fast)
    if key in _cache:
        return _cache[key]
    ref = os.path.join(_HERE, "_ref", "libsfe_ref_fast.so" if fast else "libsfe_ref.so")
    port = os.path.join(_HERE, "libsfe_oracle_fast.so" if fast else "libsfe_oracle.so")
    if kind in ("auto", "reference") and os.path.exists(ref):
        path = ref
    elif kind == "reference":
        raise FileNotFoundError(f"{ref} not built (needs /root/reference: make -C oracle ref)")
    else:
        if not os.path.exists(port):
            build()
        path = port
    lib = _bind(C.CDLL(path))
    _cache[key] = lib
    return lib


def _fp(a: np.ndarray):
    return a.ctypes.data_as(_f32p)


def _taps(t):
    return np.ascontiguousarray(np.asarray(t, dtype=np.float32))


class Blkconv:
    """libdsp/blkconv.h:35-48; the process buffer is exposed as a numpy view."""

    def __init__(self, taps, fft_len: int, lib=None):
        self._lib = lib or load()
        t = _taps(taps)
        self._h = self._lib.sfeo_blkconv_new(_fp(t), len(t), int(fft_len))
        self.blksize = self._lib.sfeo_blkconv_blksize(self._h)
        p = self._lib.sfeo_blkconv_buf(self._h)
        self.buf = np.ctypeslib.as_array(p, shape=(int(fft_len) + 2,))

    def get_blksize(self):
        return self.blksize

    def process(self):
        self._lib.sfeo_blkconv_process(self._h)

    def filter(self, x: np.ndarray) -> np.ndarray:
        """Stream x through the object block by block (zero fill of the last block)."""
        x = np.asarray(x, dtype=np.float32)
        y = np.empty_like(x)
        b = self.blksize
        for off in range(0, len(x), b):
            m = min(b, len(x) - off)
            self.buf[:m] = x[off:off + m]
            self.buf[m:b] = 0.0
            self.process()
            y[off:off + m] = self.buf[:m]
        return y

    def __del__(self):
        if getattr(self, "_h", None):
            self._lib.sfeo_blkconv_free(self._h)
            self._h = None


class _Poly:
    _name = ""

    def __init__(self, taps, upsample: int, blksize: int, lib=None):
        self._lib = lib or load()
        t = _taps(taps)
        self._h = getattr(self._lib, f"sfeo_{self._name}_new")(_fp(t), len(t), int(upsample), int(blksize))
        self._proc = getattr(self._lib, f"sfeo_{self._name}_process")
        self.blksize = int(blksize)

    def process(self, x, out_len: int, rate: float):
        x = np.ascontiguousarray(np.asarray(x, dtype=np.float32))
        y = np.zeros(int(out_len), dtype=np.float32)
        n = self._proc(self._h, _fp(x), len(x), _fp(y), int(out_len), float(rate))
        return n, y

    def stream(self, x, rate: float, blk: int | None = None) -> np.ndarray:
        """Feed x in calls of `blk` samples, out_len = floor(blk/rate)+2."""
        blk = blk or self.blksize
        out_len = int(np.floor(np.float32(blk) / np.float32(rate))) + 2
        outs = []
        for off in range(0, len(x), blk):
            n, y = self.process(x[off:off + blk], out_len, rate)
            outs.append(y[:n])
        return np.concatenate(outs) if outs else np.zeros(0, np.float32)

    def __del__(self):
        if getattr(self, "_h", None):
            getattr(self._lib, f"sfeo_{self._name}_free")(self._h)
            self._h = None


class Decimate(_Poly):
    """libdsp/decimate.h:33-46"""
    _name = "decimate"


class Resample(_Poly):
    """libdsp/resample.h:33-47"""
    _name = "resample"


def i16_to_f32(x: np.ndarray, lib=None) -> np.ndarray:
    lib = lib or load()
    x = np.ascontiguousarray(x, dtype=np.int16)
    y = np.empty(x.shape, np.float32)
    lib.sfeo_i16_to_f32(x.ctypes.data, _fp(y), x.size)
    return y


def u8_to_f32(x: np.ndarray, lib=None) -> np.ndarray:
    lib = lib or load()
    x = np.ascontiguousarray(x, dtype=np.uint8)
    y = np.empty(x.shape, np.float32)
    lib.sfeo_u8_to_f32(x.ctypes.data, _fp(y), x.size)
    return y


def f32_to_dac10(x: np.ndarray, lib=None) -> np.ndarray:
    """float samples -> 10-bit DAC words, 4 samples -> 5 bytes (examples/bpsk/bpsk.cxx:76-101; the reference's own
    convert_samples_to_bytes when oracle/_ref is loaded).  len(x) must be a multiple of 4."""
    lib = lib or load()
    x = np.ascontiguousarray(x, dtype=np.float32).ravel()
    assert len(x) % 4 == 0
    out = np.zeros(len(x) // 4 * 5, np.uint8)
    n = lib.sfeo_f32_to_dac10(_fp(x), len(x), out.ctypes.data)
    assert n == len(out)
    return out


def ref_tx_packets(x: np.ndarray, capacity: int, pkt: int, lib=None):
    """reference only: ring_buffer<float>(capacity) filled with x, drained by bpsk.cxx's tx_callback in packets of
    `pkt` bytes (bpsk.cxx:104-119); None when the loaded checker is the port"""
    lib = lib or load()
    if not hasattr(lib, "sfeo_ref_tx_packets"):
        return None
    x = np.ascontiguousarray(x, dtype=np.float32).ravel()
    out = np.zeros(len(x) // 4 * 5 + pkt, np.uint8)
    n = lib.sfeo_ref_tx_packets(_fp(x), len(x), int(capacity), int(pkt), out.ctypes.data, len(out))
    assert n >= 0
    return out[:n]


IN_I16, IN_U8, IN_F32 = 0, 1, 2
KLASS_NONE, KLASS_DECIMATE, KLASS_RESAMPLE = 0, 1, 2


def chain_run(iq: np.ndarray, in_fmt: int, blkconv=None, stage_a=None, stage_b=None,
              n_threads: int = 1, lib=None, out_cap: int | None = None):
    """Run [blkconv] -> [stage_a] -> [stage_b] on every stream of ``iq``.

    iq: (S, N, 2) int16 / uint8 / float32 according to in_fmt.
    blkconv: (taps, fft_len) or None.  stage_*: dict(klass=, taps=, up=, blk=, rate=) or None.
    Returns (out complex64 (S, n_out), n_out per stream).
    """
    lib = lib or load()
    dt = {IN_I16: np.int16, IN_U8: np.uint8, IN_F32: np.float32}[in_fmt]
    iq = np.ascontiguousarray(iq, dtype=dt)
    S, N, two = iq.shape
    assert two == 2
    cfg = _ChainCfg()
    cfg.in_fmt = in_fmt
    keep = []
    if blkconv is not None:
        t = _taps(blkconv[0])
        keep.append(t)
        cfg.bc_taps, cfg.bc_ntaps, cfg.bc_fft = _fp(t), len(t), int(blkconv[1])
    ratio = 1.0
    for name, st in (("a", stage_a), ("b", stage_b)):
        if st is None:
            continue
        t = _taps(st["taps"])
        keep.append(t)
        s = getattr(cfg, name)
        s.klass, s.taps, s.n_taps = int(st["klass"]), _fp(t), len(t)
        s.up, s.blk, s.rate = int(st["up"]), int(st["blk"]), float(st["rate"])
        ratio *= float(st["rate"])
    if out_cap is None:
        out_cap = int(N / ratio) + 64 + 4 * (N // 1024)
    out = np.zeros((S, out_cap, 2), np.float32)
    n_out = np.zeros(S, np.int32)
    rc = lib.sfeo_chain_run(C.byref(cfg), iq.ctypes.data, S, N, _fp(out), out_cap,
                            n_out.ctypes.data_as(C.POINTER(C.c_int)), int(n_threads))
    if rc != 0:
        raise RuntimeError("oracle chain_run: out_cap too small or I/Q count mismatch")
    n = int(n_out[0])
    assert (n_out == n).all()
    return out[:, :n, :].copy().view(np.complex64)[..., 0], n_out


def ring_script(capacity: int, ops, lib=None) -> np.ndarray:
    lib = lib or load()
    ops = np.ascontiguousarray(np.asarray(ops, dtype=np.int32).reshape(-1, 2))
    cap = 3 * len(ops) + int(ops[:, 1].sum()) + 16
    tr = np.zeros(cap, np.float32)
    n = lib.sfeo_ring_script(int(capacity), ops.ctypes.data_as(C.POINTER(C.c_int)), len(ops), _fp(tr), cap)
    assert n >= 0
    return tr[:n]
