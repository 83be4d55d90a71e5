"""Batched many-stream chain: [blkconv] -> [stage a] -> [stage b] on S IQ streams sharing one filter
set (additive API over sfe_chain_* in include/sfe_gpu.h)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from ._lib import ChainDesc, SfeError, check, last_error, lib, _f32p

IN_I16, IN_U8, IN_F32 = 0, 1, 2
KLASS_NONE, KLASS_DECIMATE, KLASS_RESAMPLE = 0, 1, 2
FLAG_EXACT, FLAG_NO_FUSE, FLAG_NO_TENSOR, FLAG_NO_BANK = 1, 2, 4, 8
_NP_DT = {IN_I16: np.int16, IN_U8: np.uint8, IN_F32: np.float32}
BYTES_PER_SAMPLE = {IN_I16: 4, IN_U8: 2, IN_F32: 8}


def stage(klass: int, taps, up: int, blk: int, rate: float) -> dict:
    return dict(klass=klass, taps=taps, up=up, blk=blk, rate=rate)


def make_desc(n_streams: int, in_fmt: int, blkconv=None, stage_a=None, stage_b=None, flags: int = 0):
    """(sfe_chain_desc, keep-alive list of the tap arrays it points to)"""
    d = ChainDesc()
    d.in_fmt, d.n_streams, d.flags = int(in_fmt), int(n_streams), int(flags)
    keep = []
    if blkconv is not None:
        t = np.ascontiguousarray(np.asarray(blkconv[0], np.float32))
        keep.append(t)
        d.bc_taps, d.bc_ntaps, d.bc_fft = t.ctypes.data_as(_f32p), len(t), int(blkconv[1])
    for name, st in (("a", stage_a), ("b", stage_b)):
        if st is None:
            continue
        t = np.ascontiguousarray(np.asarray(st["taps"], np.float32))
        keep.append(t)
        s = getattr(d, name)
        s.klass, s.taps, s.n_taps = int(st["klass"]), t.ctypes.data_as(_f32p), len(t)
        s.up, s.blk, s.rate = int(st["up"]), int(st["blk"]), float(st["rate"])
    return d, keep


class RxFeed:
    """Streaming front door (sfe_rx_feed_*): device callback -> pinned ring -> GPU chain."""

    def __init__(self, ring_bytes: int, block_samples: int, blkconv=None, stage_a=None, stage_b=None, flags: int = 0):
        self._l = lib()
        d, keep = make_desc(1, IN_U8, blkconv, stage_a, stage_b, flags)
        self._h = self._l.sfe_rx_feed_create(C.byref(d), int(ring_bytes), int(block_samples))
        if not self._h:
            raise SfeError(f"sfe_rx_feed_create: {last_error()}")

    def callback(self, packet: bytes) -> int:
        """what the USB thread does: sfe_rx_feed_callback(buffer, length, feed)"""
        buf = (C.c_ubyte * len(packet)).from_buffer_copy(packet)
        return self._l.sfe_rx_feed_callback(C.addressof(buf), len(packet), self._h)

    def pump(self, cap: int, wait_ms: int = 0) -> np.ndarray:
        out = np.zeros(int(cap), np.complex64)
        n = self._l.sfe_rx_feed_pump(self._h, out.ctypes.data, int(cap), int(wait_ms))
        check(int(n), "sfe_rx_feed_pump")
        return out[:n]

    def stat(self, which: int) -> int:
        return int(self._l.sfe_rx_feed_stat(self._h, int(which)))

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            self._l.sfe_rx_feed_destroy(h)


class Chain:
    def __init__(self, n_streams: int, in_fmt: int = IN_I16, blkconv=None, stage_a=None, stage_b=None, flags: int = 0,
                 device: int | None = None):
        self._l = lib()
        if device is not None:
            check(self._l.sfe_set_device(int(device)), "sfe_set_device")
        d, keep = make_desc(n_streams, in_fmt, blkconv, stage_a, stage_b, flags)
        self._h = self._l.sfe_chain_create(C.byref(d))
        if not self._h:
            raise SfeError(f"sfe_chain_create: {last_error()}")
        self.n_streams, self.in_fmt = int(n_streams), int(in_fmt)

    @property
    def plan(self) -> str:
        return self._l.sfe_chain_plan(self._h).decode()

    def profile(self, on=True):
        """bracket the main kernels of every `on`-th process call with CUDA events (True/1 = every call, 0 = off)"""
        check(self._l.sfe_chain_profile_enable(self._h, int(on)), "sfe_chain_profile_enable")

    def profile_read(self):
        """[(op name, mean kernel ms, launches)] since profiling was enabled / last read"""
        out = []
        for i in range(self._l.sfe_chain_n_ops(self._h)):
            ms, n = C.c_float(0), C.c_int(0)
            check(self._l.sfe_chain_profile_read(self._h, i, C.byref(ms), C.byref(n)), "sfe_chain_profile_read")
            out.append((self._l.sfe_chain_op_name(self._h, i).decode(), float(ms.value), int(n.value)))
        return out

    def reset(self):
        check(self._l.sfe_chain_reset(self._h), "sfe_chain_reset")

    def max_out(self, n_in: int) -> int:
        return int(self._l.sfe_chain_max_out(self._h, int(n_in)))

    def process_dev(self, d_in: int, n_in: int, in_pitch: int, d_out: int, out_pitch: int, stream: int = 0) -> int:
        """Raw device pointers (ints); enqueues on `stream` (cudaStream_t as int) and returns n_out."""
        n_out = C.c_long(0)
        check(self._l.sfe_chain_process_dev(self._h, d_in, int(n_in), int(in_pitch), d_out, int(out_pitch),
                                            C.byref(n_out), stream), "sfe_chain_process_dev")
        return int(n_out.value)

    def process_host_raw(self, h_in: int, n_in: int, in_pitch: int, h_out: int, out_pitch: int) -> int:
        n_out = C.c_long(0)
        check(self._l.sfe_chain_process_host(self._h, h_in, int(n_in), int(in_pitch), h_out, int(out_pitch),
                                             C.byref(n_out)), "sfe_chain_process_host")
        return int(n_out.value)

    def process(self, iq: np.ndarray) -> np.ndarray:
        """iq: (S, N, 2) host array in the chain's input format -> complex64 (S, n_out)."""
        iq = np.ascontiguousarray(iq, dtype=_NP_DT[self.in_fmt])
        S, N, two = iq.shape
        if S != self.n_streams or two != 2:
            raise ValueError("iq must be (n_streams, N, 2)")
        cap = self.max_out(N)
        out = np.zeros((S, cap), np.complex64)
        n = self.process_host_raw(iq.ctypes.data, N, N, out.ctypes.data, cap)
        return out[:, :n]

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            self._l.sfe_chain_destroy(h)


class Box:
    """The batched chain over the GPUs of one box (sfe_box_* in include/sfe_gpu.h): contiguous stream ranges per
    device, one worker thread and one double-buffered copy pipeline per device, no collective."""

    def __init__(self, n_streams: int, in_fmt: int = IN_I16, blkconv=None, stage_a=None, stage_b=None, flags: int = 0,
                 n_devices: int = 0):
        self._l = lib()
        d, keep = make_desc(n_streams, in_fmt, blkconv, stage_a, stage_b, flags)
        self._h = self._l.sfe_box_create(C.byref(d), int(n_devices))
        if not self._h:
            raise SfeError(f"sfe_box_create: {last_error()}")
        self.n_streams, self.in_fmt = int(n_streams), int(in_fmt)
        self.n_devices = int(self._l.sfe_box_n_devices(self._h))

    @property
    def plan(self) -> str:
        return self._l.sfe_box_plan(self._h).decode()

    def stream_range(self, index: int):
        """(device ordinal, first stream, stream count) of shard `index`"""
        a, n = C.c_int(0), C.c_int(0)
        dev = check(self._l.sfe_box_stream_range(self._h, int(index), C.byref(a), C.byref(n)), "sfe_box_stream_range")
        return dev, int(a.value), int(n.value)

    def reset(self):
        check(self._l.sfe_box_reset(self._h), "sfe_box_reset")

    def sync(self):
        check(self._l.sfe_box_sync(self._h), "sfe_box_sync")

    def max_out(self, n_in: int) -> int:
        return int(self._l.sfe_box_max_out(self._h, int(n_in)))

    def process_host_raw(self, h_in: int, n_in: int, in_pitch: int, h_out: int, out_pitch: int) -> int:
        n_out = C.c_long(0)
        check(self._l.sfe_box_process_host(self._h, h_in, int(n_in), int(in_pitch), h_out, int(out_pitch), C.byref(n_out)),
              "sfe_box_process_host")
        return int(n_out.value)

    def process_dev(self, d_in, n_in: int, in_pitch: int, d_out, out_pitch: int, streams=None) -> int:
        """d_in / d_out / streams: one raw pointer (int) per device ordinal"""
        G = self.n_devices
        a = (C.c_void_p * G)(*[int(x) for x in d_in])
        o = (C.c_void_p * G)(*[int(x) for x in d_out])
        st = (C.c_void_p * G)(*[int(x) for x in streams]) if streams is not None else None
        n_out = C.c_long(0)
        check(self._l.sfe_box_process_dev(self._h, a, int(n_in), int(in_pitch), o, int(out_pitch), C.byref(n_out), st),
              "sfe_box_process_dev")
        return int(n_out.value)

    def process(self, iq: np.ndarray) -> np.ndarray:
        """iq: (S, N, 2) host array in the box's input format -> complex64 (S, n_out)."""
        iq = np.ascontiguousarray(iq, dtype=_NP_DT[self.in_fmt])
        S, N, two = iq.shape
        if S != self.n_streams or two != 2:
            raise ValueError("iq must be (n_streams, N, 2)")
        cap = max(self.max_out(N), 1)
        out = np.zeros((S, cap), np.complex64)
        n = self.process_host_raw(iq.ctypes.data, N, N, out.ctypes.data, cap)
        return out[:, :n]

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            self._l.sfe_box_destroy(h)
