"""Batched BPSK transmit path (additive; include/sfe_gpu.h section 4): what examples/bpsk/bpsk.cxx:122-174
does for one channel -- bits -> impulses -> RRC blkconv -> float samples / 10-bit DAC words -- for many."""
from __future__ import annotations

import numpy as np

from ._lib import SfeError, check, last_error, lib, _f32p

TX_F32, TX_DAC10 = 0, 1
SCALING_FACTOR = float(np.float32(0.85) / np.float32(1.35))   # bpsk.cxx:47


def select_bits(words: np.ndarray, blk_size: int, sps: int, n_blocks: int) -> np.ndarray:
    """The symbol bits the reference's producer loop sends for a sequence of rand() words (bpsk.cxx:136-159):
    every block of `blk_size` samples first finishes the symbol cut by the previous block, then draws fresh
    words and uses their bits LSB first -- the unused bits of a block's last word are thrown away.
    Host-side integer bookkeeping (the reference's caller code), returned as one uint8 bit per symbol."""
    words = np.asarray(words, dtype=np.uint64)
    out, wi, n_phase = [], 0, 0
    for _ in range(n_blocks):
        n_input = 0
        if n_phase > 0:
            n_input += sps - n_phase
            n_phase = 0
        while n_input < blk_size:
            word = int(words[wi])
            wi += 1
            j = 0
            while j < 32 and n_input < blk_size:
                out.append((word >> j) & 1)
                n_input += 1
                n_phase = 1
                while n_phase < sps and n_input < blk_size:
                    n_input += 1
                    n_phase += 1
                j += 1
        if n_phase == sps:
            n_phase = 0
    return np.array(out, dtype=np.uint8)


def pack_bits(bits: np.ndarray) -> np.ndarray:
    """(C, n_sym) 0/1 -> (C, ceil(n_sym/32)) uint32, bit j of word w = symbol 32 w + j"""
    bits = np.atleast_2d(np.asarray(bits, dtype=np.uint8))
    C, n = bits.shape
    pad = (-n) % 32
    b = np.pad(bits, ((0, 0), (0, pad))).reshape(C, -1, 32).astype(np.uint64)
    return (b << np.arange(32, dtype=np.uint64)).sum(axis=2).astype(np.uint32)


class BpskTx:
    def __init__(self, taps, sps: int, n_channels: int, amp: float = SCALING_FACTOR, device: int | None = None):
        self._l = lib()
        if device is not None:
            check(self._l.sfe_set_device(int(device)), "sfe_set_device")
        t = np.ascontiguousarray(np.asarray(taps, np.float32))
        self._h = self._l.sfe_bpsk_tx_create(t.ctypes.data_as(_f32p), len(t), int(sps), float(amp), int(n_channels))
        if not self._h:
            raise SfeError(f"sfe_bpsk_tx_create: {last_error()}")
        self.sps, self.n_channels = int(sps), int(n_channels)

    def reset(self):
        check(self._l.sfe_bpsk_tx_reset(self._h), "sfe_bpsk_tx_reset")

    def process_dev(self, d_bits: int, n_sym: int, pitch_words: int, d_out: int, out_pitch_bytes: int, out_fmt: int = TX_F32,
                    stream: int = 0):
        check(self._l.sfe_bpsk_tx_process_dev(self._h, d_bits, int(n_sym), int(pitch_words), d_out, int(out_pitch_bytes),
                                              int(out_fmt), stream), "sfe_bpsk_tx_process_dev")

    def process(self, bits: np.ndarray, out_fmt: int = TX_F32) -> np.ndarray:
        """bits: (C, n_sym) 0/1 array -> (C, n_sym*sps) float32 or (C, n_sym*sps/4*5) uint8 (host convenience)."""
        bits = np.atleast_2d(bits)
        C, n_sym = bits.shape
        if C != self.n_channels:
            raise ValueError("bits must be (n_channels, n_sym)")
        words = np.ascontiguousarray(pack_bits(bits))
        n_samp = n_sym * self.sps
        row = n_samp * 4 if out_fmt == TX_F32 else n_samp // 4 * 5
        pitch = (row + 15) // 16 * 16
        l = self._l
        d_bits, d_out = l.sfe_dev_alloc(max(words.nbytes, 4)), l.sfe_dev_alloc(max(pitch * C, 16))
        if not d_bits or not d_out:
            raise SfeError(f"sfe_dev_alloc: {last_error()}")
        try:
            check(l.sfe_memcpy_h2d(d_bits, words.ctypes.data, words.nbytes, None), "h2d")
            self.process_dev(d_bits, n_sym, words.shape[1], d_out, pitch, out_fmt)
            raw = np.zeros((C, pitch), np.uint8)
            check(l.sfe_memcpy_d2h(raw.ctypes.data, d_out, raw.nbytes, None), "d2h")
            check(l.sfe_stream_sync(None), "sync")
        finally:
            l.sfe_dev_free(d_bits)
            l.sfe_dev_free(d_out)
        raw = raw[:, :row]
        return np.ascontiguousarray(raw).view(np.float32) if out_fmt == TX_F32 else raw

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            self._l.sfe_bpsk_tx_destroy(h)
