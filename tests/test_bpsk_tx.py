"""Batched BPSK transmit path against the reference's own chain:
impulse train (examples/bpsk/bpsk.cxx:136-159) -> blkconv(RRC) (bpsk.cxx:125,162) -> DAC words (bpsk.cxx:76-101)."""
import numpy as np
import pytest

import fixtures as F


def _impulses(bits, sps, amp):
    x = np.zeros(len(bits) * sps, np.float32)
    x[::sps] = np.where(bits == 1, -np.float32(amp), np.float32(amp))
    return x


def _unpack_codes(b):
    """5 bytes -> four 10-bit codes (the inverse of bpsk.cxx:76-101), for counting code differences"""
    b = b.reshape(-1, 5).astype(np.int32)
    return np.stack([((b[:, 0] >> (2 * k)) & 3) << 8 | b[:, 1 + k] for k in range(4)], axis=1).ravel()


def test_select_bits_follows_the_reference_block_structure():
    """bpsk.cxx:136-159 with blk 1938, 10 samples per symbol: block 0 starts 194 symbols (the last one cut after
    8 samples), block 1 pads 2 zeros and starts 194 more ...; each block discards the rest of its last word."""
    from simplefe_b200.bpsk import pack_bits, select_bits
    rng = np.random.default_rng(7)
    words = rng.integers(0, 2 ** 31, size=64, dtype=np.uint64)       # rand() returns 31 bits
    bits = select_bits(words, 1938, 10, 3)
    # symbols started per block: ceil of the samples left after finishing the cut symbol, over 10
    left, want = 0, []
    for _ in range(3):
        n = 1938 - left
        k = -(-n // 10)
        want.append(k)
        left = k * 10 - n
    assert len(bits) == sum(want) and want == [194, 194, 194]
    # first block: 194 symbols = 6 whole words + 2 bits of the 7th; the second block starts with word 7
    w = [int(v) for v in words]
    exp0 = [(w[i // 32] >> (i % 32)) & 1 for i in range(194)]
    assert list(bits[:194]) == exp0
    assert list(bits[194:194 + 32]) == [(w[7] >> j) & 1 for j in range(32)]
    p = pack_bits(bits[None, :64])
    assert p.shape == (1, 2) and int(p[0, 0]) == sum(int(b) << j for j, b in enumerate(bits[:32]))


@pytest.mark.gpu
@pytest.mark.parametrize("tapname,sps,fft", [("rrc111", 10, 2048), ("rrc551", 50, 8192), ("rrc111", 7, 2048)])   # even sps: packed two-phase loop; odd: scalar
def test_bpsk_tx_vs_reference_chain(gpu_lib, oracle_lib, tapname, sps, fft):
    from oracle import oracle as O
    from simplefe_b200.bpsk import SCALING_FACTOR, TX_DAC10, TX_F32, BpskTx, select_bits
    taps = F.taps(tapname)
    C = 3
    rng = np.random.default_rng(11)
    blk = fft + 1 - len(taps)
    allbits = np.stack([select_bits(rng.integers(0, 2 ** 31, size=4000, dtype=np.uint64), blk, sps, (1136 * sps) // blk + 2) for _ in range(C)])
    n1, n2 = 1000, 136                                 # two calls: the waveform continues across them
    allbits = allbits[:, :n1 + n2]
    tx = BpskTx(taps, sps, C)
    y = np.concatenate([tx.process(allbits[:, :n1], TX_F32), tx.process(allbits[:, n1:], TX_F32)], axis=1)
    assert y.shape == (C, (n1 + n2) * sps)
    for c in range(C):
        ref = O.Blkconv(taps, fft, oracle_lib).filter(_impulses(allbits[c], sps, SCALING_FACTOR))[: y.shape[1]]
        nrms, peak = F.err_metrics(y[c], ref)
        assert nrms <= F.NRMS_TOL and peak <= F.PEAK_TOL, (c, nrms, peak)
    # fused DAC words.  (1) the kernel's packing of its own waveform is the reference's convert_samples_to_bytes
    # (bpsk.cxx:76-101, compiled into oracle/_ref) bit for bit; (2) against the words of the REFERENCE chain
    # (blkconv -> convert_samples_to_bytes): the two float waveforms differ by rounding (~1e-7), so a sample that sits
    # on a code boundary of (short)(x*511) may land one code apart -- counted, bounded, never more than one code
    tx.reset()
    d = np.concatenate([tx.process(allbits[:, :n1], TX_DAC10), tx.process(allbits[:, n1:], TX_DAC10)], axis=1)
    off_by_one = 0
    for c in range(C):
        assert np.array_equal(d[c], O.f32_to_dac10(y[c], lib=oracle_lib))
        ref = O.Blkconv(taps, fft, oracle_lib).filter(_impulses(allbits[c], sps, SCALING_FACTOR))[: y.shape[1]]
        want = O.f32_to_dac10(ref, lib=oracle_lib)
        dc = np.abs(_unpack_codes(d[c]) - _unpack_codes(want))
        assert dc.max() <= 1, "never more than one DAC code from the reference chain's words"
        off_by_one += int((dc == 1).sum())
    assert off_by_one <= 1e-3 * C * y.shape[1], off_by_one   # boundary cases: a handful in 10^4 samples
    # one call == two calls (state carry), and a fresh object reproduces it
    one = BpskTx(taps, sps, C).process(allbits, TX_F32)
    assert np.array_equal(one, y)


@pytest.mark.gpu
def test_bpsk_tx_arg_errors(gpu_lib):
    from simplefe_b200 import SfeError
    from simplefe_b200.bpsk import TX_DAC10, BpskTx
    with pytest.raises(SfeError):
        BpskTx(np.ones(10, np.float32), 0, 1)
    with pytest.raises(SfeError):
        BpskTx(np.ones(400, np.float32), 10, 1)        # more than 32 taps per phase
    tx = BpskTx(np.ones(5, np.float32), 3, 2)
    with pytest.raises(SfeError):
        tx.process(np.zeros((2, 5), np.uint8), TX_DAC10)   # 15 samples: not a multiple of 4
    assert tx.process(np.zeros((2, 0), np.uint8)).shape == (2, 0)
