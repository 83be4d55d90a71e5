"""GPU parity tests (-m gpu): the CUDA path, called through the C ABI (ctypes), against the CPU oracle
on identical seeded inputs, against the committed golden fixtures, and through size-independent
properties (chunk invariance, linearity).  Tolerances are north_star's: conversion and indexing
bit-exact; float32 outputs nRMS <= 1e-5 and peak-normalised max error <= 1e-4."""
import os

import numpy as np
import pytest

import fixtures as F

pytestmark = pytest.mark.gpu

KNOWN = dict(np.load(os.path.join(F.GOLDEN, "known.npz")))
CHAIN = dict(np.load(os.path.join(F.GOLDEN, "chain.npz")))


def _sine():
    return np.sin(0.02 * np.pi * np.arange(1024)).astype(np.float32)


# ---- single-object drop-in API (pydsp call shapes) -------------------------------------------------

def test_blkconv_known_answer(gpu_lib):
    from simplefe_b200 import pydsp
    bc = pydsp.blkconv([1, 1, 1, 1, 1], 32)          # libdsp/test/test_blkconv.cxx:6-34
    assert bc.get_blksize() == 28
    buf = bc.get_process_buf()
    buf[:28] = 1.0
    bc.process()
    np.testing.assert_allclose(buf[:28], [1, 2, 3, 4] + [5] * 24, atol=1e-5)
    buf[:28] = 0.0
    bc.process()
    np.testing.assert_allclose(buf[:28], [4, 3, 2, 1] + [0] * 24, atol=1e-5)


@pytest.mark.parametrize("taps,fft", [("lp255", 4096), ("rrc111", 2048), ("rrc551", 8192), ("taps31", 64), ("taps31", 128),
                                      ("taps31", 256), ("rrc111", 512), ("rrc111", 1024)])
def test_blkconv_object_vs_oracle(gpu_lib, oracle_lib, taps, fft):
    from oracle import oracle as O
    from simplefe_b200 import pydsp
    t = F.taps(taps)
    rng = np.random.default_rng(11)
    bc = pydsp.blkconv(t, fft)
    blk = bc.get_blksize()
    x = rng.normal(0, 0.25, 5 * blk + 17).astype(np.float32)
    ref = O.Blkconv(t, fft, lib=oracle_lib).filter(x)
    got = np.empty_like(x)
    buf = bc.get_process_buf()
    for off in range(0, len(x), blk):
        m = min(blk, len(x) - off)
        buf[:m] = x[off:off + m]
        buf[m:blk] = 0.0
        bc.process()
        got[off:off + m] = buf[:m]
    nrms, peak = F.err_metrics(got, ref)
    assert nrms <= F.NRMS_TOL and peak <= F.PEAK_TOL, (nrms, peak)


@pytest.mark.parametrize("name,cls,rate", [("resample_077", "resample", 0.77), ("resample_177", "resample", 1.77),
                                            ("decimate_177", "decimate", 1.77), ("decimate_8", "decimate", 8.0)])
@pytest.mark.parametrize("exact", [False, True])
def test_polyphase_objects_known_answers(gpu_lib, name, cls, rate, exact):
    from simplefe_b200 import pydsp
    x0 = _sine()
    o = getattr(pydsp, cls)(F.taps("taps31"), 4, 128)   # libdsp/test/test_resample.py:17
    if exact:
        o.set_exact(True)
    ys, cnt = [], []
    for b in range(8):
        n, y = o.process(x0[b * 128:(b + 1) * 128], 4 * 128, rate)   # test_resample.py:24
        assert y.shape == (512,)
        ys.append(y[:n])
        cnt.append(n)
    y = np.concatenate(ys)
    assert cnt == list(KNOWN[name + "_counts"]), "output counts per call are bit-exact"
    ref = KNOWN[name + "_y"]
    if exact:
        assert np.array_equal(y, ref), "exact mode reproduces the reference bit for bit"
    else:
        nrms, peak = F.err_metrics(y, ref)
        assert nrms <= F.NRMS_TOL and peak <= F.PEAK_TOL, (nrms, peak)


def test_polyphase_arg_errors(gpu_lib, capfd):
    from simplefe_b200 import pydsp
    d = pydsp.decimate(F.taps("taps31"), 4, 128)
    x = np.zeros(128, np.float32)
    assert d.process(x, 512, 0.5)[0] == 0                              # decimate.cxx:75-78
    assert d.process(np.zeros(129, np.float32), 512, 2.0)[0] == 0      # decimate.cxx:79-82
    assert d.process(x, 10, 2.0)[0] == 0                               # decimate.cxx:84-87
    r = pydsp.resample(F.taps("taps31"), 4, 128)
    assert r.process(x, 512, 0.2)[0] == 0                              # resample.cxx:91-94
    out = capfd.readouterr().out
    assert "rate should be larger than 1.0" in out and "output buffer is not large enough" in out


@pytest.mark.parametrize("cls,up,rate,blk,tapname", [("decimate", 1, 8.0, 4096, "dec128"), ("decimate", 3, 208.33333, 1000, "rs512_l3"),
                                                     ("decimate", 4, 2.5003, 777, "taps31"), ("resample", 4, 0.77, 128, "taps31"),
                                                     ("resample", 24, 26.041666, 4096, "rs512_l24"), ("resample", 4, 1.77, 7, "taps31")])
def test_polyphase_objects_random_vs_oracle(gpu_lib, oracle_lib, cls, up, rate, blk, tapname):
    """ragged block sizes, fractional rates (mu != 0, held-over outputs), exact mode -> bit-exact"""
    from oracle import oracle as O
    from simplefe_b200 import pydsp
    rng = np.random.default_rng(5)
    x = rng.normal(0, 0.3, 20011).astype(np.float32)
    t = F.taps(tapname)
    ref = getattr(O, cls.capitalize())(t, up, blk, lib=oracle_lib).stream(x, rate)
    g = getattr(pydsp, cls)(t, up, blk).set_exact(True)
    out_len = int(np.floor(np.float32(blk) / np.float32(rate))) + 2
    ys = []
    for off in range(0, len(x), blk):
        n, y = g.process(x[off:off + blk], out_len, rate)
        ys.append(y[:n])
    got = np.concatenate(ys)
    assert len(got) == len(ref)
    assert np.array_equal(got, ref)


# ---- batched chain API ----------------------------------------------------------------------------

def _chain(cfg, flags=0, S=None):
    from simplefe_b200.chain import Chain
    return Chain(S or cfg["S"], cfg["in_fmt"], cfg.get("blkconv"), cfg.get("stage_a"), cfg.get("stage_b"), flags=flags)


@pytest.mark.parametrize("name", ["cfg1_small", "cfg2_small", "cfg3_small", "cfg4_small", "cfg5_small"])
@pytest.mark.parametrize("flags", [0, 8, 8 + 2, 1])
def test_chain_vs_golden(gpu_lib, name, flags):
    """small cases of every BASELINE config against the committed reference outputs (tests/golden/chain.npz);
    flags: 0 = default plan (one fused polyphase bank where the schedules are integral), 8 = NO_BANK (staged plan,
    blkconv folded into the decimator), 8+2 = NO_BANK|NO_FUSE (FFT blkconv + separate stages), 1 = EXACT"""
    cfg = F.CONFIGS[name]
    iq = F.make_input(cfg)
    ch = _chain(cfg, flags)
    got = ch.process(iq)
    ref = CHAIN[name]
    assert got.shape == ref.shape, (got.shape, ref.shape, ch.plan)
    nrms, peak = F.err_metrics(got, ref)
    assert nrms <= F.NRMS_TOL and peak <= F.PEAK_TOL, (ch.plan, nrms, peak)
    if flags == 1 and cfg.get("blkconv") is None:
        assert np.array_equal(got, ref), "EXACT plan without FFT stage is bit-exact"


@pytest.mark.parametrize("name", ["cfg2_small", "cfg5_small", "cfg1_small"])
def test_chain_chunk_invariance(gpu_lib, name):
    """integral-step configs: any split of the stream into calls gives the same concatenated output"""
    cfg = F.CONFIGS[name]
    iq = F.make_input(cfg)
    one = _chain(cfg).process(iq)
    ch = _chain(cfg)
    parts, off = [], 0
    for n in (1, 7, 1000, 4096, 12345, 3, 10 ** 9):
        n = min(n, cfg["N"] - off)
        if n <= 0:
            break
        parts.append(ch.process(iq[:, off:off + n]))
        off += n
    many = np.concatenate(parts, axis=1)
    assert many.shape == one.shape
    nrms, peak = F.err_metrics(many, one)
    assert nrms <= 2e-6 and peak <= 1e-5, (nrms, peak)


def test_chain_int16_conversion_bit_exact(gpu_lib):
    """a 1-tap unit filter exposes the in-kernel conversion: must equal (float)s * (1.0f/32767.0f) exactly"""
    from simplefe_b200.chain import Chain
    s = np.arange(-32768, 32768, dtype=np.int32).astype(np.int16)
    iq = np.stack([s, s[::-1]], axis=1)[None]
    ch = Chain(1, F.IN_I16, None, dict(klass=1, taps=np.array([1.0], np.float32), up=1, blk=4096, rate=1.0), flags=1)
    got = ch.process(iq)[0]
    want = (iq[0].astype(np.float32) * np.float32(1.0 / 32767.0)).view(np.complex64)[:, 0]
    # rate 1.0: the last sample's interpolation partner is in the next call, so the reference holds it over
    assert len(got) == len(want) - 1 and np.array_equal(got, want[:-1])
    u = np.arange(256, dtype=np.uint8)
    iq8 = np.stack([u, u[::-1]], axis=1)[None]
    ch8 = Chain(1, F.IN_U8, None, dict(klass=1, taps=np.array([1.0], np.float32), up=1, blk=4096, rate=1.0), flags=1)
    want8 = ((iq8[0].astype(np.int32) - 128).astype(np.float32) * np.float32(1.0 / 127.0)).view(np.complex64)[:, 0]
    assert np.array_equal(ch8.process(iq8)[0], want8[:-1])


def test_chain_linearity_full_size_stream(gpu_lib):
    """size-independent property at BASELINE length (2^20 samples/stream): chain(a) + chain(b) == chain(a+b)"""
    cfg = dict(F.CONFIGS["cfg5"], S=2)
    rng = np.random.default_rng(99)
    a = rng.integers(-8000, 8000, size=(2, cfg["N"], 2)).astype(np.int16)
    b = rng.integers(-8000, 8000, size=(2, cfg["N"], 2)).astype(np.int16)
    ya, yb, yab = (_chain(cfg, S=2).process(v) for v in (a, b, (a + b).astype(np.int16)))
    assert ya.shape[1] == 5034 or ya.shape[1] == 5033
    nrms, peak = F.err_metrics(ya + yb, yab)
    assert nrms <= 2e-6 and peak <= 1e-5, (nrms, peak)


@pytest.mark.parametrize("name,S,fmt,flags", [("cfg1", 1, None, 0), ("cfg2", 1, None, 0), ("cfg3", 1, None, 0), ("cfg4", 1, None, 0),
                                              ("cfg5", 48, None, 0), ("cfg5", 40, F.IN_U8, 0), ("cfg5", 40, None, 8), ("cfg2", 40, None, 0),
                                              ("cfg1", 3, F.IN_I16, 0)])
def test_baseline_configs_at_full_size_vs_oracle(gpu_lib, oracle_lib, name, S, fmt, flags):
    """every BASELINE config at the length it is worded on (2^20 / 10^7 samples per stream; cfg 5 on 48 of its 1024
    streams -- a ragged last group of 32 -- the oracle needs ~0.1 s per stream), fed in one call, against the
    reference classes: same output counts, errors within north_star's tolerances"""
    from oracle import oracle as O
    cfg = dict(F.CONFIGS[name], S=S)
    if fmt is not None:          # the device's u8 wire format / int16 rows into the FFT convolver
        cfg["in_fmt"] = fmt
    if name == "cfg2" and S > 1:
        cfg["N"] = 1 << 20       # the many-streams flavour of cfg 2 (tcgen05 Toeplitz plan) as bench_stages.py times it
    iq = F.make_input(cfg)
    ref, _ = O.chain_run(iq, cfg["in_fmt"], cfg.get("blkconv"), cfg.get("stage_a"), cfg.get("stage_b"), lib=oracle_lib, n_threads=8)
    ch = _chain(cfg, flags)      # flags 8 = NO_BANK: fused blkconv*decimate on the tensor cores + generic resampler
    got = ch.process(iq)
    assert got.shape == ref.shape, (ch.plan, got.shape, ref.shape)
    nrms, peak = F.err_metrics(got, ref)
    assert nrms <= F.NRMS_TOL and peak <= F.PEAK_TOL, (ch.plan, nrms, peak)


def test_chain_dev_pointers_and_streams(gpu_lib, oracle_lib):
    """device-resident entry point on a caller-provided CUDA stream (torch only supplies memory and the stream)"""
    import torch
    from oracle import oracle as O
    cfg = dict(F.CONFIGS["cfg5_small"], S=5)
    iq = F.make_input(cfg)
    ref, _ = O.chain_run(iq, cfg["in_fmt"], cfg["blkconv"], cfg["stage_a"], cfg["stage_b"], lib=oracle_lib)
    ch = _chain(cfg)
    d_in = torch.from_numpy(iq).cuda()
    cap = ch.max_out(cfg["N"])
    d_out = torch.zeros((cfg["S"], cap, 2), dtype=torch.float32, device="cuda")
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        n = ch.process_dev(d_in.data_ptr(), cfg["N"], cfg["N"], d_out.data_ptr(), cap, st.cuda_stream)
    st.synchronize()
    got = d_out.cpu().numpy()[:, :n].view(np.complex64)[..., 0]
    assert got.shape == ref.shape
    nrms, peak = F.err_metrics(got, ref)
    assert nrms <= F.NRMS_TOL and peak <= F.PEAK_TOL, (nrms, peak)


def test_empty_and_tiny_inputs(gpu_lib):
    cfg = F.CONFIGS["cfg5_small"]
    ch = _chain(cfg)
    iq = F.make_input(dict(cfg, N=5))
    y = ch.process(iq[:, :0])
    assert y.shape == (cfg["S"], 0)
    y = ch.process(iq)          # 5 samples: one decimated sample, one resampled output
    assert y.shape[1] in (0, 1)


def test_dac10_packing(gpu_lib, oracle_lib):
    """float -> 10-bit DAC words, integer exact against the reference's own convert_samples_to_bytes
    (examples/bpsk/bpsk.cxx:76-101, compiled verbatim into oracle/_ref; its restatement on boxes without the reference)"""
    import torch
    from oracle import oracle as O
    rng = np.random.default_rng(1)
    x = np.concatenate([rng.uniform(-1, 1, 4096), np.arange(-1022, 1023) / 1022.0, [1.0, -1.0, 0.0, 0.999]]).astype(np.float32)
    x = x[: len(x) // 4 * 4]
    want = O.f32_to_dac10(x, lib=oracle_lib).reshape(-1, 5)
    d_in = torch.from_numpy(x).cuda()
    d_out = torch.zeros(len(x) // 4 * 5, dtype=torch.uint8, device="cuda")
    assert gpu_lib.sfe_pack_dac10_dev(d_in.data_ptr(), len(x), d_out.data_ptr(), None) == 0
    torch.cuda.synchronize()
    assert np.array_equal(d_out.cpu().numpy().reshape(-1, 5), want)


# ---- tensor-core Toeplitz path (needs >= 8 streams sharing one filter) ------------------------------

@pytest.mark.parametrize("base,S,N", [("cfg5_small", 40, 70001), ("cfg2_small", 33, 50003), ("cfg5_small", 8, 4099)])
def test_chain_tensor_path_vs_oracle(gpu_lib, oracle_lib, base, S, N):
    from oracle import oracle as O
    cfg = dict(F.CONFIGS[base], S=S, N=N)
    iq = F.make_input(cfg)
    iq[0, :100] = 32767          # full-scale corners of the byte-plane split
    iq[1, :100] = -32768
    iq[2, 100:300, 0] = -1
    ref, _ = O.chain_run(iq, cfg["in_fmt"], cfg.get("blkconv"), cfg.get("stage_a"), cfg.get("stage_b"), lib=oracle_lib, n_threads=4)
    ch = _chain(cfg, flags=8)
    assert "toeplitz_i8_tcgen05" in ch.plan, ch.plan
    got = ch.process(iq)
    assert got.shape == ref.shape, (got.shape, ref.shape)
    nrms, peak = F.err_metrics(got, ref)
    assert nrms <= F.NRMS_TOL and peak <= F.PEAK_TOL, (ch.plan, nrms, peak)
    # and against the CUDA-core plan of the same chain
    alt = _chain(cfg, flags=8 + 4).process(iq)
    nrms, peak = F.err_metrics(got, alt)
    assert nrms <= 2e-6 and peak <= 1e-5, (nrms, peak)


def test_chain_tensor_path_chunk_invariance(gpu_lib):
    cfg = dict(F.CONFIGS["cfg5_small"], S=9, N=40000)
    iq = F.make_input(cfg)
    one = _chain(cfg, flags=8).process(iq)
    ch = _chain(cfg, flags=8)
    assert "toeplitz_i8_tcgen05" in ch.plan
    parts, off = [], 0
    for n in (1, 7, 1000, 4096, 12345, 3, 10 ** 9):
        n = min(n, cfg["N"] - off)
        if n <= 0:
            break
        parts.append(ch.process(iq[:, off:off + n]))
        off += n
    many = np.concatenate(parts, axis=1)
    assert many.shape == one.shape
    nrms, peak = F.err_metrics(many, one)
    assert nrms <= 2e-6 and peak <= 1e-5, (nrms, peak)


# ---- K3: the whole chain as one fused polyphase bank (TMA-staged int16 rows, one stream per lane) ------

@pytest.mark.parametrize("base,S,N", [("cfg5_small", 70, 70001), ("cfg5_small", 1, 30011), ("cfg3_small", 37, 200001),
                                      ("cfg5_small", 33, 4099)])
def test_chain_bank_path_vs_oracle(gpu_lib, oracle_lib, base, S, N):
    from oracle import oracle as O
    cfg = dict(F.CONFIGS[base], S=S, N=N)
    iq = F.make_input(cfg)
    iq[0, :100] = 32767
    iq[-1, :100] = -32768
    ref, _ = O.chain_run(iq, cfg["in_fmt"], cfg.get("blkconv"), cfg.get("stage_a"), cfg.get("stage_b"), lib=oracle_lib, n_threads=4)
    ch = _chain(cfg)
    assert ch.plan.startswith("bank_fused"), ch.plan
    got = ch.process(iq)
    assert got.shape == ref.shape, (got.shape, ref.shape)
    nrms, peak = F.err_metrics(got, ref)
    assert nrms <= F.NRMS_TOL and peak <= F.PEAK_TOL, (ch.plan, nrms, peak)
    for s in range(S):                      # every stream on its own: a swapped or stale row must not hide in the mean
        nrms, peak = F.err_metrics(got[s], ref[s])
        assert nrms <= F.NRMS_TOL and peak <= F.PEAK_TOL, (s, nrms, peak)
    alt = _chain(cfg, flags=8).process(iq)  # and against the staged plan of the same chain
    nrms, peak = F.err_metrics(got, alt)
    assert nrms <= 2e-6 and peak <= 1e-5, (nrms, peak)


@pytest.mark.parametrize("base,S", [("cfg5_small", 35), ("cfg3_small", 3)])
def test_chain_bank_path_chunk_invariance(gpu_lib, base, S):
    """ragged call sizes: history rows, held-over outputs and changing sample alignment (n_first mod 4)"""
    cfg = dict(F.CONFIGS[base], S=S, N=60013)
    iq = F.make_input(cfg)
    one = _chain(cfg).process(iq)
    ch = _chain(cfg)
    assert ch.plan.startswith("bank_fused")
    parts, off = [], 0
    for n in (1, 7, 1000, 4096, 12345, 3, 2, 5001, 10 ** 9):
        n = min(n, cfg["N"] - off)
        if n <= 0:
            break
        parts.append(ch.process(iq[:, off:off + n]))
        off += n
    many = np.concatenate(parts, axis=1)
    assert many.shape == one.shape
    nrms, peak = F.err_metrics(many, one)
    assert nrms <= 2e-6 and peak <= 1e-5, (nrms, peak)


def test_chain_bank_unaligned_device_rows(gpu_lib):
    """caller rows that TMA cannot address (odd pitch, 4-byte aligned base) are realigned on the device"""
    import torch
    cfg = dict(F.CONFIGS["cfg5_small"], S=6, N=30001)
    iq = F.make_input(cfg)
    want = _chain(cfg).process(iq)
    ch = _chain(cfg)
    pitch = cfg["N"] + 3
    buf = torch.zeros(cfg["S"] * pitch * 2 + 2, dtype=torch.int16, device="cuda")
    view = buf[2:].view(cfg["S"], pitch, 2)
    view[:, :cfg["N"]] = torch.from_numpy(iq).cuda()
    cap = ch.max_out(cfg["N"])
    d_out = torch.zeros((cfg["S"], cap, 2), dtype=torch.float32, device="cuda")
    n = ch.process_dev(buf.data_ptr() + 4, cfg["N"], pitch, d_out.data_ptr(), cap, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    got = d_out.cpu().numpy()[:, :n].view(np.complex64)[..., 0]
    assert got.shape == want.shape      # (the host path launches stream groups: other work-item cuts, equal to rounding)
    nrms, peak = F.err_metrics(got, want)
    assert nrms <= 1e-6 and peak <= 5e-6, (nrms, peak)


# ---- device wire format: u8 offset-binary I/Q (gr-simplefe/lib/source_c_impl.cc:121-132) ---------------

@pytest.mark.parametrize("base", ["cfg5_small", "cfg2_small", "cfg4_small"])
def test_chain_u8_wire_format_vs_oracle(gpu_lib, oracle_lib, base):
    """the bytes the board sends, (b-128)*(1/127) fused into the first kernel's load"""
    from oracle import oracle as O
    from simplefe_b200.chain import Chain
    cfg = dict(F.CONFIGS[base], in_fmt=F.IN_U8, S=5, N=40003)
    iq = F.make_input(cfg)
    assert iq.dtype == np.uint8
    ref, _ = O.chain_run(iq, cfg["in_fmt"], cfg.get("blkconv"), cfg.get("stage_a"), cfg.get("stage_b"), lib=oracle_lib)
    got = Chain(cfg["S"], cfg["in_fmt"], cfg.get("blkconv"), cfg.get("stage_a"), cfg.get("stage_b")).process(iq)
    assert got.shape == ref.shape
    nrms, peak = F.err_metrics(got, ref)
    assert nrms <= F.NRMS_TOL and peak <= F.PEAK_TOL, (nrms, peak)


@pytest.mark.parametrize("S,N,splits", [(33, 70001, (1, 4097, 30000, 30001, 70001)), (2, 9001, (64, 65, 4000, 9001))])
def test_chain_u8_bank_chunk_invariance(gpu_lib, oracle_lib, S, N, splits):
    """u8 rows through the fused bank kernel (8 samples per 16-byte chunk): any split of the stream into calls
    gives the same samples, and they match the reference classes"""
    from oracle import oracle as O
    from simplefe_b200.chain import Chain
    cfg = dict(F.CONFIGS["cfg5_small"], in_fmt=F.IN_U8, S=S, N=N)
    iq = F.make_input(cfg)
    ch = Chain(S, F.IN_U8, cfg["blkconv"], cfg["stage_a"], cfg["stage_b"])
    assert ch.plan.startswith("bank_fused")
    whole = ch.process(iq)
    ch.reset()
    parts, a = [], 0
    for b in splits:
        parts.append(ch.process(iq[:, a:b]))
        a = b
    pieces = np.concatenate(parts, axis=1)
    assert pieces.shape == whole.shape
    nrms, peak = F.err_metrics(pieces, whole)
    assert nrms <= 2e-6 and peak <= 1e-5, (nrms, peak)
    ref, _ = O.chain_run(iq, F.IN_U8, cfg["blkconv"], cfg["stage_a"], cfg["stage_b"], lib=oracle_lib)
    nrms, peak = F.err_metrics(whole, ref)
    assert whole.shape == ref.shape and nrms <= F.NRMS_TOL and peak <= F.PEAK_TOL, (nrms, peak)


# ---- generic polyphase kernel: tiled (dense outputs) and warp-per-output (sparse outputs) flavours ------------

@pytest.mark.parametrize("cls,up,rate,blk,tapname", [("decimate", 3, 208.4, 4096, "rs512_l3"), ("resample", 3, 57.31, 1000, "rs512_l3"),
                                                     ("decimate", 1, 2.5003, 4096, "taps31"), ("resample", 4, 1.77, 333, "taps31"),
                                                     ("resample", 24, 26.7, 4096, "rs512_l24"), ("decimate", 1, 1.0, 512, "dec128")])
def test_polyphase_objects_fast_order_vs_oracle(gpu_lib, oracle_lib, cls, up, rate, blk, tapname):
    """default (non-exact) summation order of the single-object path: counts bit-exact, values within tolerance"""
    from oracle import oracle as O
    from simplefe_b200 import pydsp
    rng = np.random.default_rng(11)
    x = rng.normal(0, 0.3, 30011).astype(np.float32)
    t = F.taps(tapname)
    ref = getattr(O, cls.capitalize())(t, up, blk, lib=oracle_lib).stream(x, rate)
    g = getattr(pydsp, cls)(t, up, blk)
    out_len = int(np.floor(np.float32(blk) / np.float32(rate))) + 2
    ys = []
    for off in range(0, len(x), blk):
        n, y = g.process(x[off:off + blk], out_len, rate)
        ys.append(y[:n])
    got = np.concatenate(ys)
    assert len(got) == len(ref)
    nrms, peak = F.err_metrics(got, ref)
    assert nrms <= F.NRMS_TOL and peak <= F.PEAK_TOL, (nrms, peak)


@pytest.mark.parametrize("up,rate,tapname,S,N,fmt", [(3, 208.4, "rs512_l3", 5, 150001, F.IN_I16), (1, 2.5003, "taps31", 3, 40001, F.IN_F32),
                                                      (4, 1.3, "taps31", 2, 9001, F.IN_U8), (24, 700.53, "rs512_l24", 3, 90001, F.IN_I16)])
def test_chain_generic_stage_fractional_rates(gpu_lib, oracle_lib, up, rate, tapname, S, N, fmt):
    """one decimate stage with a non-integral step over several streams: the tiled and the warp-per-output kernels,
    fed in two calls (history and held-over outputs cross the seam)"""
    from oracle import oracle as O
    st = dict(klass=F.KLASS_DECIMATE, taps=F.taps(tapname), up=up, blk=4096, rate=rate)
    cfg = dict(in_fmt=fmt, S=S, N=N, seed=77, stage_a=st)
    iq = F.make_input(cfg)
    ch = _chain(cfg)
    assert "polyphase_generic" in ch.plan
    cut = (N // 3) // 4096 * 4096 + 4096
    got = np.concatenate([ch.process(iq[:, :cut]), ch.process(iq[:, cut:])], axis=1)
    ref, _ = O.chain_run(iq, fmt, None, st, None, lib=oracle_lib)
    assert got.shape == ref.shape, (got.shape, ref.shape)
    nrms, peak = F.err_metrics(got, ref)
    assert nrms <= F.NRMS_TOL and peak <= F.PEAK_TOL, (ch.plan, nrms, peak)


@pytest.mark.parametrize("base,N,fmt", [("cfg3_small", 2_000_001, F.IN_I16), ("cfg5_small", 600_001, F.IN_I16), ("cfg5_small", 900_003, F.IN_U8)])
def test_chain_bank_single_stream_time_split(gpu_lib, oracle_lib, base, N, fmt):
    """one long stream: the bank launch cuts the row into pieces that run as the rows of one launch (lanes would idle
    otherwise).  Must equal the unsplit kernel to rounding (same stream fed as two rows) and the oracle within tolerance;
    fed in two ragged calls so that history, alignment and a held-over schedule cross the seam."""
    from oracle import oracle as O
    cfg = dict(F.CONFIGS[base], S=1, N=N, in_fmt=fmt)
    iq = F.make_input(cfg)
    ref, _ = O.chain_run(iq, fmt, cfg.get("blkconv"), cfg.get("stage_a"), cfg.get("stage_b"), lib=oracle_lib)
    cut = N // 2 + 1237
    one = _chain(cfg)
    assert one.plan.startswith("bank_fused"), one.plan
    got = np.concatenate([one.process(iq[:, :cut]), one.process(iq[:, cut:])], axis=1)
    two = _chain(cfg, S=2)
    iq2 = np.concatenate([iq, iq], axis=0)
    dup = np.concatenate([two.process(iq2[:, :cut]), two.process(iq2[:, cut:])], axis=1)
    assert got.shape == ref.shape, (got.shape, ref.shape)
    # the two rows of one launch are bit-identical; the time-split launch cuts the row at other places than the row
    # launch does, and a window that straddles a cut is summed in two parts (older + newer): same numbers to rounding
    assert np.array_equal(dup[0], dup[1])
    nrms, peak = F.err_metrics(got[0], dup[0])
    assert nrms <= 1e-6 and peak <= 5e-6, (nrms, peak)
    nrms, peak = F.err_metrics(got, ref)
    assert nrms <= F.NRMS_TOL and peak <= F.PEAK_TOL, (nrms, peak)


@pytest.mark.parametrize("up,rate,tapname,S,N,fmt", [(1, 2.5003, "taps31", 40, 40001, F.IN_F32), (4, 1.77, "taps31", 33, 20001, F.IN_I16),
                                                      (3, 57.31, "rs512_l3", 17, 150001, F.IN_I16), (4, 0.77, "taps31", 35, 5001, F.IN_U8),
                                                      (24, 26.7, "rs512_l24", 64, 60001, F.IN_I16)])
def test_chain_generic_stage_many_streams(gpu_lib, oracle_lib, up, rate, tapname, S, N, fmt):
    """>= 16 streams sharing a fractional-rate filter run the lane-per-stream kernel (K2s): ragged last group,
    two calls (history, held-over output across the seam), decimation and interpolation"""
    from oracle import oracle as O
    klass = F.KLASS_RESAMPLE if rate < 1.0 else F.KLASS_DECIMATE
    st = dict(klass=klass, taps=F.taps(tapname), up=up, blk=4096, rate=rate)
    cfg = dict(in_fmt=fmt, S=S, N=N, seed=91, stage_a=st)
    iq = F.make_input(cfg)
    ch = _chain(cfg)
    assert "polyphase_generic" in ch.plan
    cut = (N // 3) // 4096 * 4096 + 4096
    got = np.concatenate([ch.process(iq[:, :cut]), ch.process(iq[:, cut:])], axis=1)
    ref, _ = O.chain_run(iq, fmt, None, st, None, lib=oracle_lib, n_threads=4)
    assert got.shape == ref.shape, (got.shape, ref.shape)
    for s in range(S):
        nrms, peak = F.err_metrics(got[s], ref[s])
        assert nrms <= F.NRMS_TOL and peak <= F.PEAK_TOL, (s, ch.plan, nrms, peak)


def test_chain_cfg4_many_streams(gpu_lib, oracle_lib):
    """the BPSK receive chain (FFT blkconv -> fractional decimate -> fractional resample) on 33 streams"""
    from oracle import oracle as O
    cfg = dict(F.CONFIGS["cfg4_small"], S=33, N=30000)
    iq = F.make_input(cfg)
    ref, _ = O.chain_run(iq, cfg["in_fmt"], cfg.get("blkconv"), cfg.get("stage_a"), cfg.get("stage_b"), lib=oracle_lib, n_threads=4)
    got = _chain(cfg).process(iq)
    assert got.shape == ref.shape
    for s in range(33):
        nrms, peak = F.err_metrics(got[s], ref[s])
        assert nrms <= F.NRMS_TOL and peak <= F.PEAK_TOL, (s, nrms, peak)
