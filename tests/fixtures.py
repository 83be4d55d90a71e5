"""Shared test/bench fixtures: BASELINE.json configs and their synthetic inputs."""
from __future__ import annotations

import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")

IN_I16, IN_U8, IN_F32 = 0, 1, 2
KLASS_DECIMATE, KLASS_RESAMPLE = 1, 2

_taps = None


def taps(name: str) -> np.ndarray:
    global _taps
    if _taps is None:
        _taps = dict(np.load(os.path.join(GOLDEN, "taps.npz")))
    return _taps[name]


def _lazy_configs():
    R3 = float(np.float32(625.0 / 3.0))
    R24 = float(np.float32(625.0 / 24.0))

    def st(klass, t, up, blk, rate):
        return dict(klass=klass, taps=taps(t), up=up, blk=blk, rate=rate)

    full = dict(
        # cfg 1: blkconv, 2^20 complex float samples, 255 taps, FFT 4096
        cfg1=dict(in_fmt=IN_F32, S=1, N=1 << 20, seed=1001, blkconv=(taps("lp255"), 4096)),
        # cfg 2: decimate by 8, 128 taps, 10 MS/s int16 IQ, one stream
        cfg2=dict(in_fmt=IN_I16, S=1, N=10_000_000, seed=1002, stage_a=st(KLASS_DECIMATE, "dec128", 1, 4096, 8.0)),
        # cfg 3: rational resample 10 MS/s -> 48 kS/s, L=3, M=625, 512 taps
        cfg3=dict(in_fmt=IN_I16, S=1, N=10_000_000, seed=1003, stage_a=st(KLASS_DECIMATE, "rs512_l3", 3, 4096, R3)),
        # cfg 4: BPSK receive chain with fractional rates (mu != 0, left-over outputs)
        cfg4=dict(in_fmt=IN_I16, S=1, N=1_000_000, seed=1004, bpsk=True, blkconv=(taps("rrc111"), 2048),
                  stage_a=st(KLASS_DECIMATE, "taps31", 1, 4096, 2.5003), stage_b=st(KLASS_RESAMPLE, "taps31", 4, 4096, 1.77)),
        # cfg 5: 1024 int16 IQ streams x 2^20 through blkconv -> decimate -> resample
        cfg5=dict(in_fmt=IN_I16, S=1024, N=1 << 20, seed=5000, blkconv=(taps("lp255"), 4096),
                  stage_a=st(KLASS_DECIMATE, "dec128", 1, 4096, 8.0), stage_b=st(KLASS_RESAMPLE, "rs512_l24", 24, 4096, R24)),
    )
    small = dict(
        cfg1_small=dict(full["cfg1"], S=2, N=20000),
        cfg2_small=dict(full["cfg2"], S=2, N=50003),
        cfg3_small=dict(full["cfg3"], S=2, N=200001),
        cfg4_small=dict(full["cfg4"], S=2, N=60000),
        cfg5_small=dict(full["cfg5"], S=3, N=70001),
    )
    full.update(small)
    return full


class _Configs(dict):
    def __missing__(self, key):
        self.update(_lazy_configs())
        return dict.__getitem__(self, key)


CONFIGS = _Configs()


def _quant16(x):
    return np.round(np.clip(x, -1.0, 1.0) * 32767.0).astype(np.int16)


def bpsk_waveform(n: int, rng) -> np.ndarray:
    """Restates the generator of examples/bpsk/bpsk.cxx:145-159 with a seeded PRNG: +-0.85/1.35 impulses every
    10 samples, shaped by the 111-tap RRC (the reference's transmit filter)."""
    sps, amp = 10, np.float32(0.85) / np.float32(1.35)
    bits = rng.integers(0, 2, size=(n + sps - 1) // sps)
    x = np.zeros(n, np.float64)
    x[::sps] = np.where(bits[: len(x[::sps])] == 1, -amp, amp)
    return np.convolve(x, taps("rrc111").astype(np.float64))[:n]


def make_input(cfg: dict, streams=None) -> np.ndarray:
    """(S, N, 2) array in cfg['in_fmt']; stream s uses seed cfg['seed'] + s."""
    S, N = cfg["S"], cfg["N"]
    idx = range(S) if streams is None else streams
    dt = {IN_I16: np.int16, IN_U8: np.uint8, IN_F32: np.float32}[cfg["in_fmt"]]
    out = np.empty((len(idx), N, 2), dt)
    for k, s in enumerate(idx):
        rng = np.random.Generator(np.random.PCG64(cfg["seed"] + s))
        if cfg.get("bpsk"):
            tx = bpsk_waveform(N, rng)
            v = np.stack([0.5 * tx + rng.normal(0, 0.01, N), rng.normal(0, 0.01, N)], axis=1)
        else:
            v = np.clip(rng.normal(0.0, 0.25, size=(N, 2)), -1.0, 1.0)
        if cfg["in_fmt"] == IN_I16:
            out[k] = _quant16(v)
        elif cfg["in_fmt"] == IN_U8:
            out[k] = np.clip(np.round(v * 127.0) + 128, 0, 255).astype(np.uint8)
        else:
            out[k] = v.astype(np.float32)
    return out


def err_metrics(got: np.ndarray, ref: np.ndarray):
    """nRMS = rms(g-r)/rms(r); peak-normalised max error = max|g-r| / max|r|   (SURVEY.md H2)."""
    g = got.astype(np.complex128).ravel()
    r = ref.astype(np.complex128).ravel()
    d = g - r
    rms_r = np.sqrt(np.mean(np.abs(r) ** 2))
    nrms = float(np.sqrt(np.mean(np.abs(d) ** 2)) / (rms_r if rms_r > 0 else 1.0))
    peak = float(np.max(np.abs(d)) / (np.max(np.abs(r)) if np.max(np.abs(r)) > 0 else 1.0))
    return nrms, peak


NRMS_TOL = 1e-5      # north_star: float32 outputs within 1e-5 normalised RMS error
PEAK_TOL = 1e-4      # and 1e-4 max relative error (peak-normalised)
