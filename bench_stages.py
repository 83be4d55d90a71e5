#!/usr/bin/env python
"""bench_stages.py -- secondary report: every BASELINE.json config and every kernel plan, one GPU.

bench.py is the contract (cfg 5, one JSON line).  This script fills BASELINE.md's table: for each config of
SURVEY.md 8-d it times the device-resident call (CUDA events on the launching stream, best-of after warm-up),
names the kernels the chain picked, and reports algorithmic GB/s against the measured HBM peak.  Single-stream
configs are launch-latency bound by construction, so each also runs batched.

    python bench_stages.py [--quick] > gpurun_out/stages.json
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import fixtures as F  # noqa: E402

FLAG_EXACT, FLAG_NO_FUSE, FLAG_NO_TENSOR, FLAG_NO_BANK = 1, 2, 4, 8


def bytes_alg(cfg, n_in, n_out):
    """first read of the input + final write of the output, per stream"""
    return n_in * {F.IN_I16: 4, F.IN_U8: 2, F.IN_F32: 8}[cfg["in_fmt"]] + n_out * 8


def run_case(name, cfg, S, N, flags, steps, warmup, peak, torch, Chain):
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev)
    g.manual_seed(cfg["seed"])
    if cfg["in_fmt"] == F.IN_I16:
        d_in = torch.empty((S, N, 2), dtype=torch.int16, device=dev)
        rows = max(1, (128 << 20) // (N * 8))
        for s0 in range(0, S, rows):
            blk = torch.randn((min(rows, S - s0), N, 2), generator=g, device=dev) * 0.25
            d_in[s0:s0 + rows] = torch.round(blk.clamp_(-1, 1) * 32767.0).to(torch.int16)
        del blk
    elif cfg["in_fmt"] == F.IN_U8:
        d_in = torch.empty((S, N, 2), dtype=torch.uint8, device=dev)
        rows = max(1, (128 << 20) // (N * 8))
        for s0 in range(0, S, rows):
            blk = torch.randn((min(rows, S - s0), N, 2), generator=g, device=dev) * 0.25
            d_in[s0:s0 + rows] = (torch.round(blk.clamp_(-1, 1) * 127.0) + 128).clamp_(0, 255).to(torch.uint8)
        del blk
    else:
        d_in = (torch.randn((S, N, 2), generator=g, device=dev) * 0.25).clamp_(-1, 1)
    ch = Chain(S, cfg["in_fmt"], cfg.get("blkconv"), cfg.get("stage_a"), cfg.get("stage_b"), flags=flags, device=0)
    cap = ch.max_out(N)
    d_out = torch.zeros((S, cap, 2), dtype=torch.float32, device=dev)
    st = torch.cuda.current_stream()

    def step():
        return ch.process_dev(d_in.data_ptr(), N, N, d_out.data_ptr(), cap, st.cuda_stream)

    for _ in range(warmup):
        n_out = step()
    torch.cuda.synchronize()
    ch.profile(True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(steps):
        n_out = step()
    e1.record(st)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    ch.profile(False)
    prof = ch.profile_read()
    b = bytes_alg(cfg, N, n_out) * S
    res = {"case": name, "streams": S, "samples_per_stream": N, "outputs_per_stream": n_out, "plan": ch.plan,
           "ms_per_step": ms, "Msamples_per_s": S * N / ms / 1e3, "bytes_alg_per_sample": b / (S * N),
           "alg_GBps": b / ms / 1e6, "frac_of_measured_hbm": b / ms / 1e6 / peak,
           "kernels": [{"name": n, "ms": m} for n, m, _ in prof]}
    del ch, d_in, d_out
    torch.cuda.empty_cache()
    return res


def cpu_table():
    """the reference's CPU code (oracle/_ref, -O3) on every BASELINE config: one stream on one thread, and one
    stream per thread on all host threads, on a bounded sample (<= 2^20 samples per stream)"""
    import time
    import numpy as np
    from oracle import oracle as O
    lib = O.load("auto", fast=True)
    T = len(os.sched_getaffinity(0))
    rows = []
    for name in ("cfg1", "cfg2", "cfg3", "cfg4", "cfg5"):
        cfg = dict(F.CONFIGS[name], S=1, N=min(F.CONFIGS[name]["N"], 1 << 20))
        x1 = F.make_input(cfg)
        xT = np.ascontiguousarray(np.repeat(x1, T, axis=0))
        best1 = bestT = 1e30
        for _ in range(3):
            t0 = time.perf_counter()
            O.chain_run(x1, cfg["in_fmt"], cfg.get("blkconv"), cfg.get("stage_a"), cfg.get("stage_b"), n_threads=1, lib=lib)
            best1 = min(best1, time.perf_counter() - t0)
            t0 = time.perf_counter()
            O.chain_run(xT, cfg["in_fmt"], cfg.get("blkconv"), cfg.get("stage_a"), cfg.get("stage_b"), n_threads=T, lib=lib)
            bestT = min(bestT, time.perf_counter() - t0)
        rows.append({"config": name, "samples_per_stream": cfg["N"], "cpu_1thread_Msamples_per_s": cfg["N"] / best1 / 1e6,
                     "cpu_all_threads_Msamples_per_s": T * cfg["N"] / bestT / 1e6, "threads": T,
                     "kind": lib.sfeo_kind().decode(), "build": lib.sfeo_build_flags().decode()})
        print(json.dumps(rows[-1]), file=sys.stderr, flush=True)
    return rows


def tx_cases(peak, torch):
    """batched BPSK transmitter: bits -> RRC interpolator -> float samples / 10-bit DAC words.
    Unit: output samples; algorithmic bytes = bits read + samples written."""
    from simplefe_b200 import bpsk as B
    out = []
    C, n_sym, sps = 1024, 1 << 17, 10
    dev = torch.device("cuda", 0)
    words = torch.randint(0, 2 ** 31 - 1, (C, n_sym // 32), dtype=torch.int32, device=dev)
    for name, fmt, bps in (("float32 samples", B.TX_F32, 4.0), ("10-bit DAC words (5 bytes per 4 samples)", B.TX_DAC10, 1.25)):
        tx = B.BpskTx(F.taps("rrc111"), sps, C)
        d_out = torch.empty((C, int(n_sym * sps * bps)), dtype=torch.uint8, device=dev)
        st = torch.cuda.current_stream()

        def step():
            tx.process_dev(words.data_ptr(), n_sym, n_sym // 32, d_out.data_ptr(), d_out.shape[1], fmt, st.cuda_stream)

        for _ in range(3):
            step()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(10):
            step()
        e1.record(st)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        n = C * n_sym * sps
        b = C * n_sym / 8 + n * bps
        out.append({"case": f"bpsk tx x{C} channels, 111-tap RRC, {sps} samples/symbol -> {name}", "channels": C, "symbols_per_channel": n_sym,
                    "ms_per_step": ms, "Msamples_per_s": n / ms / 1e3, "bytes_alg_per_sample": b / n, "alg_GBps": b / ms / 1e6,
                    "frac_of_measured_hbm": b / ms / 1e6 / peak})
        print(json.dumps(out[-1]), file=sys.stderr, flush=True)
        del tx, d_out
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--only", default="")
    ap.add_argument("--cpu", action="store_true", help="only time the reference CPU code on every config")
    args = ap.parse_args()
    if args.cpu:
        print(json.dumps({"cpu": cpu_table()}, indent=1))
        return
    import torch
    import simplefe_b200  # noqa: F401
    from simplefe_b200.chain import Chain
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        peak = 6650.0
    C = F.CONFIGS
    big = 1 << 20
    cases = [
        # name, config, S, N, flags
        ("cfg1 blkconv c64 (single stream)", C["cfg1"], 1, big, 0),
        ("cfg1 blkconv c64 x256 streams", C["cfg1"], 256, big, 0),
        ("cfg1' blkconv int16 x512 streams", dict(C["cfg1"], in_fmt=F.IN_I16), 512, big, 0),
        ("cfg2 decimate/8 int16 (single stream, 10 MS)", C["cfg2"], 1, 10_000_000, 0),
        ("cfg2 decimate/8 int16 x1024 streams [tcgen05 Toeplitz]", C["cfg2"], 1024, big, 0),
        ("cfg2 decimate/8 int16 x1024 streams [CUDA-core FIR]", C["cfg2"], 1024, big, FLAG_NO_TENSOR),
        ("cfg3 resample 3/625 int16 (single stream, 10 MS)", C["cfg3"], 1, 10_000_000, 0),
        ("cfg3 resample 3/625 int16 x1024 streams [bank]", C["cfg3"], 1024, big, 0),
        ("cfg3 resample 3/625 int16 x1024 streams [generic]", C["cfg3"], 1024, big, FLAG_NO_BANK),
        ("cfg4 bpsk chain, fractional rates (single stream)", C["cfg4"], 1, 1_000_000, 0),
        ("cfg4 bpsk chain, fractional rates x256 streams", C["cfg4"], 256, 1_000_000, 0),
        ("cfg5 chain x1024 [one fused bank kernel]", C["cfg5"], 1024, big, 0),
        ("cfg5 chain x1024, u8 device wire format [one fused bank kernel]", dict(C["cfg5"], in_fmt=F.IN_U8), 1024, big, 0),
        ("cfg5 chain x1024 [fused blkconv*decim on tcgen05 + generic resample]", C["cfg5"], 1024, big, FLAG_NO_BANK),
        ("cfg5 chain x1024 [fused blkconv*decim CUDA-core FIR + generic]", C["cfg5"], 1024, big, FLAG_NO_BANK | FLAG_NO_TENSOR),
        ("cfg5 chain x256 [stage by stage: FFT blkconv, FIR, generic]", C["cfg5"], 256, big, FLAG_NO_BANK | FLAG_NO_FUSE | FLAG_NO_TENSOR),
    ]
    out = []
    for name, cfg, S, N, flags in cases:
        if args.only and args.only not in name:
            continue
        steps, warmup = (3, 2) if args.quick else (10, 3)
        try:
            r = run_case(name, cfg, S, N, flags, steps, warmup, peak, torch, Chain)
        except Exception as ex:  # report, keep going
            r = {"case": name, "error": repr(ex)}
        out.append(r)
        print(json.dumps(r), file=sys.stderr, flush=True)
    if not args.only or "tx" in args.only:
        try:
            out.extend(tx_cases(peak, torch))
        except Exception as ex:
            out.append({"case": "bpsk tx", "error": repr(ex)})
    print(json.dumps({"hbm_peak_GBps": peak, "cases": out}, indent=1))


if __name__ == "__main__":
    main()
