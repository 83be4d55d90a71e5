#!/usr/bin/env python
"""bench.py -- headline benchmark of the libdsp hot path on B200.

Workload (BASELINE.json configs[4], SURVEY.md 8-d cfg 5): 1024 independent int16 IQ streams x 2^20
samples through blkconv(255 taps, FFT 4096) -> decimate(128 taps, /8) -> resample(512 taps, L=24,
rate 625/24), one shared filter set.  The 1024 streams are SPLIT over the N GPUs (strong scaling:
1024 / 512 / 256 / 128 streams per GPU at N = 1 / 2 / 4 / 8); there is no collective:
the path never exchanges data between streams.  Metric: Msamples/s, one sample = one complex input sample.

    python bench.py [--gpus N --steps K --warmup W] [--impl reference] [--config cfg1..cfg5]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One JSON line on stdout (rank 0).  `value` = device-resident throughput (CUDA events, max over
ranks); `e2e` = the same through the host-buffer C-ABI call (pinned host in, pinned host out,
copies inside the timed region) next to the box's measured concurrent H2D ceiling; `roofline` = the
dominant kernel against the measured HBM peak; `parity` = a sample of the output rows of every rank
against the reference's own CPU code (outside the timed region); `cpu_baseline` = that CPU code timed
on this box's host cores; `weak` (N > 1) = the secondary figure with 1024 streams on every GPU.
`--config cfg1..cfg4` runs the other BASELINE.json configs through the same harness (same keys).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402

import fixtures as F  # noqa: E402

BYTES_ALG_PER_SAMPLE = 4.0 + 8.0 * 3.0 / 625.0   # SURVEY.md 8-d cfg 5: int16 IQ in + complex64 out at 48k/10M
METRIC = "Msamples/s"


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# stdout carries exactly one JSON line.  Libraries write there too (NCCL prints its version line from C), so file
# descriptor 1 is pointed at stderr for the whole run and the line goes out through a private duplicate.
_JSON_FD = None


def capture_stdout():
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line: dict):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


# ---------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe).  The sampler is started
    early (nvidia-smi needs ~0.1-0.3 s before its first line) and every line carries a timestamp, so samples are
    attributed to the window they were taken in."""
    Q = ("timestamp,index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.gpu), "-lms", "20"], stdout=open(self.path, "w"),
                                         stderr=subprocess.DEVNULL)
            t0 = time.time()
            while time.time() - t0 < 3.0 and os.path.getsize(self.path) == 0:   # wait for the first line
                time.sleep(0.02)
        except Exception:
            self.proc = None

    def stop(self, t_lo: float, t_hi: float) -> dict:
        """summary of the samples whose timestamp lies in [t_lo, t_hi] (time.time() seconds)"""
        import datetime
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons, n_all = [], [], [], set(), 0
        for line in open(self.path):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 10:
                continue
            n_all += 1
            try:
                ts = datetime.datetime.strptime(f[0], "%Y/%m/%d %H:%M:%S.%f").timestamp()
                if ts < t_lo - 0.011 or ts > t_hi + 0.011:
                    continue
                sm.append(float(f[2]))
                mx.append(float(f[3]))
                pw.append(float(f[4]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[6:10]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.path)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"], "samples_total": n_all}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "power_w_max": float(max(pw)),
                "samples": len(sm), "reasons": sorted(reasons)}


def measured_peaks():
    try:
        d = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, burst copy)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic(name, plan=None):
    """dram bytes per launch of the dominant kernel from the committed ncu capture of THIS build's kernel, if any
    (profiles/dominant_kernel_traffic.json is regenerated by scripts/ncu_traffic.py from an `ncu --set full` capture;
    an entry is used only when its workload and kernel plan match the run)"""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "dominant_kernel_traffic.json")))
        e = d.get(name) if isinstance(d.get(name), dict) else (d if name == "cfg5" and "dram_bytes_per_launch" in d else None)
        if e and plan and e.get("plan") and e["plan"] != plan:
            return {"dram_bytes_per_launch": None, "source": f"stale: capture was of plan {e['plan']}"}
        return e
    except Exception:
        return None


# ---------------------------------------------------------------------------------------------------
# workloads: BASELINE.json configs.  cfg5 is the headline; cfg1..cfg4 run through the same harness.
FLAG_EXACT, FLAG_NO_FUSE, FLAG_NO_TENSOR, FLAG_NO_BANK = 1, 2, 4, 8
EB = {F.IN_I16: 4, F.IN_U8: 2, F.IN_F32: 8}
WORKLOADS = {
    # name: (fixture, streams batched, samples, streams as worded, samples as worded, text)
    "cfg1": ("cfg1", 256, 1 << 20, 1, 1 << 20, "cfg1: blkconv 255 taps, FFT 4096 on complex-float streams of 2^20 samples"),
    "cfg2": ("cfg2", 1024, 1 << 20, 1, 10_000_000, "cfg2: decimate /8, 128-tap polyphase FIR on int16 IQ streams"),
    "cfg3": ("cfg3", 1024, 1 << 20, 1, 10_000_000, "cfg3: rational resample 10 MS/s -> 48 kS/s (L=3, M=625, 512 taps) on int16 IQ streams"),
    "cfg4": ("cfg4", 256, 1_000_000, 1, 1_000_000, "cfg4: BPSK receive chain blkconv(111 RRC, FFT 2048) -> decimate(31 taps, rate 2.5003) -> "
                                                     "resample(31 taps, L=4, rate 1.77), fractional rates"),
    "cfg5": ("cfg5", 1024, 1 << 20, 1024, 1 << 20, "cfg5: 1024 int16 IQ streams x 2^20 samples, blkconv(255 taps, FFT 4096) -> decimate(128 taps, /8) "
                                                     "-> resample(512 taps, L=24, rate 625/24)"),
}


def bytes_alg_per_sample(name, cfg, N, n_out):
    """first read of the input + final write of the output, per complex input sample"""
    if name == "cfg5" and cfg["in_fmt"] == F.IN_I16:
        return BYTES_ALG_PER_SAMPLE
    return EB[cfg["in_fmt"]] + 8.0 * n_out / max(N, 1)


def cpu_chain(cfg, iq, n_threads, fast=True):
    from oracle import oracle as O
    lib = O.load("auto", fast=fast)
    t0 = time.perf_counter()
    y, n_out = O.chain_run(iq, cfg["in_fmt"], cfg.get("blkconv"), cfg.get("stage_a"), cfg.get("stage_b"),
                           n_threads=n_threads, lib=lib)
    return time.perf_counter() - t0, y, lib


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_build_note(lib, cfg):
    b = lib.sfeo_build_flags().decode()
    if cfg.get("blkconv") is not None:
        b += " | blkconv's FFT = this repo's float Stockham shim (oracle/fftw_shim.cxx), NOT FFTW: the image has no libfftw3f"
    return b


def strong_cpu_fft(cfg, make_rows, budget_s=4.0):
    """SURVEY.md 8-d 'strong CPU' yardstick for the FFT stage: scipy float32 pocketfft overlap-add of blkconv's stage
    only (int16->float conversion included), vectorised over blocks, all host threads.  It bounds how much the
    shim-backed reference build understates a CPU with a real FFT library."""
    if cfg.get("blkconv") is None:
        return None
    try:
        import scipy.fft as sfft
    except Exception as ex:
        return {"error": repr(ex)}
    T = host_threads()
    taps, nfft = cfg["blkconv"]
    taps = np.asarray(taps, np.float32)
    blk = nfft + 1 - len(taps)
    Hf = sfft.rfft(np.concatenate([taps, np.zeros(nfft - len(taps), np.float32)])).astype(np.complex64)
    N = cfg["N"]

    def run(iq):
        S = iq.shape[0]
        if cfg["in_fmt"] == F.IN_I16:
            x = iq.astype(np.float32) * np.float32(1.0 / 32767.0)
        elif cfg["in_fmt"] == F.IN_U8:
            x = (iq.astype(np.float32) - np.float32(128)) * np.float32(1.0 / 127.0)
        else:
            x = iq
        x = np.ascontiguousarray(np.moveaxis(x, 2, 1)).reshape(2 * S, N)          # one real row per component
        nb = (N + blk - 1) // blk
        xb = np.zeros((2 * S, nb, nfft), np.float32)
        xb.reshape(2 * S, nb * nfft)  # view check
        pad = np.zeros((2 * S, nb * blk), np.float32)
        pad[:, :N] = x
        xb[:, :, :blk] = pad.reshape(2 * S, nb, blk)
        Y = sfft.rfft(xb, axis=-1, workers=T)
        Y *= Hf
        yb = sfft.irfft(Y, n=nfft, axis=-1, workers=T)
        out = np.zeros((2 * S, nb * blk + nfft), np.float32)
        out[:, :nb * blk] = yb[:, :, :blk].reshape(2 * S, nb * blk)
        tail = yb[:, :, blk:]
        for j in range(nb):                                                       # overlap-add of the tails
            out[:, (j + 1) * blk:(j + 1) * blk + nfft - blk] += tail[:, j]
        return out[:, :N]

    t0 = time.perf_counter()
    run(make_rows(1))
    t1 = time.perf_counter() - t0
    S = int(max(1, min(4 * T, budget_s / max(t1, 1e-3))))
    iq = make_rows(S)
    t0 = time.perf_counter()
    run(iq)
    dt = time.perf_counter() - t0
    return {"value": S * N / dt / 1e6, "unit": METRIC, "cores": T, "stage": "blkconv only (the chain's later stages excluded)",
            "kind": "scipy.fft float32 pocketfft overlap-add, vectorised over blocks",
            "sample": f"{S} streams x {N} samples, {dt:.2f} s"}


def cpu_baseline(cfg, make_rows, budget_s=12.0):
    """The reference's CPU path on a bounded sample of the same workload, all host threads."""
    T = host_threads()
    N = cfg["N"]
    t1, _, lib = cpu_chain(cfg, make_rows(1), 1)            # one stream on one thread: calibration
    per_thread = max(1, int(budget_s / max(t1, 1e-3)))
    S = int(min(max(cfg["S"], T), max(T, per_thread * T)))
    S = max(T, S // T * T) if S >= T else S
    iq = make_rows(S)
    dt, _, _ = cpu_chain(cfg, iq, T)
    v = S * N / dt / 1e6
    out = {"value": v, "unit": METRIC, "cores": T, "kind": lib.sfeo_kind().decode(),
           "sample": f"{S} streams x {N} samples ({cfg['S']} in the workload), {T} threads, {dt:.2f} s; 1 thread: {N / t1 / 1e6:.2f} {METRIC}",
           "build": cpu_build_note(lib, cfg), "one_thread_value": N / t1 / 1e6}
    try:
        sc = strong_cpu_fft(cfg, make_rows)
        if sc is not None:
            out["strong_cpu_fft"] = sc
    except Exception as ex:
        out["strong_cpu_fft"] = {"error": repr(ex)}
    return out


def host_rows(cfg, n_extra=8192):
    """S distinct rows in the config's input format as windows of one long seeded base signal (cheap on the host)"""
    rng = np.random.Generator(np.random.PCG64(cfg["seed"]))
    v = np.clip(rng.normal(0.0, 0.25, size=(cfg["N"] + n_extra, 2)), -1, 1)
    if cfg["in_fmt"] == F.IN_I16:
        base = F._quant16(v)
    elif cfg["in_fmt"] == F.IN_U8:
        base = np.clip(np.round(v * 127.0) + 128, 0, 255).astype(np.uint8)
    else:
        base = v.astype(np.float32)

    def make_rows(S):
        return np.stack([base[(37 * s) % n_extra:(37 * s) % n_extra + cfg["N"]] for s in range(S)])
    return make_rows


# ---------------------------------------------------------------------------------------------------
def run_reference(args, name, cfg, rank, world):
    """--impl reference: the reference's own CPU implementation (oracle/_ref), rank 0 only."""
    if rank != 0:
        return
    T = host_threads()
    make_rows = host_rows(cfg)
    t1, _, lib = cpu_chain(cfg, make_rows(1), 1)
    S = int(min(max(cfg["S"], T), max(T, int(8.0 / max(t1, 1e-3)) * T)))
    S = max(T, S // T * T) if S >= T else S
    iq = make_rows(S)
    for _ in range(args.warmup):
        cpu_chain(cfg, iq, T)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_chain(cfg, iq, T)
    dt = time.perf_counter() - t0
    v = S * cfg["N"] * args.steps / dt / 1e6
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": METRIC, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": args.scaling,
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(name, cfg, args.gpus, args.scaling),
        "cpu_baseline": {"value": v, "unit": METRIC, "cores": T, "kind": lib.sfeo_kind().decode(),
                         "sample": f"{S} streams x {cfg['N']} samples per step ({cfg['S']} in the workload), {T} threads",
                         "build": cpu_build_note(lib, cfg)},
        "e2e": {"value": v, "unit": METRIC, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def workload_config(name, cfg, n_gpus, scaling):
    """identical in both arms (ours / reference): the driver compares the two `config` objects"""
    S = cfg["S"]
    fmt_name = {F.IN_I16: "int16 IQ", F.IN_U8: "u8 offset-binary IQ", F.IN_F32: "complex64"}[cfg["in_fmt"]]
    return {"workload": WORKLOADS[name][5], "input_format": fmt_name, "streams_total": S, "samples_per_stream": cfg["N"],
            "streams_per_gpu": S // n_gpus if scaling == "strong" else S,
            "partition": f"independent streams split over {n_gpus} GPU(s) ({scaling}), no collective",
            "l2": "inputs larger than L2 (no flush needed)" if S * cfg["N"] * EB[cfg["in_fmt"]] // max(n_gpus, 1) > (256 << 20)
                  else "L2 flushed between timed steps (write of a 256 MiB buffer)"}


def device_rows(torch, cfg, S, N, dev, seed):
    """synthetic rows generated on the device: N(0, 0.25^2) clipped, quantised to the config's input format"""
    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    fmt = cfg["in_fmt"]
    dt = {F.IN_I16: torch.int16, F.IN_U8: torch.uint8, F.IN_F32: torch.float32}[fmt]
    d_in = torch.empty((S, N, 2), dtype=dt, device=dev)
    rows = max(1, (256 << 20) // (N * 2 * 4))
    for s0 in range(0, S, rows):
        blk = (torch.randn((min(rows, S - s0), N, 2), generator=g, device=dev) * 0.25).clamp_(-1, 1)
        if fmt == F.IN_I16:
            d_in[s0:s0 + rows] = torch.round(blk * 32767.0).to(torch.int16)
        elif fmt == F.IN_U8:
            d_in[s0:s0 + rows] = (torch.round(blk * 127.0) + 128).clamp_(0, 255).to(torch.uint8)
        else:
            d_in[s0:s0 + rows] = blk
        del blk
    return d_in


def check_rows(S):
    """rows compared with the oracle after the timed region: first/last of the first 32-stream group, first of the
    second, the middle, the last, plus one row of every eighth group"""
    rows = {0, min(31, S - 1), min(32, S - 1), S // 2, S - 1}
    for g in range(0, (S + 31) // 32, 8):
        rows.add(min(S - 1, 32 * g + (5 * g + 3) % 32))
    return sorted(rows)


# ---------------------------------------------------------------------------------------------------
def run_box(args, name, cfg):
    """--box: the same workload on N GPUs from one process.  The library partitions the streams (sfe_box_create) and
    drives every GPU from its own worker thread; torch only provides per-device memory, streams and events."""
    import torch
    import simplefe_b200
    from simplefe_b200.chain import Box
    lib = simplefe_b200.lib()
    G, N, eb = args.gpus, cfg["N"], EB[cfg["in_fmt"]]
    assert args.scaling == "strong", "--box runs the split workload"
    box = Box(cfg["S"], cfg["in_fmt"], cfg.get("blkconv"), cfg.get("stage_a"), cfg.get("stage_b"), flags=args.flags, n_devices=G)
    assert box.n_devices == G, (box.n_devices, G)
    cap = max(box.max_out(N), 1)
    ranges = [box.stream_range(g) for g in range(G)]
    d_in, d_out, streams = [], [], []
    for dev_i, s0, ns in ranges:
        dev = torch.device("cuda", dev_i)
        with torch.cuda.device(dev):
            d_in.append(device_rows(torch, cfg, ns, N, dev, 5000 + s0))
            d_out.append(torch.zeros((ns, cap, 2), dtype=torch.float32, device=dev))
            streams.append(torch.cuda.current_stream(dev))
    pin, pout, pst = [t.data_ptr() for t in d_in], [t.data_ptr() for t in d_out], [s.cuda_stream for s in streams]

    def step():
        return box.process_dev(pin, N, N, pout, cap, pst)

    def sync_all():
        for g in range(G):
            torch.cuda.synchronize(g)

    sampler = ClockSampler(0)
    sampler.start()                                        # before the warm-up: no idle gap in front of the timed region
    ev = []
    for g in range(G):
        with torch.cuda.device(g):
            ev.append((torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)))
    for _ in range(args.warmup):
        n_out = step()
    sync_all()
    launches0 = lib.sfe_kernel_launches()
    w0 = time.time()
    for g in range(G):
        ev[g][0].record(streams[g])
    h0 = time.perf_counter()
    for _ in range(args.steps):
        n_out = step()
    host_us = (time.perf_counter() - h0) / args.steps * 1e6
    for g in range(G):
        ev[g][1].record(streams[g])
    sync_all()
    w1 = time.time()
    ms = max(a.elapsed_time(b) for a, b in ev)              # device time, max over the GPUs
    launches = lib.sfe_kernel_launches() - launches0
    while time.time() - w0 < 0.25:
        for _ in range(20):
            step()
        sync_all()
    w2 = time.time()
    clocks = sampler.stop(w0, w2)
    clocks["window"] = f"timed region ({(w1 - w0) * 1e3:.1f} ms) + untimed continuation of the same loop to {(w2 - w0) * 1e3:.0f} ms (GPU 0)"
    value = cfg["S"] * N * args.steps / (ms * 1e-3) / 1e6

    parity = None
    if not args.no_parity:
        from oracle import oracle as O
        olib = O.load("auto")
        box.reset()
        n_d = step()
        sync_all()
        nr = pk = 0.0
        rows_all = []
        for g, (dev_i, s0, ns) in enumerate(ranges):
            rows = check_rows(ns)
            idx = torch.tensor(rows, device=d_in[g].device)
            got = d_out[g].index_select(0, idx)[:, :n_d].cpu().numpy().view(np.complex64)[..., 0]
            x = d_in[g].index_select(0, idx).cpu().numpy()
            ref, _ = O.chain_run(x, cfg["in_fmt"], cfg.get("blkconv"), cfg.get("stage_a"), cfg.get("stage_b"), lib=olib,
                                 n_threads=max(1, min(len(rows), host_threads())))
            if got.shape != ref.shape:
                nr = pk = float("inf")
                continue
            for k in range(len(rows)):
                a, b = F.err_metrics(got[k], ref[k])
                nr, pk = max(nr, a), max(pk, b)
            rows_all.append([s0 + r for r in rows])
        parity = {"rows": rows_all, "devices": G, "outputs_per_stream": int(n_d), "nrms_max": nr, "peak_max": pk,
                  "tol": {"nrms": F.NRMS_TOL, "peak": F.PEAK_TOL}, "ok": bool(nr <= F.NRMS_TOL and pk <= F.PEAK_TOL),
                  "oracle": olib.sfeo_kind().decode()}

    e2e = None
    if not args.no_e2e:
        S = cfg["S"]
        nb_in, nb_out = S * N * eb, S * cap * 8
        p_in, p_out = lib.sfe_pinned_alloc(nb_in), lib.sfe_pinned_alloc(nb_out)
        assert p_in and p_out, simplefe_b200.last_error()
        import ctypes as C
        h_in = np.ctypeslib.as_array(C.cast(p_in, C.POINTER(C.c_uint8)), shape=(nb_in,))
        for g, (dev_i, s0, ns) in enumerate(ranges):
            h_in[s0 * N * eb:(s0 + ns) * N * eb] = d_in[g].cpu().numpy().view(np.uint8).reshape(-1)
        for _ in range(2):
            box.process_host_raw(p_in, N, N, p_out, cap)
        k_e2e = max(2, min(args.steps, 5))
        t0 = time.perf_counter()
        for _ in range(k_e2e):
            n_h = box.process_host_raw(p_in, N, N, p_out, cap)
        dt = time.perf_counter() - t0
        e2e = {"value": S * N * k_e2e / dt / 1e6, "unit": METRIC, "h2d_bytes_per_step": int(nb_in), "d2h_bytes_per_step": int(S * n_h * 8),
               "steps": k_e2e, "api": "sfe_box_process_host (pinned host in/out; per GPU: worker thread, 2 CUDA streams, stream groups double-buffered)"}
        e2e["h2d_achieved_gbs"] = e2e["value"] * 1e6 * eb / 1e9
        lib.sfe_pinned_free(p_in)
        lib.sfe_pinned_free(p_out)

    peak, peak_src = measured_peaks()
    bps = bytes_alg_per_sample(name, cfg, N, n_out)
    line = {
        "metric": METRIC, "value": value, "unit": METRIC, "n_gpus": G, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": workload_config(name, cfg, G, "strong"),
        "launcher": "one process: sfe_box_* (library-side stream partition, one host worker thread per GPU)",
        "plan": box.plan, "outputs_per_stream": int(n_out), "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "peak": peak, "unit": "GB/s", "peak_source": peak_src, "bytes_alg_per_sample": bps,
                     "whole_step_frac": bps * value * 1e6 / G / 1e9 / peak, "achieved": bps * value * 1e6 / G / 1e9,
                     "frac": bps * value * 1e6 / G / 1e9 / peak, "traffic": None,
                     "note": "whole step per GPU (no per-kernel brackets in box mode)"},
        "parity": parity, "host_enqueue_us_per_step": host_us, "cpu_baseline": None,
    }
    emit(line)


# ---------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="cfg5", choices=sorted(WORKLOADS), help="BASELINE.json config (cfg5 = the headline workload)")
    ap.add_argument("--as-worded", action="store_true", help="cfg1..cfg4: the single-stream shape BASELINE.json words (launch bound) "
                                                              "instead of the batched one")
    ap.add_argument("--streams", type=int, default=0, help="total streams (default: the workload's)")
    ap.add_argument("--samples", type=int, default=0)
    ap.add_argument("--in-fmt", default="", choices=["", "i16", "u8", "f32"], help="override the workload's input format "
                    "(cfg1 i16 = the int16 IQ flavour of blkconv, 12 B/sample; cfg5 u8 = the device wire format)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-weak", action="store_true", help="N > 1: skip the secondary weak-scaling figure")
    ap.add_argument("--profile-every", type=int, default=10,
                    help="bracket the kernels of every n-th timed step with CUDA events (0 = never).  A bracketed step cannot chain "
                         "its launch behind the previous call's (the event record sits between them), so it is a few us slower")
    ap.add_argument("--flags", type=int, default=0, help="SFE_FLAG_* for the chain (debug)")
    ap.add_argument("--box", action="store_true",
                    help="N GPUs from ONE process through sfe_box_* (a host worker thread per GPU inside the library) instead of "
                         "one torchrun rank per GPU; same workload, same keys")
    ap.add_argument("--scaling", default="strong", choices=["weak", "strong"],
                    help="strong (default, SURVEY.md 8-e): the workload's streams are split over the N GPUs; "
                         "weak: every GPU runs the full shard (box = N x streams)")
    args = ap.parse_args()
    capture_stdout()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    name = args.config
    fx, S_b, N_b, S_w, N_w, _ = WORKLOADS[name]
    S_def, N_def = (S_w, N_w) if args.as_worded else (S_b, N_b)
    cfg = dict(F.CONFIGS[fx], S=args.streams or S_def, N=args.samples or N_def)
    if args.in_fmt:
        cfg["in_fmt"] = {"i16": F.IN_I16, "u8": F.IN_U8, "f32": F.IN_F32}[args.in_fmt]

    if args.impl == "reference":
        run_reference(args, name, cfg, rank, world)
        return

    if args.box:
        run_box(args, name, cfg)
        return

    import torch
    import torch.distributed as dist
    import simplefe_b200
    from simplefe_b200.chain import Chain

    assert world == args.gpus or world == 1, f"WORLD_SIZE={world} but --gpus {args.gpus}"
    n_ranks = world
    torch.cuda.set_device(local_rank)
    if n_ranks > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    lib = simplefe_b200.lib()
    simplefe_b200._lib.check(lib.sfe_set_device(local_rank), "sfe_set_device")
    dev = torch.device("cuda", local_rank)
    N = cfg["N"]
    eb = EB[cfg["in_fmt"]]
    from simplefe_b200.shard import shard_range

    def barrier():
        if n_ranks > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if n_ranks > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def measure(S, seed, scaling_label, with_clocks):
        """device-resident throughput of S streams on this rank: (ms over K steps max over ranks, details)"""
        d_in = device_rows(torch, cfg, S, N, dev, seed)
        ch = Chain(S, cfg["in_fmt"], cfg.get("blkconv"), cfg.get("stage_a"), cfg.get("stage_b"), flags=args.flags, device=local_rank)
        cap = max(ch.max_out(N), 1)
        d_out = torch.zeros((S, cap, 2), dtype=torch.float32, device=dev)
        stream = torch.cuda.current_stream()
        small = S * N * eb <= (256 << 20)       # input does not exceed L2 by a wide margin: flush between timed steps

        def step():   # one pass of the hot path over one batch: the next N samples of every stream
            return ch.process_dev(d_in.data_ptr(), N, N, d_out.data_ptr(), cap, stream.cuda_stream)

        # The clock sampler needs 0.1-3 s to deliver its first line: it is started BEFORE the warm-up steps, so that the GPUs
        # do not sit idle (and drop their clocks) between the warm-up and the timed region -- at 0.1 ms per step a timed
        # region of K = 20 steps is 2 ms, shorter than a clock ramp.
        sampler = ClockSampler(local_rank)
        if rank == 0 and with_clocks:
            sampler.start()
        barrier()
        # every tenth step is bracketed with events for the roofline; the others run as a caller's would (a bracketed
        # step costs two event records and cannot start as a programmatic dependent of the previous call's kernel)
        ch.profile(0)
        for _ in range(args.warmup):
            n_out = step()
        ch.profile(args.profile_every)
        launches0 = lib.sfe_kernel_launches()
        barrier()
        w0 = time.time()
        host_s = 0.0
        if small:
            evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
            for a, b in evs:
                flush.fill_(1)
                a.record(stream)
                h0 = time.perf_counter()
                n_out = step()
                host_s += time.perf_counter() - h0
                b.record(stream)
            barrier()
            ms = sum(a.elapsed_time(b) for a, b in evs)
        else:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            h0 = time.perf_counter()
            for _ in range(args.steps):
                n_out = step()
            host_s = time.perf_counter() - h0
            e1.record(stream)
            barrier()
            ms = e0.elapsed_time(e1)
        w1 = time.time()
        launches = lib.sfe_kernel_launches() - launches0
        ch.profile(False)
        # K short steps can be shorter than one nvidia-smi period: the same loop keeps running, untimed, until the
        # sampler has seen at least 0.25 s of this load; the clocks line covers the timed region plus that continuation
        clocks = None
        if rank == 0 and with_clocks:
            while time.time() - w0 < 0.25:
                for _ in range(20):
                    step()
                torch.cuda.synchronize()
            w2 = time.time()
            clocks = sampler.stop(w0, w2)
            clocks["window"] = f"timed region ({(w1 - w0) * 1e3:.1f} ms) + untimed continuation of the same loop to {(w2 - w0) * 1e3:.0f} ms"
        barrier()
        prof = ch.profile_read()
        rank_ms = [ms]
        if n_ranks > 1:           # every rank's own device time over the K steps: shows whether one rank is the straggler
            tt = torch.zeros(n_ranks, dtype=torch.float64, device=dev)
            tt[rank] = ms
            dist.all_reduce(tt, op=dist.ReduceOp.SUM)
            rank_ms = [float(x) for x in tt.tolist()]
        return dict(ms=max_ranks(ms), rank_ms=rank_ms, ch=ch, d_in=d_in, d_out=d_out, cap=cap, n_out=n_out, launches=launches, prof=prof,
                    clocks=clocks, host_us=host_s / args.steps * 1e6, small=small)

    # ---- the workload: streams split over the ranks (strong), or the full shard everywhere (weak) ---------------
    if args.scaling == "strong":
        assert cfg["S"] % n_ranks == 0, "the strong-scaling bench expects a stream count divisible by the GPU count"
        lo, hi = shard_range(cfg["S"], rank, n_ranks)
        S = hi - lo
        S_total = cfg["S"]
    else:
        lo, S = rank * cfg["S"], cfg["S"]
        S_total = S * n_ranks
    m = measure(S, 5000 + lo, args.scaling, True)
    ch, d_in, d_out, cap, n_out = m["ch"], m["d_in"], m["d_out"], m["cap"], m["n_out"]
    value = S_total * N * args.steps / (m["ms"] * 1e-3) / 1e6
    plan = ch.plan
    stream = torch.cuda.current_stream()

    def step():
        return ch.process_dev(d_in.data_ptr(), N, N, d_out.data_ptr(), cap, stream.cuda_stream)

    # ---- parity of what was just timed: a sample of this rank's rows against the reference's CPU code ---------------
    parity = None
    if not args.no_parity:
        from oracle import oracle as O
        olib = O.load("auto")
        rows = check_rows(S)
        ch.reset()
        n_d = step()
        torch.cuda.synchronize()
        idx = torch.tensor(rows, device=dev)
        got = d_out.index_select(0, idx)[:, :n_d].cpu().numpy().view(np.complex64)[..., 0]
        x = d_in.index_select(0, idx).cpu().numpy()
        ref, _ = O.chain_run(x, cfg["in_fmt"], cfg.get("blkconv"), cfg.get("stage_a"), cfg.get("stage_b"), lib=olib,
                             n_threads=max(1, min(len(rows), host_threads() // max(n_ranks, 1))))
        same_shape = got.shape == ref.shape
        nr = pk = float("inf")
        if same_shape:
            errs = [F.err_metrics(got[k], ref[k]) for k in range(len(rows))] if got.shape[1] else [(0.0, 0.0)]
            nr, pk = max(e[0] for e in errs), max(e[1] for e in errs)
        nr, pk = max_ranks(nr), max_ranks(pk)
        parity = {"rows": rows, "ranks": n_ranks, "streams_per_rank": S, "outputs_per_stream": int(n_d),
                  "nrms_max": nr, "peak_max": pk, "tol": {"nrms": F.NRMS_TOL, "peak": F.PEAK_TOL},
                  "ok": bool(nr <= F.NRMS_TOL and pk <= F.PEAK_TOL), "oracle": olib.sfeo_kind().decode(),
                  "what": "the timed launch shape (same chain object, same rows, one call from a fresh stream state) vs the reference classes"}

    # ---- end to end through the host-buffer C-ABI call --------------------------------------------
    e2e = None
    if not args.no_e2e:
        h_in = torch.empty(d_in.shape, dtype=d_in.dtype, pin_memory=True)
        h_in.copy_(d_in)
        h_out = torch.zeros((S, cap, 2), dtype=torch.float32, pin_memory=True)
        torch.cuda.synchronize()

        def step_host():
            return ch.process_host_raw(h_in.data_ptr(), N, N, h_out.data_ptr(), cap)

        for _ in range(2):
            step_host()
        k_e2e = max(2, min(args.steps, 5))
        barrier()
        t0 = time.perf_counter()
        for _ in range(k_e2e):
            n_h = step_host()
        torch.cuda.synchronize()
        dt = max_ranks(time.perf_counter() - t0)
        e2e = {"value": S_total * N * k_e2e / dt / 1e6, "unit": METRIC,
               "h2d_bytes_per_step": int(S * N * eb), "d2h_bytes_per_step": int(S * n_h * 8), "steps": k_e2e,
               "api": "sfe_chain_process_host (pinned host in/out, stream groups double-buffered over 2 CUDA streams)"}
        # what the box can do at best: one plain pinned H2D copy of the same bytes per GPU, all GPUs at once
        barrier()
        t0 = time.perf_counter()
        for _ in range(3):
            d_in.copy_(h_in, non_blocking=True)
        torch.cuda.synchronize()
        dtc = max_ranks(time.perf_counter() - t0)
        ceil_gbs = 3 * S * N * eb * n_ranks / dtc / 1e9
        e2e["h2d_ceiling_gbs"] = ceil_gbs
        e2e["h2d_ceiling_what"] = f"{n_ranks} concurrent cudaMemcpyAsync H2D of {S * N * eb >> 20} MiB each from pinned memory, aggregate"
        e2e["h2d_achieved_gbs"] = e2e["value"] * 1e6 * eb / 1e9
        e2e["frac_of_h2d_ceiling"] = e2e["h2d_achieved_gbs"] / ceil_gbs
        # the device-resident and the host path must agree from the same stream state
        ch.reset()
        n_d = step()
        torch.cuda.synchronize()
        ref_dev = d_out.cpu()[:, :n_d].clone()
        ch.reset()
        n_h2 = step_host()
        # (to rounding: the host path launches stream groups, whose work items are cut at other places than the full launch's)
        if n_d == n_h2 and n_d > 0:
            dn, dp = F.err_metrics(h_out[:, :n_h2].numpy().view(np.complex64)[..., 0], ref_dev.numpy().view(np.complex64)[..., 0])
        else:
            dn, dp = (0.0, 0.0) if n_d == n_h2 else (float("inf"), float("inf"))
        e2e["matches_device_path"] = bool(n_d == n_h2 and dn <= 1e-6 and dp <= 5e-6)
        e2e["vs_device_path"] = {"nrms": dn, "peak": dp, "bit_identical": bool(n_d == n_h2 and torch.equal(h_out[:, :n_h2], ref_dev))}
        del h_in, h_out

    # ---- secondary figure at N > 1: the other scaling mode, device resident only ---------------------------------
    other = None
    if n_ranks > 1 and not args.no_weak:
        del d_in, d_out, ch
        m["ch"] = m["d_in"] = m["d_out"] = None
        torch.cuda.empty_cache()
        if args.scaling == "strong":
            mo = measure(cfg["S"], 7000 + rank * cfg["S"], "weak", False)
            tot = cfg["S"] * n_ranks
        else:
            lo2, hi2 = shard_range(cfg["S"], rank, n_ranks)
            mo = measure(hi2 - lo2, 7000 + lo2, "strong", False)
            tot = cfg["S"]
        other = {"scaling": "weak" if args.scaling == "strong" else "strong", "value": tot * N * args.steps / (mo["ms"] * 1e-3) / 1e6,
                 "unit": METRIC, "streams_total": tot, "ms_per_step": mo["ms"] / args.steps}
        mo["ch"] = mo["d_in"] = mo["d_out"] = None

    if rank != 0:
        if n_ranks > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel ------------------------------------------------------------
    peak, peak_src = measured_peaks()
    prof = m["prof"]
    dom = max(prof, key=lambda p: p[1]) if prof else ("", 0.0, 0)
    bps = bytes_alg_per_sample(name, cfg, N, n_out)
    ach = bps * S * N / (dom[1] * 1e-3) / 1e9 if dom[1] > 0 else None
    # the capture is per launch of the full single-GPU workload; other shard sizes have no capture
    tr = ncu_traffic(name, plan) if args.flags == 0 and S == WORKLOADS[name][1] and N == WORKLOADS[name][2] and not args.as_worded else None
    roofline = {"bound": "hbm", "kernel": dom[0], "kernel_ms": dom[1], "launches_timed": dom[2], "timed_every": args.profile_every,
                "achieved": ach, "peak": peak, "unit": "GB/s", "frac": (ach / peak) if ach else None,
                "traffic": (tr or {}).get("dram_bytes_per_launch"), "traffic_source": (tr or {}).get("source"),
                "peak_source": peak_src, "bytes_alg_per_sample": bps,
                "note": "achieved = bytes_alg_per_sample x the samples one launch of this kernel covers / its mean launch duration "
                        "(a multi-kernel plan's later stages read intermediates: see whole_step_frac).  kernel_ms comes from the "
                        "bracketed steps, whose launch cannot chain behind the previous call's kernel: it includes the launch "
                        "latency the other steps hide, so it can exceed ms_per_step by a few us",
                "kernels": [{"name": n, "ms": ms_, "launches": k} for n, ms_, k in prof],
                "whole_step_frac": bps * value * 1e6 / n_ranks / 1e9 / peak}

    # ---- CPU baseline: the reference's own code on this box's cores ------------------------------
    cpu = None
    if not args.no_cpu and n_ranks == 1:
        try:
            cpu = cpu_baseline(cfg, host_rows(cfg))
        except Exception as ex:  # the baseline is a report, not a dependency of the GPU numbers
            cpu = {"value": None, "error": repr(ex)}

    line = {
        "metric": METRIC, "value": value, "unit": METRIC, "n_gpus": n_ranks, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": m["ms"] / args.steps, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": workload_config(name, cfg, n_ranks, args.scaling),
        "plan": plan, "outputs_per_stream": int(n_out),
        "clocks": m["clocks"], "e2e": e2e, "gpu_launches": int(m["launches"]), "roofline": roofline, "parity": parity,
        "host_enqueue_us_per_step": m["host_us"], "rank_ms_per_step": [x / args.steps for x in m["rank_ms"]], "cpu_baseline": cpu,
    }
    if other is not None:
        line[other["scaling"]] = other
    emit(line)
    if n_ranks > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
