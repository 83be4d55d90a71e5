#!/usr/bin/env python
"""bench.py -- headline benchmark of the libdsp hot path on B200.

Workload (BASELINE.json configs[4], SURVEY.md 8-d cfg 5): 1024 independent int16 IQ streams x 2^20
samples through blkconv(255 taps, FFT 4096) -> decimate(128 taps, /8) -> resample(512 taps, L=24,
rate 625/24), one shared filter set, streams partitioned over the N GPUs (no collective: the path
never exchanges data between streams).  Metric: Msamples/s, one sample = one complex input sample.

    python bench.py [--gpus N --steps K --warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One JSON line on stdout (rank 0).  `value` = device-resident throughput (CUDA events, max over
ranks); `e2e` = the same through the host-buffer C-ABI call (pinned host in, pinned host out,
copies inside the timed region); `roofline` = the dominant kernel against the measured HBM peak;
`cpu_baseline` = the reference's own CPU code (oracle/_ref) on this box's host cores.
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


def ncu_traffic():
    """dram bytes per launch of the dominant kernel from the committed ncu capture, if any"""
    try:
        d = json.load(open(os.path.join(ROOT, "profiles", "dominant_kernel_traffic.json")))
        return d
    except Exception:
        return None


# ---------------------------------------------------------------------------------------------------
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


def cpu_baseline(cfg, make_rows, budget_s=12.0):
    """The reference's CPU path on a bounded sample of the same workload, all host threads."""
    T = host_threads()
    N = cfg["N"]
    t1, _, lib = cpu_chain(cfg, make_rows(1), 1)            # one stream on one thread: calibration
    per_thread = max(1, int(budget_s / max(t1, 1e-3)))
    S = int(min(cfg["S"], max(T, per_thread * T)))
    S = max(T, S // T * T) if S >= T else S
    iq = make_rows(S)
    dt, _, _ = cpu_chain(cfg, iq, T)
    v = S * N / dt / 1e6
    return {"value": v, "unit": METRIC, "cores": T, "kind": lib.sfeo_kind().decode(),
            "sample": f"{S} of {cfg['S']} streams x {N} samples, {T} threads, {dt:.2f} s; 1 thread: {N / t1 / 1e6:.2f} {METRIC}",
            "build": lib.sfeo_build_flags().decode(), "one_thread_value": N / t1 / 1e6}


# ---------------------------------------------------------------------------------------------------
def run_reference(args, cfg, rank, world):
    """--impl reference: the reference's own CPU implementation (oracle/_ref), rank 0 only."""
    if rank != 0:
        return
    T = host_threads()
    rng = np.random.Generator(np.random.PCG64(cfg["seed"]))
    base = F._quant16(np.clip(rng.normal(0.0, 0.25, size=(cfg["N"] + 4096, 2)), -1, 1))

    def make_rows(S):
        return np.stack([base[(37 * s) % 4096:(37 * s) % 4096 + cfg["N"]] for s in range(S)])

    t1, _, lib = cpu_chain(cfg, make_rows(1), 1)
    S = int(min(cfg["S"], max(T, int(8.0 / max(t1, 1e-3)) * T)))
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
        "config": workload_config(cfg, args.gpus),
        "cpu_baseline": {"value": v, "unit": METRIC, "cores": T, "kind": lib.sfeo_kind().decode(),
                         "sample": f"{S} of {cfg['S']} streams x {cfg['N']} samples per step, {T} threads",
                         "build": lib.sfeo_build_flags().decode()},
        "e2e": {"value": v, "unit": METRIC, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def workload_config(cfg, n_gpus, streams_per_gpu=None, scaling="weak"):
    spg = streams_per_gpu if streams_per_gpu is not None else cfg["S"]
    return {"workload": "cfg5: 1024 int16 IQ streams x 2^20 samples per GPU, blkconv(255 taps, FFT 4096) -> decimate(128 taps, /8) "
                        "-> resample(512 taps, L=24, rate 625/24)",
            "streams_total": spg * n_gpus, "samples_per_stream": cfg["N"], "streams_per_gpu": spg,
            "partition": f"independent streams over {n_gpus} GPU(s) ({scaling}), no collective",
            "l2": "inputs larger than L2 (no flush needed)"}


# ---------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--streams", type=int, default=1024, help="total streams (BASELINE config: 1024)")
    ap.add_argument("--samples", type=int, default=1 << 20)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--profile-every", type=int, default=3, help="bracket the kernels of every n-th timed step with CUDA events")
    ap.add_argument("--flags", type=int, default=0, help="SFE_FLAG_* for the chain (debug)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: every GPU runs the full 1024-stream shard (box = N x 1024 streams); "
                         "strong: the 1024 streams are split over the N GPUs")
    args = ap.parse_args()
    capture_stdout()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    cfg = dict(F.CONFIGS["cfg5"], S=args.streams, N=args.samples)

    if args.impl == "reference":
        run_reference(args, cfg, rank, world)
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

    N = cfg["N"]
    # the path shards by stream with no data-path collective.  weak (default): each GPU carries the
    # full single-GPU shard of cfg["S"] streams; strong: cfg["S"] streams are split over the ranks.
    from simplefe_b200.shard import shard_range
    if args.scaling == "weak":
        S = cfg["S"]
    else:
        lo, hi = shard_range(cfg["S"], rank, n_ranks)
        S = hi - lo
        assert cfg["S"] % n_ranks == 0, "strong scaling bench expects streams divisible by the GPU count"
    S_total = S * n_ranks
    dev = torch.device("cuda", local_rank)

    # synthetic int16 IQ, generated on the device (N(0, 0.25^2) clipped, round(x * 32767)); rank-distinct seed
    g = torch.Generator(device=dev)
    g.manual_seed(5000 + rank)
    d_in = torch.empty((S, N, 2), dtype=torch.int16, device=dev)
    rows = max(1, (256 << 20) // (N * 2 * 4))
    for s0 in range(0, S, rows):
        blk = torch.randn((min(rows, S - s0), N, 2), generator=g, device=dev) * 0.25
        d_in[s0:s0 + rows] = torch.round(blk.clamp_(-1, 1) * 32767.0).to(torch.int16)
    del blk

    ch = Chain(S, cfg["in_fmt"], cfg["blkconv"], cfg["stage_a"], cfg["stage_b"], flags=args.flags, device=local_rank)
    cap = ch.max_out(N)
    d_out = torch.zeros((S, cap, 2), dtype=torch.float32, device=dev)
    stream = torch.cuda.current_stream()

    def step():   # one pass of the hot path over one batch: the next 2^20 samples of every stream
        return ch.process_dev(d_in.data_ptr(), N, N, d_out.data_ptr(), cap, stream.cuda_stream)

    def barrier():
        if n_ranks > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput -----------------------------------------------------------------
    for _ in range(args.warmup):
        n_out = step()
    barrier()
    # every third step is bracketed with events for the roofline; the others run as a caller's would (a bracketed
    # step costs two event records and keeps the state carry from running beside the main kernel)
    ch.profile(args.profile_every)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = lib.sfe_kernel_launches()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    w0 = time.time()
    e0.record(stream)
    for _ in range(args.steps):
        n_out = step()
    e1.record(stream)
    barrier()
    w1 = time.time()
    ms = e0.elapsed_time(e1)
    launches = lib.sfe_kernel_launches() - launches0
    ch.profile(False)
    # K steps of 0.7 ms are shorter than one nvidia-smi period: the same loop keeps running, untimed, until the
    # sampler has seen at least 0.25 s of this load; the clocks line covers the timed region plus that continuation
    w2 = w1
    if rank == 0:
        while time.time() - w0 < 0.25:
            for _ in range(20):
                step()
            torch.cuda.synchronize()
        w2 = time.time()
    clocks = sampler.stop(w0, w2) if rank == 0 else None
    if clocks is not None:
        clocks["window"] = f"timed region ({(w1 - w0) * 1e3:.1f} ms) + untimed continuation of the same loop to {(w2 - w0) * 1e3:.0f} ms"
    barrier()
    ch.profile(False)
    prof = ch.profile_read()
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if n_ranks > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    value = S_total * N * args.steps / (ms_max * 1e-3) / 1e6

    # ---- end to end through the host-buffer C-ABI call --------------------------------------------
    e2e = None
    if not args.no_e2e:
        h_in = torch.empty((S, N, 2), dtype=torch.int16, pin_memory=True)
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
        dt = time.perf_counter() - t0
        tt = torch.tensor([dt], dtype=torch.float64, device=dev)
        if n_ranks > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt.item())
        e2e = {"value": S_total * N * k_e2e / dt / 1e6, "unit": METRIC,
               "h2d_bytes_per_step": int(S * N * 4), "d2h_bytes_per_step": int(S * n_h * 8), "steps": k_e2e,
               "api": "sfe_chain_process_host (pinned host in/out, 2 CUDA streams, stream groups double-buffered)"}
        # the device-resident and the host path must agree bit for bit from the same stream state
        ch.reset()
        n_d = step()
        torch.cuda.synchronize()
        ref_dev = d_out.cpu()[:, :n_d].clone()
        ch.reset()
        n_h2 = step_host()
        e2e["matches_device_path"] = bool(n_d == n_h2 and torch.equal(h_out[:, :n_h2], ref_dev))

    if rank != 0:
        if n_ranks > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel ------------------------------------------------------------
    peak, peak_src = measured_peaks()
    dom = max(prof, key=lambda p: p[1]) if prof else ("", 0.0, 0)
    per_launch_samples = S * N
    ach = BYTES_ALG_PER_SAMPLE * per_launch_samples / (dom[1] * 1e-3) / 1e9 if dom[1] > 0 else None
    tr = ncu_traffic()
    roofline = {"bound": "hbm", "kernel": dom[0], "kernel_ms": dom[1], "launches_timed": dom[2], "timed_every": args.profile_every,
                "achieved": ach, "peak": peak, "unit": "GB/s", "frac": (ach / peak) if ach else None,
                "traffic": (tr or {}).get("dram_bytes_per_launch"), "peak_source": peak_src,
                "bytes_alg_per_sample": BYTES_ALG_PER_SAMPLE,
                "kernels": [{"name": n, "ms": m, "launches": k} for n, m, k in prof],
                "whole_step_frac": BYTES_ALG_PER_SAMPLE * value * 1e6 / n_ranks / 1e9 / peak}

    # ---- CPU baseline: the reference's own code on this box's cores ------------------------------
    cpu = None
    if not args.no_cpu and n_ranks == 1:
        try:
            src = d_in[: min(S, 1024)].cpu().numpy()
            cpu = cpu_baseline(cfg, lambda k: src[np.arange(k) % len(src)])
        except Exception as ex:  # the baseline is a report, not a dependency of the GPU numbers
            cpu = {"value": None, "error": repr(ex)}

    line = {
        "metric": METRIC, "value": value, "unit": METRIC, "n_gpus": n_ranks, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": dict(workload_config(cfg, n_ranks, S, args.scaling), plan=ch.plan, outputs_per_stream=n_out),
        "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu,
    }
    emit(line)
    if n_ranks > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
