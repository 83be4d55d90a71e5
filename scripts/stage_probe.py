#!/usr/bin/env python
"""stage_probe.py -- ad-hoc timing of one chain configuration: cfg name, input format, streams, samples, flags.
usage: scripts/stage_probe.py cfg1 i16|u8|f32 S N [flags]   (prints plan, ms per step, kernel times)"""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import fixtures as F
import bench_stages as B
import torch
from simplefe_b200.chain import Chain
name, fmt, S, N = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
flags = int(sys.argv[5]) if len(sys.argv) > 5 else 0
cfg = dict(F.CONFIGS[name], in_fmt={"i16": F.IN_I16, "u8": F.IN_U8, "f32": F.IN_F32}[fmt])
r = B.run_case(f"{name} {fmt} x{S} N={N} flags={flags}", cfg, S, N, flags, 10, 3, 6531.3, torch, Chain)
print(json.dumps({k: r[k] for k in ("case", "plan", "ms_per_step", "Msamples_per_s", "frac_of_measured_hbm", "kernels")}))
