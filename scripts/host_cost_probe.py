#!/usr/bin/env python
"""host_cost_probe.py -- host time of one sfe_chain_process_dev call (no synchronise) against the whole step, cfg 4 single stream"""
import os, sys, time
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import numpy as np, torch
import fixtures as F
from simplefe_b200.chain import Chain
cfg = dict(F.CONFIGS["cfg4"], S=1)
iq = torch.from_numpy(F.make_input(cfg)).cuda()
ch = Chain(1, cfg["in_fmt"], cfg.get("blkconv"), cfg.get("stage_a"), cfg.get("stage_b"))
N = cfg["N"]; cap = ch.max_out(N)
out = torch.zeros((1, cap, 2), dtype=torch.float32, device="cuda")
st = torch.cuda.current_stream()
for _ in range(5): ch.process_dev(iq.data_ptr(), N, N, out.data_ptr(), cap, st.cuda_stream)
torch.cuda.synchronize()
host, total = [], []
for _ in range(20):
    t0 = time.perf_counter(); ch.process_dev(iq.data_ptr(), N, N, out.data_ptr(), cap, st.cuda_stream); t1 = time.perf_counter()
    torch.cuda.synchronize(); t2 = time.perf_counter()
    host.append(t1 - t0); total.append(t2 - t0)
print("host ms median %.3f  step ms median %.3f  plan %s" % (1e3 * np.median(host), 1e3 * np.median(total), ch.plan))
