#!/usr/bin/env python
"""ncu_traffic.py -- regenerate profiles/dominant_kernel_traffic.json from `ncu --set full` captures made by
scripts/ncu_capture.sh (DRAM bytes read + written per launch of the dominant kernel of a bench.py workload).
bench.py quotes an entry as roofline.traffic only when the workload AND the kernel plan of the run match the capture's,
so a capture of an older build can never be reported for a newer kernel.

usage: scripts/ncu_traffic.py cfg5=r02_k3 cfg1=r02_k1_c64 ...     (tags of gpurun_out/<tag>_summary.json / <tag>_plain.json)
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
out_path = os.path.join(ROOT, "profiles", "dominant_kernel_traffic.json")
try:
    out = json.load(open(out_path))
    if "dram_bytes_per_launch" in out:      # the round-1 single-entry layout
        out = {}
except Exception:
    out = {}
for arg in sys.argv[1:]:
    name, tag = arg.split("=")
    summ = json.load(open(os.path.join(ROOT, "gpurun_out", tag + "_summary.json")))
    plain = json.loads(open(os.path.join(ROOT, "gpurun_out", tag + "_plain.json")).readline())
    (rep, v), = summ.items()
    rd = v["dram__bytes_read.sum"] * UNIT[v["dram__bytes_read.sum [unit]"]]
    wr = v["dram__bytes_write.sum"] * UNIT[v["dram__bytes_write.sum [unit]"]]
    cfg = plain["config"]
    alg = plain["roofline"]["bytes_alg_per_sample"] * cfg["streams_per_gpu"] * cfg["samples_per_stream"]
    out[name] = {"kernel": v["kernel"], "plan": plain["plan"], "workload": cfg["workload"], "streams": cfg["streams_per_gpu"],
                 "samples_per_stream": cfg["samples_per_stream"], "dram_bytes_per_launch": rd + wr, "dram_bytes_read": rd,
                 "dram_bytes_write": wr, "algorithmic_bytes_per_launch": alg, "ratio": (rd + wr) / alg,
                 "source": f"profiles/{tag}_summary.json (ncu --set full --clock-control none, one launch; scripts/ncu_capture.sh)"}
    print(name, out[name]["kernel"][:60], "traffic / algorithmic =", round(out[name]["ratio"], 4))
json.dump(out, open(out_path, "w"), indent=1)
