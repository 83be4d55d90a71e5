#!/usr/bin/env python
"""one line per bench.py JSON file: the fields a tuning loop looks at"""
import json
import sys

def lines(paths):
    for path in paths:
        try:
            for k, l in enumerate(open(path)):
                if l.strip():
                    yield f"{path}:{k}", json.loads(l)
        except Exception as ex:
            print(path, "unreadable:", ex)


for path, d in lines(sys.argv[1:]):
    if "value" not in d:
        print(path, d)
        continue
    r, e, p, c = d.get("roofline") or {}, d.get("e2e") or {}, d.get("parity") or {}, d.get("clocks") or {}
    print(path.split("/")[-1], (d.get("config") or {}).get("workload", "")[:5], "S", (d.get("config") or {}).get("streams_total"), "| value", round(d["value"]), "ms/step", round(d["ms_per_step"], 4), "kernel_ms", round(r.get("kernel_ms") or 0, 4),
          "frac", round(r.get("frac") or 0, 3), "whole", round(r.get("whole_step_frac") or 0, 3), "host_us", round(d.get("host_enqueue_us_per_step") or 0, 1),
          "| parity", p.get("ok"), "%.2e" % (p.get("nrms_max") or 0), "| e2e", round(e.get("value") or 0), "of ceiling", round(e.get("frac_of_h2d_ceiling") or 0, 3),
          "match", e.get("matches_device_path"), (e.get("vs_device_path") or {}).get("nrms"), "| clk", c.get("sm_mhz"), c.get("reasons"))
