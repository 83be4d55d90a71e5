#!/bin/bash
# One `ncu --set full` capture of one kernel of a bench.py command (after the command has run clean without ncu).
#   scripts/ncu_capture.sh <tag> <kernel-regex> <launch-skip> <bench.py args...>     (GPU box, one GPU)
# writes gpurun_out/<tag>.ncu-rep and gpurun_out/<tag>_summary.json (scripts/ncu_summary.py)
tag=$1; rx=$2; skip=$3; shift 3
python bench.py "$@" --no-cpu --no-e2e --no-parity --steps 3 > gpurun_out/${tag}_plain.json 2> gpurun_out/${tag}_plain.err || { echo "$tag: plain run failed"; tail -3 gpurun_out/${tag}_plain.err; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:$rx -s $skip -c 1 -f -o gpurun_out/$tag \
    python bench.py "$@" --no-cpu --no-e2e --no-parity --steps 3 > gpurun_out/${tag}_ncu.log 2>&1
python scripts/ncu_summary.py gpurun_out/$tag.ncu-rep > gpurun_out/${tag}_summary.json 2>/dev/null
# gpurun copies at most 64 MiB back: the report itself (3-10 MB) is kept only on request (KEEP_REP=1)
[ "${KEEP_REP:-0}" = 1 ] || rm -f gpurun_out/$tag.ncu-rep
python - <<PY
import json
d=json.load(open("gpurun_out/${tag}_summary.json"))
for k,v in d.items():
    print("$tag", v.get("kernel","")[:70], "ms", v.get("gpu__time_duration.sum"), "issue%", v.get("smsp__issue_active.avg.pct_of_peak_sustained_active"), "fma%", v.get("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active"), "l1%", v.get("l1tex__throughput.avg.pct_of_peak_sustained_elapsed"), "dram%", v.get("dram__throughput.avg.pct_of_peak_sustained_elapsed"), "tensor%", v.get("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"))
PY
