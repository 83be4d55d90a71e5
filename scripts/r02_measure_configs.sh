#!/bin/bash
# Reproducible measurements of every BASELINE.json config through bench.py (same harness, same keys: clocks, e2e,
# roofline, parity, same-run cpu_baseline).  Run on the GPU box:  gpurun -- scripts/r02_measure_configs.sh [tag]
# One JSON line per case in gpurun_out/<tag>_configs.jsonl
tag=${1:-r02}
out=gpurun_out/${tag}_configs.jsonl
: > $out
run() { echo "== $*" >&2; python bench.py "$@" >> $out 2>> gpurun_out/${tag}_configs.err || echo "{\"failed\": \"$*\"}" >> $out; }
run --config cfg1 --steps 10
run --config cfg1 --steps 10 --as-worded --no-cpu
run --config cfg2 --steps 10
run --config cfg2 --steps 10 --flags 4 --no-cpu          # CUDA-core FIR instead of the tcgen05 Toeplitz kernel
run --config cfg2 --steps 10 --as-worded --no-cpu
run --config cfg3 --steps 10
run --config cfg3 --steps 10 --as-worded --no-cpu
run --config cfg4 --steps 10
run --config cfg4 --steps 10 --as-worded --no-cpu
run --config cfg5 --steps 20
python scripts/bench_summary.py $out >&2
