#!/bin/bash
# GPU box: time one bench.py workload with each tuning build.   scripts/lib_ab.sh "<bench args>" <tag> ...
args=$1; shift
for tag in "$@"; do
  cp simplefe_b200/build/variants/libsfegpu_$tag.so simplefe_b200/libsfegpu.so
  python bench.py $args --no-cpu --no-e2e --steps 20 > gpurun_out/ab_${tag}.json 2>> gpurun_out/ab.err
done
python scripts/bench_summary.py gpurun_out/ab_*.json
