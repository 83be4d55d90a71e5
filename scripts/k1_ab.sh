#!/bin/bash
# GPU box: A/B of the K1 tuning builds made by scripts/k1_variants.sh.   scripts/k1_ab.sh <tag> ...   (results: gpurun_out/k1ab_<tag>_*.json)
for tag in "$@"; do
  cp simplefe_b200/build/variants/libsfegpu_$tag.so simplefe_b200/libsfegpu.so
  python bench.py --config cfg1 --no-cpu --no-e2e --steps 20 > gpurun_out/k1ab_${tag}_c64.json 2>> gpurun_out/k1ab.err
  python bench.py --config cfg1 --in-fmt i16 --streams 512 --no-cpu --no-e2e --steps 20 > gpurun_out/k1ab_${tag}_i16.json 2>> gpurun_out/k1ab.err
  python bench.py --config cfg4 --no-cpu --no-e2e --steps 20 > gpurun_out/k1ab_${tag}_cfg4.json 2>> gpurun_out/k1ab.err
done
python scripts/bench_summary.py gpurun_out/k1ab_*.json
