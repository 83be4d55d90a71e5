#!/bin/bash
# One ncu --set full capture per stage kernel (bench_stages.py cases). Run under gpurun; reports in gpurun_out/.
# usage: scripts/ncu_stages.sh tag "case substring" kernel-regex [more triples...]
tag=$1; shift
while [ $# -ge 2 ]; do
  only=$1; rx=$2; shift 2
  name=$(echo "$rx" | tr -c 'a-zA-Z0-9' '_')
  python bench_stages.py --quick --only "$only" > gpurun_out/${tag}_${name}_plain.json 2> gpurun_out/${tag}_${name}_plain.err &&
  ncu --set full --clock-control none --import-source on -k regex:$rx -s 2 -c 1 -f -o gpurun_out/${tag}_${name} \
      python bench_stages.py --quick --only "$only" > gpurun_out/${tag}_${name}_ncu.log 2>&1
  echo "$only / $rx -> rc $?"
done
