#!/bin/bash
# Tuning sweep of the fused bank kernel (K3): SFE_BANK_VARIANT = warps,slots,conv.  Run under gpurun.
# The sweep flavours are compiled only with SFE_BANK_SWEEP=1 (python -m simplefe_b200.build): build that library first.
export SFE_BANK_SWEEP=1; python -m simplefe_b200.build --force >/dev/null || exit 1
out=gpurun_out/bank_sweep.txt
: > $out
for v in ${VARIANTS:-8,4,0 12,2,0 12,3,0 16,2,0 20,2,0 16,2,1 16,2,2 16,2,3 12,3,1 12,3,2 12,3,3 20,2,3 20,2,2}; do
  SFE_BANK_VARIANT=$v python bench.py --steps 10 --warmup 3 --no-cpu --no-e2e 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); r=d['roofline']
print('$v', 'value', round(d['value']/1e3,1), 'Gs/s  ms_step', round(d['ms_per_step'],4), 'kernel_ms', round(r['kernel_ms'],4), 'frac', round(r['frac'],3), 'whole', round(r['whole_step_frac'],3))" >> $out 2>&1 || echo "$v FAILED" >> $out
done
cat $out
