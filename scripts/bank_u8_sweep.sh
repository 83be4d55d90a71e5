#!/bin/bash
# K3 on the u8 wire format: SFE_BANK_VARIANT_U8 = warps,conversion flavour,boxes per ring slot.  Run under gpurun.
# The sweep flavours are compiled only with SFE_BANK_SWEEP=1 (python -m simplefe_b200.build): build that library first.
export SFE_BANK_SWEEP=1; python -m simplefe_b200.build --force >/dev/null || exit 1
for v in 16,12,1 16,10,1 12,12,1 8,12,2 8,10,2; do
  echo "variant $v"
  SFE_BANK_VARIANT_U8=$v python bench_stages.py --only "u8 device" 2>&1 | grep -o '"ms_per_step": [0-9.]*\|"frac_of_measured_hbm": [0-9.]*\|"ms": [0-9.]*' | head -3 | paste - - -
done
