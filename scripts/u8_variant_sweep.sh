#!/bin/bash
# K3 on the u8 wire format: warps per CTA x conversion flavour (SFE_BANK_VARIANT = warps,slots,conv16,conv8)
for v in 8,2,2,12 8,2,2,10 8,2,2,11 10,2,2,12 12,2,2,12 12,2,2,10 16,2,2,12 16,2,2,10; do
  echo "variant $v"
  SFE_BANK_VARIANT=$v python bench_stages.py --only "u8 device" 2>&1 | grep -o '"ms_per_step": [0-9.]*\|"frac_of_measured_hbm": [0-9.]*\|"ms": [0-9.]*' | paste - - -
done
