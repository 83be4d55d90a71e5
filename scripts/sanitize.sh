#!/bin/bash
# compute-sanitizer over a small -m gpu subset that reaches every kernel family: K1 (shared-memory FFT exchanges), K2d/K2g/K2s,
# K2t (tcgen05, TMEM alloc/dealloc), K3 (TMA + mbarrier ring, seam exchange), save_tail / expand_sched, the pinned ring.
# Run on the GPU box: gpurun -- scripts/sanitize.sh ; summaries land in gpurun_out/r02_sanitizer_{memcheck,racecheck}.txt
SEL='test_chain_vs_golden or (test_chain_tensor_path_vs_oracle and 4099) or (test_chain_bank_path_vs_oracle and 4099) or test_blkconv_known_answer or test_dropin_classes or test_chain_cfg4_many_streams or (test_blkconv_any_fft_len_vs_oracle and 1000)'
for tool in memcheck racecheck; do
  out=gpurun_out/r02_sanitizer_$tool
  timeout 1500 compute-sanitizer --tool $tool --print-limit 20 --log-file $out.log \
      python -m pytest tests -m gpu -q -x -k "$SEL" > $out.pytest 2>&1
  rc=$?
  { echo "tool: $tool   pytest rc: $rc   selection: $SEL"; tail -3 $out.pytest; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard|Invalid|out of bounds|misaligned" $out.log | sort | uniq -c | head -20; } > $out.txt
  cat $out.txt
done
