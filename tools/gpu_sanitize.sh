#!/bin/bash
# compute-sanitizer memcheck + racecheck over every launch mode; logs under
# gpurun_out/sanitizer/, summaries copied to profiles/ by hand afterwards.
mkdir -p gpurun_out/sanitizer
CS=/usr/local/cuda/bin/compute-sanitizer
for tool in memcheck racecheck; do
  for c in c1_fused c1_nofuse split1024 slabs_local pure_lb getters; do
    timeout 600 $CS --tool $tool --print-limit 20 --log-file gpurun_out/sanitizer/${tool}_${c}.log \
      python tools/sanitize_cases.py $c > gpurun_out/sanitizer/${tool}_${c}.out 2>&1
    echo "$tool $c exit $? : $(tail -n 2 gpurun_out/sanitizer/${tool}_${c}.log | tr '\n' ' ')"
  done
done
