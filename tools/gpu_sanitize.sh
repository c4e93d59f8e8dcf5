#!/bin/bash
# Range-checked build over every launch mode (stands in for compute-sanitizer, closed on this pool).
mkdir -p gpurun_out
IBLB_LIB=$PWD/cuda_iblb_11_b200/libiblb_checked.so timeout 900 python tools/sanitize_cases.py \
  c1_fused c1_nofuse split1024 slabs_local pure_lb getters > gpurun_out/${IBLB_TAG:-r02}_checked_build.log 2>&1
echo "checked build rc=$?"; cat gpurun_out/${IBLB_TAG:-r02}_checked_build.log
