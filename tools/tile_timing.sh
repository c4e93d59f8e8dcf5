#!/bin/bash
# Development (GPU box): phase stamps of lb_ib_tile_step at C3 and C1 from the timing build
#   nvcc -DIBLB_TILE_TIMING ... -o cuda_iblb_11_b200/libiblb_timing.so   
# printed by iblb_destroy for the last launch of a 264-step run.
for args in "576 192 12 48" "288 192 6 48"; do
  echo "== $args"
  IBLB_LIB=$PWD/cuda_iblb_11_b200/libiblb_timing.so python - $args <<'PY'
import sys
sys.path.insert(0, ".")
from cuda_iblb_11_b200 import Params, Solver
X, Y, cn, cs = [int(v) for v in sys.argv[1:5]]
s = Solver(Params.channel(X, Y, c_num=cn, c_space=cs, c_fraction=cn))
s.step(264)
s.sync()
s.close()
PY
done
