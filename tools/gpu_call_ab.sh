# GPU-box A/B on ONE box: libiblb.so (new) against libiblb_old.so (previous build), alternating
set -x
cd cuda_iblb_11_b200; cp libiblb.so libiblb_new.so; cd ..
for rep in 1 2; do
  for v in new old; do
    cp cuda_iblb_11_b200/libiblb_$v.so cuda_iblb_11_b200/libiblb.so
    echo "== $v"
    python tools/diag_geometry.py 2>&1 | grep " x "
    python tools/time_small.py 2 2>&1 | grep "cilia:"
  done
done
cp cuda_iblb_11_b200/libiblb_new.so cuda_iblb_11_b200/libiblb.so
python -m pytest tests -m gpu -q --timeout 900 -x 2>&1 | tail -4
