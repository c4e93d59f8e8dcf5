# GPU-box batch: tile width 6 against 8 on one box (IBLB_TILE_W override), then the GPU suite
set -x
for w in 8 6 auto; do
  echo "== tile width $w"
  if [ $w = auto ]; then unset IBLB_TILE_W; else export IBLB_TILE_W=$w; fi
  python tools/time_small.py 4 2>&1 | grep "cilia:"
done
unset IBLB_TILE_W
python -m pytest tests -m gpu -q --timeout 900 -x 2>&1 | tail -4
IBLB_TILE_W=8 python -m pytest tests -m gpu -q --timeout 900 -x -k "fused or graph or checkpoint or golden" 2>&1 | tail -3
