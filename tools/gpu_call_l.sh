# GPU-box batch L: phase stamps (development build), parity of fused vs unfused
set -x
mkdir -p gpurun_out
bash tools/gpu_call_j.sh 2>&1 | grep -E "tile timing|worst"
python -m pytest tests -m gpu -q --timeout 900 -x -k "fused or graph or determin" 2>&1 | tail -3
python tools/time_small.py 4 2>&1 | grep -v event-timed
