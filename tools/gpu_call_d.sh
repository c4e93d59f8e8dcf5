set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --timeout 900 -k "fused or cilia or slabs_on_one or in_phase or golden" > gpurun_out/pytest_gpu_d.log 2>&1
python tools/time_small.py 4 > gpurun_out/time_small_d.log 2>&1
tail -n 12 gpurun_out/pytest_gpu_d.log gpurun_out/time_small_d.log
