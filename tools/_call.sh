mkdir -p gpurun_out
python tools/ab_lib.py cuda_iblb_11_b200/libiblb_r2start.so cuda_iblb_11_b200/libiblb.so > gpurun_out/r02_ab_lib4.log 2>&1; cat gpurun_out/r02_ab_lib4.log
timeout 900 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/r02_pytest_gpu_i.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/r02_pytest_gpu_i.log
IBLB_TAG=r02i bash tools/gpu_sanitize.sh | tail -7
timeout 600 python tools/time_small.py 6 > gpurun_out/r02_time_small_i.log 2>&1; grep -E "forms|cilia:" gpurun_out/r02_time_small_i.log
