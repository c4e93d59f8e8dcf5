mkdir -p gpurun_out
timeout 300 python tools/time_small.py 4 > gpurun_out/r02_time_small_roles.log 2>&1; echo "small rc=$?"; cat gpurun_out/r02_time_small_roles.log
timeout 900 python -m pytest tests -m gpu -q --timeout 600 -x > gpurun_out/r02_pytest_gpu_roles.log 2>&1; echo "pytest rc=$?"; tail -n 25 gpurun_out/r02_pytest_gpu_roles.log
IBLB_TAG=r02 bash tools/gpu_sanitize.sh
