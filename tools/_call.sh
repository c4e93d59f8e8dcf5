mkdir -p gpurun_out
python tools/ab_lib.py cuda_iblb_11_b200/libiblb_r2start.so cuda_iblb_11_b200/libiblb.so > gpurun_out/r02_ab_lib3.log 2>&1; cat gpurun_out/r02_ab_lib3.log
timeout 900 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/r02_pytest_gpu_g.log 2>&1; echo "pytest rc=$?"; tail -n 6 gpurun_out/r02_pytest_gpu_g.log
IBLB_TAG=r02g bash tools/gpu_sanitize.sh | tail -7
