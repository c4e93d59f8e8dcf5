mkdir -p gpurun_out
python tools/ab_lib.py cuda_iblb_11_b200/libiblb_before.so cuda_iblb_11_b200/libiblb.so > gpurun_out/r02_ab_y.log 2>&1; cat gpurun_out/r02_ab_y.log
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q --timeout 600 -x -k "fused or block_orders or in_phase or golden or tall or geometry or carpet" > gpurun_out/r02_pytest_y.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/r02_pytest_y.log
