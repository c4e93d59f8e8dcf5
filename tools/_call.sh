mkdir -p gpurun_out
python tools/ab_lib.py cuda_iblb_11_b200/libiblb_before.so cuda_iblb_11_b200/libiblb.so > gpurun_out/r02_ab_order2.log 2>&1; cat gpurun_out/r02_ab_order2.log
timeout 600 python tools/time_small.py 6 > gpurun_out/r02_time_small_l.log 2>&1; grep -E "cilia:" gpurun_out/r02_time_small_l.log
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q --timeout 600 -x -k "fused or block_orders or in_phase or tall or geometry" > gpurun_out/r02_pytest_order2.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/r02_pytest_order2.log
