mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q --timeout 600 -k "four_node or tall or rest_state" > gpurun_out/r02_pytest_v4.log 2>&1; echo "pytest rc=$?"; tail -n 8 gpurun_out/r02_pytest_v4.log
timeout 600 python tools/ab_c2.py > gpurun_out/r02_ab_v2_v4.log 2>&1; cat gpurun_out/r02_ab_v2_v4.log
