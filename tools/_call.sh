mkdir -p gpurun_out
python tools/ab_lib.py cuda_iblb_11_b200/libiblb_noprep.so cuda_iblb_11_b200/libiblb.so > gpurun_out/r02_ab_prep.log 2>&1; cat gpurun_out/r02_ab_prep.log
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q --timeout 600 -x -k "fused or every_form or in_phase or golden or geometry" > gpurun_out/r02_pytest_prep.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/r02_pytest_prep.log
