mkdir -p gpurun_out
timeout 600 python tools/time_small.py 6 > gpurun_out/r02_time_small_j.log 2>&1; grep -E "forms|cilia:" gpurun_out/r02_time_small_j.log
