mkdir -p gpurun_out
tools/micro/copy_peak > gpurun_out/r02_copy_peak2.log 2>&1; cat gpurun_out/r02_copy_peak2.log
