mkdir -p gpurun_out
bash tools/tile_timing.sh > gpurun_out/r02_tile_timing.log 2>&1; cat gpurun_out/r02_tile_timing.log
