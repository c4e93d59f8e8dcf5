# ncu --set full of the single-launch IB+LB kernel on the warm C3 case
set -x
mkdir -p gpurun_out
ncu --set full --cache-control none --clock-control none --import-source on -k regex:lb_ib_tile_step -s 40 -c 2 -f -o gpurun_out/prof_tile_c3 \
    python bench.py --workload c3 --steps 64 --warmup 64 --no-cpu-baseline > gpurun_out/ncu_i.log 2>&1
tail -3 gpurun_out/ncu_i.log
ls -la gpurun_out/*.ncu-rep
