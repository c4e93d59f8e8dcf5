# GPU-box batch O: ncu of the fused IB+LB kernel on the warm C3 case (durations, then --set full)
set -x
mkdir -p gpurun_out
python bench.py --workload c3 --steps 64 --warmup 64 --no-cpu-baseline > gpurun_out/bench_c3_o.json 2> gpurun_out/bench_c3_o.err || exit 1
ncu --metrics gpu__time_duration.sum --cache-control none --clock-control none -s 200 -c 300 --csv --log-file gpurun_out/launches_c3_warm_o.csv \
    python bench.py --workload c3 --steps 64 --warmup 64 --no-cpu-baseline > gpurun_out/ncu_o1.log 2>&1
ncu --set full --cache-control none --clock-control none --import-source on -k regex:lb_ib_tile_step -s 40 -c 2 -f -o gpurun_out/prof_tile_c3_o \
    python bench.py --workload c3 --steps 64 --warmup 64 --no-cpu-baseline > gpurun_out/ncu_o2.log 2>&1
tail -3 gpurun_out/ncu_o2.log
ls -la gpurun_out/*.ncu-rep
