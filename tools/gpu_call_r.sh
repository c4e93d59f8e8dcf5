# GPU-box batch R: ncu evidence for the default bench command at the final HEAD
set -x
mkdir -p gpurun_out
python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c2_r.json 2> gpurun_out/bench_c2_r.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c2_r.csv \
    python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_r1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:lb_fused_v2 -s 4 -c 1 -f -o gpurun_out/prof_lb_fused_v2_r \
    python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_r2.log 2>&1
tail -2 gpurun_out/ncu_r2.log; ls -la gpurun_out/*.ncu-rep
