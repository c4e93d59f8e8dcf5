# GPU-box batch C: fused IB+LB step — tests, timing, C4 on one GPU, ncu of the fused kernel.
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/pytest_gpu_c.log 2>&1
python tools/time_small.py > gpurun_out/time_small.log 2>&1
timeout 900 python bench.py --gpus 1 --workload c4 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c4_n1_fused.json 2> gpurun_out/bench_c4_n1_fused.err
python bench.py --workload c3 --steps 4000 --warmup 100 --no-cpu-baseline > gpurun_out/bench_c3_fused.json 2> gpurun_out/bench_c3_fused.err
python bench.py --workload c1 --steps 4000 --warmup 100 --no-cpu-baseline > gpurun_out/bench_c1_fused.json 2> gpurun_out/bench_c1_fused.err
ncu --set full --clock-control none --import-source on -k regex:"lb_band_step|ib_markers" -s 40 -c 4 -f -o gpurun_out/prof_fused_c3 \
    python bench.py --workload c3 --steps 40 --warmup 10 --no-cpu-baseline > gpurun_out/ncu_fused.log 2>&1
ncu -i gpurun_out/prof_fused_c3.ncu-rep --page raw --csv > gpurun_out/prof_fused_c3_raw.csv 2>> gpurun_out/ncu_fused.log
tail -n 8 gpurun_out/pytest_gpu_c.log gpurun_out/time_small.log gpurun_out/bench_c*_fused.json
