# GPU-box batch M: PDL chain of lb_ib_tile_step — full GPU suite, small-grid timing, C3/C1 bench
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --timeout 900 -x > gpurun_out/pytest_gpu_m.log 2>&1; echo "pytest rc=$?"
tail -n 8 gpurun_out/pytest_gpu_m.log
python tools/time_small.py 4 2>&1 | grep -v event-timed | tee gpurun_out/time_small_m.log
python bench.py --workload c3 --steps 4000 --warmup 100 --no-cpu-baseline > gpurun_out/bench_c3_m.json 2> gpurun_out/bench_c3_m.err
python bench.py --workload c1 --steps 4000 --warmup 100 --no-cpu-baseline > gpurun_out/bench_c1_m.json 2> gpurun_out/bench_c1_m.err
cut -c1-200 gpurun_out/bench_c3_m.json gpurun_out/bench_c1_m.json
