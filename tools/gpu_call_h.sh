# GPU-box batch H: single-launch IB+LB iteration — tests, small-grid timing, C3/C1 bench.
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --timeout 900 -x > gpurun_out/pytest_gpu_h.log 2>&1; echo "pytest rc=$?"
python tools/time_small.py 4 > gpurun_out/time_small_h.log 2>&1
python bench.py --workload c3 --steps 4000 --warmup 100 --no-cpu-baseline > gpurun_out/bench_c3_h.json 2> gpurun_out/bench_c3_h.err
python bench.py --workload c1 --steps 4000 --warmup 100 --no-cpu-baseline > gpurun_out/bench_c1_h.json 2> gpurun_out/bench_c1_h.err
tail -n 15 gpurun_out/pytest_gpu_h.log; cat gpurun_out/time_small_h.log; cut -c1-200 gpurun_out/bench_c3_h.json gpurun_out/bench_c1_h.json
