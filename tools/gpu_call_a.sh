# GPU-box batch A (round 1): diagnostics vs the reference CUDA build, golden regeneration, the
# GPU test-suite, smoke, bench (ours + reference arm), ncu launch list + one full capture.
set -x
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus.log 2>&1
python tools/diag_ref_cuda.py > gpurun_out/diag.log 2>&1
python tests/golden/make_golden.py --kind ref_cuda --out gpurun_out/golden > gpurun_out/golden.log 2>&1
python -m pytest tests -m gpu -q -x --timeout 900 > gpurun_out/pytest_gpu.log 2>&1
python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/pytest_gpu_all.log 2>&1
python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1
python bench.py > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err
python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
python bench.py --workload c1 --steps 2000 --warmup 100 --no-cpu-baseline > gpurun_out/bench_c1.json 2> gpurun_out/bench_c1.err
python bench.py --workload c3 --steps 2000 --warmup 100 --no-cpu-baseline > gpurun_out/bench_c3.json 2> gpurun_out/bench_c3.err
python tools/gpu_check.py > gpurun_out/check.log 2>&1
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
      python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:lb_fused_v2 -s 3 -c 2 -f -o gpurun_out/prof_lb \
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu2.log 2>&1
ncu -i gpurun_out/prof_lb.ncu-rep --page raw --csv > gpurun_out/prof_lb_raw.csv 2> gpurun_out/ncu3.log
tail -5 gpurun_out/*.log gpurun_out/*.json gpurun_out/*.err
