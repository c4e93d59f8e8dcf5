# GPU-box batch F: re-verify HEAD — full GPU test suite, smoke, default bench (both arms), ncu launch list.
set -x
mkdir -p gpurun_out
nproc; nvidia-smi -L
python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/pytest_gpu_f.log 2>&1; echo "pytest rc=$?"
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_f.log 2>&1; echo "smoke rc=$?"
python bench.py > gpurun_out/bench_c2_f.json 2> gpurun_out/bench_c2_f.err; echo "bench rc=$?"
python bench.py --impl reference > gpurun_out/bench_ref_c2_f.json 2> gpurun_out/bench_ref_c2_f.err; echo "ref rc=$?"
python bench.py --workload c3 --steps 4000 --warmup 100 --no-cpu-baseline > gpurun_out/bench_c3_f.json 2> gpurun_out/bench_c3_f.err
python bench.py --workload c3 --impl reference --steps 200 --warmup 10 > gpurun_out/bench_ref_c3_f.json 2> gpurun_out/bench_ref_c3_f.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c2_f.csv \
    python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_f.log 2>&1
tools/micro/lat > gpurun_out/lat_f.log 2>&1
tail -n 6 gpurun_out/pytest_gpu_f.log gpurun_out/smoke_f.log gpurun_out/bench_*_f.json gpurun_out/lat_f.log
