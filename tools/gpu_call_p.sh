# GPU-box batch P: HEAD verification — GPU suite, smoke, default bench (C2) + reference arm, C3/C1
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --timeout 900 -x > gpurun_out/pytest_gpu_p.log 2>&1; echo "pytest rc=$?"
tail -n 4 gpurun_out/pytest_gpu_p.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke_p.log 2>&1; echo "smoke rc=$?"; tail -n 3 gpurun_out/smoke_p.log
python bench.py --impl reference > gpurun_out/bench_ref_c2_p.json 2> gpurun_out/bench_ref_c2_p.err; echo "ref rc=$?"
python bench.py > gpurun_out/bench_c2_p.json 2> gpurun_out/bench_c2_p.err; echo "bench rc=$?"
python bench.py --workload c3 --steps 4000 --warmup 100 > gpurun_out/bench_c3_p.json 2> gpurun_out/bench_c3_p.err
python bench.py --workload c3 --impl reference --steps 200 --warmup 10 > gpurun_out/bench_ref_c3_p.json 2> gpurun_out/bench_ref_c3_p.err
python bench.py --workload c1 --steps 4000 --warmup 100 --no-cpu-baseline > gpurun_out/bench_c1_p.json 2> gpurun_out/bench_c1_p.err
cut -c1-900 gpurun_out/bench_ref_c2_p.json gpurun_out/bench_c2_p.json gpurun_out/bench_c3_p.json gpurun_out/bench_ref_c3_p.json gpurun_out/bench_c1_p.json
