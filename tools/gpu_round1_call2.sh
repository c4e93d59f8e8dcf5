set -x
python tests/golden/make_golden.py --kind ref_cuda --out gpurun_out/golden > gpurun_out/golden.log 2>&1
python tools/gpu_check.py > gpurun_out/check2.log 2>&1
python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1
python bench.py > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err
python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:lb_fused_v2 -s 3 -c 3 -o gpurun_out/prof_lb python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu2.log 2>&1
tail -3 gpurun_out/*.log gpurun_out/*.json gpurun_out/*.err
