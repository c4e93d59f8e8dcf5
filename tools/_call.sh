mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/r02_pytest_gpu_k.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/r02_pytest_gpu_k.log
IBLB_TAG=r02k bash tools/gpu_sanitize.sh | tail -7
python tools/ab_lib.py cuda_iblb_11_b200/libiblb_r2start.so cuda_iblb_11_b200/libiblb.so > gpurun_out/r02_ab_lib5.log 2>&1; cat gpurun_out/r02_ab_lib5.log
timeout 600 python tools/time_small.py 6 > gpurun_out/r02_time_small_k.log 2>&1; grep -E "order|cilia:" gpurun_out/r02_time_small_k.log
python bench.py --no-cpu-baseline > gpurun_out/r02k_bench_c2.json 2> gpurun_out/r02k_bench_c2.err; echo "bench rc=$?"
python - <<PY
import json
for line in open("gpurun_out/r02k_bench_c2.json"):
    if line.startswith("{"):
        d=json.loads(line); print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["parity"]["ok"], d["c4"].get("value"), d["ib_share"])
PY
