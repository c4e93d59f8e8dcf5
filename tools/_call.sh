mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/r02_pytest_gpu_u.log 2>&1; echo "pytest rc=$?"; tail -n 5 gpurun_out/r02_pytest_gpu_u.log
IBLB_TAG=r02u bash tools/gpu_sanitize.sh | tail -7
for k in 20 200; do
python bench.py --steps $k --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/r02u_bench_c2_$k.json 2> gpurun_out/r02u_bench_c2_$k.err; echo "bench rc=$?"
python - <<PY
import json
for line in open("gpurun_out/r02u_bench_c2_$k.json"):
    if line.startswith("{"):
        d=json.loads(line); print("steps $k:", d["value"], d["ms_per_step"], "e2e", d["e2e"]["value"], d["parity"]["ok"])
PY
done
