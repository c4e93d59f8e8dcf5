mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29655 tools/slab_ipc_check.py > gpurun_out/r02_slab_ipc_2b.log 2>&1; echo "ipc rc=$?"; grep -E "case|IPC" gpurun_out/r02_slab_ipc_2b.log
timeout 900 python -m pytest tests/test_gpu_slabs_fullsize.py -m gpu -q --timeout 600 -k "slabs or gpus" > gpurun_out/r02_pytest_2gpu_b.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/r02_pytest_2gpu_b.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29656 bench.py --gpus 2 --steps 100 --warmup 10 > gpurun_out/r02_bench_c2_n2b.json 2> gpurun_out/r02_bench_c2_n2b.err; echo "bench2 rc=$?"
python - <<PY
import json
for line in open("gpurun_out/r02_bench_c2_n2b.json"):
    if line.startswith("{"):
        d=json.loads(line); print(d["value"], d["ms_per_step"], d["parity"]["ok"], d["parity"]["c4s"].get("bitwise_equal"), d["c4"])
PY
