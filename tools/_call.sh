mkdir -p gpurun_out
for n in 8 4; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2967$n bench.py --gpus $n --steps 100 --warmup 10 > gpurun_out/r02_bench_c2_n$n.json 2> gpurun_out/r02_bench_c2_n$n.err; echo "bench$n rc=$?"; cut -c1-400 gpurun_out/r02_bench_c2_n$n.json; python - <<PY
import json
d=json.load(open("gpurun_out/r02_bench_c2_n$n.json"))
print("N=$n value", d["value"], "parity", d["parity"]["ok"], d["parity"]["c4s"].get("how"), d["parity"]["c4s"].get("bitwise_equal"), "c4", d["c4"].get("value"), d["c4"].get("ms_per_step"), "e2e", d["e2e"]["value"])
PY
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29699 tools/slab_ipc_check.py > gpurun_out/r02_slab_ipc_8.log 2>&1; echo "ipc rc=$?"; grep -E "case|IPC" gpurun_out/r02_slab_ipc_8.log
