mkdir -p gpurun_out
for n in 8 2; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2968$n bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/r02v_scale_n$n.json 2> gpurun_out/r02v_scale_n$n.err; echo "bench$n rc=$?"
python - <<PY
import json
for line in open("gpurun_out/r02v_scale_n$n.json"):
    if line.startswith("{"):
        d=json.loads(line); print("N=$n", d["value"], d["ms_per_step"], "e2e", d["e2e"]["value"], "parity", d["parity"]["ok"], "c4s", d["parity"]["c4s"].get("bitwise_equal"), "c4", d["c4"].get("value"))
PY
done
