# GPU-box batch (N GPUs): C5 ensemble only (N replicas of C1 under torchrun)
N=${1:-8}
set -x
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 \
    bench.py --gpus $N --workload c1 --steps 2000 --warmup 100 > gpurun_out/bench_c5_n$N.json 2> gpurun_out/bench_c5_n$N.err; echo "c5 rc=$?"
tail -n 1 gpurun_out/bench_c5_n$N.json | cut -c1-160
