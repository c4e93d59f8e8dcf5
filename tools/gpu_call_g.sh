# GPU-box batch G (N GPUs): default bench under torchrun (weak C2) and C4 strong at N; 2-GPU tests.
N=${1:-4}
set -x
mkdir -p gpurun_out
nvidia-smi -L | wc -l
if [ "$N" = "2" ]; then python -m pytest tests -m gpu -q --timeout 900 -k "ipc or two_gpu or 2gpu or multi" > gpurun_out/pytest_gpu_g_n$N.log 2>&1; fi
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_c2_n$N.json 2> gpurun_out/bench_c2_n$N.err; echo "c2 rc=$?"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
    bench.py --gpus $N --workload c4 --steps 10 --warmup 3 > gpurun_out/bench_c4_n$N.json 2> gpurun_out/bench_c4_n$N.err; echo "c4 rc=$?"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 \
    bench.py --gpus $N --workload c1 --steps 2000 --warmup 100 > gpurun_out/bench_c5_n$N.json 2> gpurun_out/bench_c5_n$N.err; echo "c5 rc=$?"
tail -n 3 gpurun_out/bench_c*_n$N.json; tail -n 5 gpurun_out/bench_c*_n$N.err
