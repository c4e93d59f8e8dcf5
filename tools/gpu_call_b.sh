# GPU-box batch B (2 GPUs): full GPU test-suite incl. the two-GPU slab tests, N=2 benches.
set -x
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus_b.log 2>&1
nvidia-smi topo -m >> gpurun_out/gpus_b.log 2>&1
free -g >> gpurun_out/gpus_b.log 2>&1
python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/pytest_gpu_b.log 2>&1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/bench_c2_n2.json 2> gpurun_out/bench_c2_n2.err
timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_c2_n1.json 2> gpurun_out/bench_c2_n1.err
timeout 900 $TR bench.py --gpus 2 --workload c4 --steps 10 --warmup 3 > gpurun_out/bench_c4_n2.json 2> gpurun_out/bench_c4_n2.err
timeout 900 python bench.py --gpus 1 --workload c4 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c4_n1.json 2> gpurun_out/bench_c4_n1.err
tail -n 5 gpurun_out/pytest_gpu_b.log gpurun_out/bench_c*_n*.json gpurun_out/bench_c*_n*.err
