mkdir -p gpurun_out
nvidia-smi topo -m | head -8
python -m pytest tests/test_gpu_slabs_fullsize.py -m gpu -q --timeout 900 -k "slabs" > gpurun_out/r02_pytest_2gpu.log 2>&1; echo "pytest rc=$?"; tail -n 12 gpurun_out/r02_pytest_2gpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29655 tools/slab_ipc_check.py > gpurun_out/r02_slab_ipc_2.log 2>&1; echo "ipc rc=$?"; grep -E "case|IPC" gpurun_out/r02_slab_ipc_2.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29656 bench.py --gpus 2 --steps 100 --warmup 10 > gpurun_out/r02_bench_c2_n2.json 2> gpurun_out/r02_bench_c2_n2.err; echo "bench2 rc=$?"; cut -c1-3000 gpurun_out/r02_bench_c2_n2.json; tail -n 5 gpurun_out/r02_bench_c2_n2.err
python tools/run_reference_program.py --out gpurun_out/r02_refprog > gpurun_out/r02_refprog.log 2>&1; echo "refprog rc=$?"; tail -n 70 gpurun_out/r02_refprog.log
