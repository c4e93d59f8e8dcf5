#!/bin/bash
# One parameterised GPU-box batch (replaces round 1's gpu_call_*.sh).  Usage, from the repo root:
#   gpurun --timeout 1500 -- 'bash tools/gpu_call.sh verify refprog sanitize small'
# stages: verify   GPU suite + smoke + default bench (+ reference arm)
#         refprog  the unmodified reference program vs ./app (tools/run_reference_program.py)
#         sanitize the range-checked build (-DIBLB_CHECKED: bounds checks on every device index, no race
#                  detection) over every launch mode (tools/gpu_sanitize.sh; compute-sanitizer is closed here)
#         small    per-step times of the cilia cases and the IB share (tools/time_small.py)
#         ncu_c2 / ncu_c3   launch list + --set full capture of the dominant kernel
#         ncu_markers       --set full capture of the marker-side and host-buffer kernels
#         driver_bench      both bench arms with the driver's arguments (--gpus 1 --steps 20 --warmup 5)
TAG=${IBLB_TAG:-r03}
mkdir -p gpurun_out
for stage in "$@"; do
  case $stage in
    verify)
      python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"
      tail -n 15 gpurun_out/${TAG}_pytest_gpu.log
      python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 3 gpurun_out/${TAG}_smoke.log
      python bench.py --impl reference > gpurun_out/${TAG}_bench_ref_c2.json 2> gpurun_out/${TAG}_bench_ref_c2.err; echo "ref rc=$?"
      python bench.py > gpurun_out/${TAG}_bench_c2.json 2> gpurun_out/${TAG}_bench_c2.err; echo "bench rc=$?"
      cut -c1-1500 gpurun_out/${TAG}_bench_ref_c2.json gpurun_out/${TAG}_bench_c2.json; tail -n 5 gpurun_out/${TAG}_bench_c2.err
      ;;
    refprog)
      python tools/run_reference_program.py --out gpurun_out/${TAG}_refprog > gpurun_out/${TAG}_refprog.log 2>&1; echo "refprog rc=$?"
      tail -n 60 gpurun_out/${TAG}_refprog.log
      ;;
    sanitize)
      bash tools/gpu_sanitize.sh
      ;;
    small)
      python tools/time_small.py ${IBLB_SMALL_CASES:-4} > gpurun_out/${TAG}_time_small.log 2>&1; echo "small rc=$?"
      cat gpurun_out/${TAG}_time_small.log
      ;;
    ncu_c2)
      ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/${TAG}_c2_ncu_launches.csv \
        python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_c2_ncu_launches.out 2>&1
      ncu --set full --clock-control none --import-source on -k regex:lb_fused -s 6 -c 1 -o gpurun_out/${TAG}_lb_fused_full -f \
        python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_c2_ncu_full.out 2>&1
      ;;
    ncu_c3)
      ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -s 200 -c 200 --csv --log-file gpurun_out/${TAG}_c3_ncu_launches.csv \
        python bench.py --workload c3 --steps 64 --warmup 64 --no-cpu-baseline > gpurun_out/${TAG}_c3_ncu_launches.out 2>&1
      ncu --set full --clock-control none --cache-control none --import-source on -k regex:lb_ib -s 40 -c 2 -o gpurun_out/${TAG}_lb_ib_full -f \
        python bench.py --workload c3 --steps 64 --warmup 64 --no-cpu-baseline > gpurun_out/${TAG}_c3_ncu_full.out 2>&1
      ;;
    ncu_markers)
      python tools/ncu_marker_kernels.py > gpurun_out/${TAG}_ncu_markers_plain.out 2>&1; echo "plain rc=$?"
      ncu --set full --clock-control none --import-source on \
        -k regex:"ib_kinematics_batch|ib_tile_lists|ib_interp|ib_spread|lb_init_t|lb_moments_t" -c 24 \
        -o gpurun_out/${TAG}_marker_kernels_full -f python tools/ncu_marker_kernels.py > gpurun_out/${TAG}_ncu_markers.out 2>&1; echo "ncu rc=$?"
      tail -n 3 gpurun_out/${TAG}_ncu_markers.out
      ;;
    driver_bench)
      python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/${TAG}_bench_ref_driver.json 2> gpurun_out/${TAG}_bench_ref_driver.err; echo "ref rc=$?"
      python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/${TAG}_bench_driver.json 2> gpurun_out/${TAG}_bench_driver.err; echo "bench rc=$?"
      cut -c1-6000 gpurun_out/${TAG}_bench_driver.json; tail -n 5 gpurun_out/${TAG}_bench_driver.err
      ;;
    *) echo "unknown stage $stage";;
  esac
done
