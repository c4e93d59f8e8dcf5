mkdir -p gpurun_out
for g in "8192 192 64 128" "2304 192 48 48"; do
  tag=$(echo $g | tr ' ' '_')
  IBLB_FORM=2 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -s 100 -c 50 --csv --log-file gpurun_out/r02_form2_${tag}.csv python tools/ncu_chain.py $g 4 > gpurun_out/r02_form2_${tag}.out 2>&1
  python - <<PY
import csv,collections
rows=[r for r in csv.reader(open("gpurun_out/r02_form2_${tag}.csv")) if len(r)>10]
hdr=rows[0]; ki=hdr.index("Kernel Name"); vi=hdr.index("Metric Value"); gi=hdr.index("Grid Size") if "Grid Size" in hdr else None
d=collections.defaultdict(list)
for r in rows[1:]:
    d[r[ki].split("(")[0] + (" grid "+r[gi] if gi is not None else "")].append(float(r[vi].replace(",","")))
print("${g}", {k:(len(v), round(sum(v)/len(v),1)) for k,v in d.items()})
PY
done
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q --timeout 600 -x -k "flux_history or four_node" > gpurun_out/r02_pytest_hist.log 2>&1; echo "pytest rc=$?"; tail -n 12 gpurun_out/r02_pytest_hist.log
python bench.py --workload c1 --steps 4096 --warmup 64 --no-cpu-baseline --no-parity > gpurun_out/r02_bench_c1_hist.json 2> gpurun_out/r02_bench_c1_hist.err; cut -c1-700 gpurun_out/r02_bench_c1_hist.json
