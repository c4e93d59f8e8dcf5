# per-kernel durations of the warm C3 step (ncu, no cache/clock control)
set -x
mkdir -p gpurun_out
ncu --metrics gpu__time_duration.sum --cache-control none --clock-control none -s 200 -c 300 --csv --log-file gpurun_out/launches_c3_warm.csv \
    python bench.py --workload c3 --steps 64 --warmup 64 --no-cpu-baseline > gpurun_out/ncu_e.log 2>&1
python - <<'PY'
import csv, collections
rows = list(csv.reader(l for l in open('gpurun_out/launches_c3_warm.csv') if l.startswith('"')))
h = rows[0]; ki, vi = h.index("Kernel Name"), h.index("Metric Value")
d = collections.defaultdict(list)
for r in rows[1:]:
    d[r[ki].split('(')[0]].append(float(r[vi].replace(',', '')))
for k, v in d.items():
    print("%-24s n=%3d  mean %.2f us  min %.2f  max %.2f" % (k, len(v), sum(v) / len(v) / 1e3, min(v) / 1e3, max(v) / 1e3))
PY
