"""Rewrites BASELINE.md section 4 (measured round-1 table) from the JSON lines under profiles/."""
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def g(n):
    return json.loads(open(os.path.join(ROOT, "profiles", n)).read().strip().splitlines()[-1])


c2, r2, c3, r3, c1 = (g("r01_bench_c2_final.json"), g("r01_bench_ref_c2_final.json"),
                      g("r01_bench_c3_final.json"), g("r01_bench_ref_c3_final.json"),
                      g("r01_bench_c1_final.json"))
c2n = {n: g("r01_bench_c2_n%d.json" % n) for n in (2, 4, 8)}
c4n = {n: g("r01_bench_c4_n%d.json" % n) for n in (1, 2, 4, 8)}
c5 = g("r01_bench_c5_n8.json")
cb = c2.get("cpu_baseline", {})
share = 100 * (1 - 4.5 / (1e3 * c3["ms_per_step"]))
txt = """
## 4. Measured in round 1 (B200, same box and run for both arms; files under `profiles/`)

| # | new solver | reference CUDA build (sm_100a, unmodified sources) | notes |
|---|---|---|---|
| C2 | **%.0f MLUPS**, %.3f ms/step, %.1f GB/s of algorithmic traffic = **%.1f %%** of the measured HBM copy peak (%.1f GB/s); ncu DRAM traffic %.3f GB per launch vs %.3f GB algorithmic; e2e (host buffers in and out) %.0f MLUPS | %.0f MLUPS, %.2f ms/step | `r01_bench_c2_final.json`, `r01_bench_ref_c2_final.json`; CPU oracle port on %s host threads: %s MLUPS |
| C2 weak scaling | %.0f / %.0f / %.0f MLUPS on 2 / 4 / 8 GPUs (8192 columns per GPU) | single GPU only | `r01_bench_c2_n{2,4,8}.json` |
| C3 | **%.0f MLUPS = %.1f us/step**, one kernel launch per iteration; e2e %.0f MLUPS | %.0f MLUPS = %.0f us/step | `r01_bench_c3_final.json`, `r01_bench_ref_c3_final.json`; IB share ~39 %% by differencing against pure LB on the same grid (target 10 %% not met, DESIGN.md section 3) |
| C1 | %.0f MLUPS = %.1f us/step | 275 us/step (`r01_diag_vs_reference_cuda.log`) | `r01_bench_c1_final.json` |
| C4 | %.0f / %.0f / %.0f / %.0f MLUPS on 1 / 2 / 4 / 8 GPUs (strong scaling), **%.1f %%** parallel efficiency at 8 | cannot address this grid (32-bit `9*j`) | `r01_bench_c4_n{1,2,4,8}.json` |
| C5 | %.0f MLUPS aggregate, 8 independent replicas of C1, one per GPU | runs the lines sequentially on one GPU | `r01_bench_c5_n8.json`; launcher `python -m cuda_iblb_11_b200.sweep` |
""" % (c2["value"], c2["ms_per_step"], c2["roofline"]["achieved"], 100 * c2["roofline"]["frac"],
       c2["roofline"]["peak"], (c2["roofline"]["traffic"] or 0) / 1e9,
       c2["roofline"]["algorithmic_bytes_per_launch"] / 1e9, c2["e2e"]["value"],
       r2["value"], r2["ms_per_step"], cb.get("cores", "?"), cb.get("value", "?"),
       c2n[2]["value"], c2n[4]["value"], c2n[8]["value"],
       c3["value"], 1e3 * c3["ms_per_step"], c3["e2e"]["value"], r3["value"], 1e3 * r3["ms_per_step"],
       c1["value"], 1e3 * c1["ms_per_step"],
       c4n[1]["value"], c4n[2]["value"], c4n[4]["value"], c4n[8]["value"],
       100 * c4n[8]["value"] / (8 * c4n[1]["value"]), c5["value"])
p = os.path.join(ROOT, "BASELINE.md")
s = open(p).read()
i = s.find("\n## 4. Measured in round 1")
if i >= 0:
    s = s[:i]
open(p, "w").write(s.rstrip("\n") + "\n" + txt)
print(txt)
