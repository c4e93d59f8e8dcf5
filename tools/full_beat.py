"""Development tool (GPU box): the product against the oracle over a WHOLE beat of the in-phase
script lines (cilia6.sh:17 `6 6 ...`, and the same with 12 cilia): T = 100 000 iterations
(main.cu:299-301), compared every 1000.  Records, per check point, the
achieved differences (not just pass/fail) and stops at the first point where the oracle's state is
no longer finite.  Writes <out>.json and <out>.md; the oracle is the checker, nothing here is
imported by the product.

    python tools/full_beat.py --out gpurun_out/r03_full_beat [--budget-s 900] [--every 1000]
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np

from checkers import REF_FLAG_PICK_LOWEST_K, Checker, make_params, rel_l2
from cuda_iblb_11_b200 import Params, Solver


def one_case(c_num, every, total, budget_s):
    X = 48 * c_num
    s = Solver(Params.channel(X, 192, c_num=c_num, c_space=48, c_fraction=c_num))
    o = Checker("oracle", make_params(X, 192, c_num=c_num, c_space=48, c_fraction=c_num,
                                      flags=REF_FLAG_PICK_LOWEST_K, threads=os.cpu_count() or 1))
    rows, t0, stop = [], time.time(), "completed the beat"
    for it in range(every, total + 1, every):
        s.step(every)
        o.step(every)
        rho, u = s.fields()
        r2, u2 = o.fields()
        m1, m2 = s.markers(), o.markers()
        fin_o = bool(np.isfinite(r2).all() and np.isfinite(u2).all() and np.isfinite(o.Q))
        fin_s = bool(np.isfinite(rho).all() and np.isfinite(u).all() and np.isfinite(s.Q))
        row = {"it": it, "oracle_finite": fin_o, "product_finite": fin_s,
               "markers_bit_equal": bool(np.array_equal(m1[0], m2[0]) and np.array_equal(m1[1], m2[1])
                                         and np.array_equal(m1[3], m2[3]))}
        if fin_o and fin_s:
            row.update(rel_l2_rho=rel_l2(rho, r2), rel_l2_u=rel_l2(u, u2),
                       rel_l2_force=rel_l2(s.force(), o.force()), rel_l2_F_s=rel_l2(m1[2], m2[2]),
                       rel_Q=abs(s.Q - o.Q) / max(abs(o.Q), 1e-300), Q=o.Q,
                       max_abs_u=float(np.abs(u2).max()), rho_min=float(r2.min()), rho_max=float(r2.max()))
        rows.append(row)
        if not fin_o or not fin_s:
            stop = "first non-finite state at the check of iteration %d (oracle finite: %s, product finite: %s)" % (
                it, fin_o, fin_s)
            break
        if time.time() - t0 > budget_s:
            stop = "time budget of %d s reached after iteration %d" % (budget_s, it)
            break
    s.close()
    o.close()
    ok = [r for r in rows if "rel_l2_u" in r]
    summary = {"cilia": c_num, "grid": [X, 192], "checked_every": every, "stop": stop,
               "last_compared_iteration": ok[-1]["it"] if ok else 0,
               "wall_s": round(time.time() - t0, 1)}
    for k in ("rel_l2_rho", "rel_l2_u", "rel_l2_force", "rel_l2_F_s", "rel_Q"):
        summary["max_" + k] = max([r[k] for r in ok] or [None])
    summary["markers_bit_equal_everywhere"] = all(r["markers_bit_equal"] for r in rows)
    return summary, rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="gpurun_out/r03_full_beat")
    ap.add_argument("--every", type=int, default=1000)
    ap.add_argument("--total", type=int, default=100000)
    ap.add_argument("--budget-s", type=int, default=900)
    ap.add_argument("--cilia", default="6,12")
    a = ap.parse_args()
    res = {}
    md = ["# Product vs oracle over a whole beat (T = %d iterations), in-phase script lines" % a.total, ""]
    for c in [int(x) for x in a.cilia.split(",")]:
        summary, rows = one_case(c, a.every, a.total, a.budget_s)
        res["c_num=%d" % c] = {"summary": summary, "rows": rows}
        md += ["## %d cilia, %dx192, `c_fraction = c_num`" % (c, 48 * c), "",
               "* " + summary["stop"] + "; %s s wall" % summary["wall_s"],
               "* maxima over the compared points (up to iteration %d): rho %.3g, u %.3g, force %.3g, F_s %.3g, Q %.3g"
               % (summary["last_compared_iteration"], *(summary["max_" + k] or float("nan") for k in
                                                        ("rel_l2_rho", "rel_l2_u", "rel_l2_force", "rel_l2_F_s", "rel_Q"))),
               "* marker positions / velocities / epsilon bit-equal at every point: %s" % summary["markers_bit_equal_everywhere"],
               "", "| it | rel L2 rho | rel L2 u | rel L2 force | rel L2 F_s | rel Q | max abs u | rho min | rho max |",
               "|---|---|---|---|---|---|---|---|---|"]
        shown = [r for i, r in enumerate(rows) if "rel_l2_u" in r and (i < 5 or i % 10 == 9 or i >= len(rows) - 5)]
        for r in shown:
            md.append("| %d | %.2e | %.2e | %.2e | %.2e | %.2e | %.3g | %.6f | %.6f |" % (
                r["it"], r["rel_l2_rho"], r["rel_l2_u"], r["rel_l2_force"], r["rel_l2_F_s"], r["rel_Q"],
                r["max_abs_u"], r["rho_min"], r["rho_max"]))
        md.append("")
        print(json.dumps(summary), flush=True)
    json.dump(res, open(a.out + ".json", "w"), indent=1)
    open(a.out + ".md", "w").write("\n".join(md) + "\n")
    return 0


if __name__ == "__main__":
    sys.exit(main())
