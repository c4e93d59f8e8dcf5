#!/usr/bin/env python
"""TEST INFRASTRUCTURE (GPU box): run the UNMODIFIED reference program and `./app` side by side.

oracle/_ref/IBLB is ptheywood/CUDA_IBLB_11 built from its own four source files with main() intact
(oracle/Makefile, target ref_bin).  This script gives it the directory tree its hard-coded output
root needs on Linux (`./C:/Users/phq16ja/Documents/Data/{Raw,Cilia}/<c_num>/<c_fraction>`, `Flux`;
main.cu:591-613), runs the same argument line through it and through the product's
`app`, and compares every file the two write:

  SimLog.txt       text-identical except the date, completion-time and runtime lines
  *-cilia.dat      text-identical except marker 32 of each cilium (the reference's write race, Q12)
  *-fluid.dat      numeric (6 significant digits in the file): relative L2 per column
  *-flux.dat       numeric; first non-finite entry reported

and times both programs on longer runs without snapshots (whole-binary ms per step).
Writes <out>/summary.json and <out>/summary.md; nothing here is imported by the product.
"""
import argparse
import glob
import json
import math
import os
import shutil
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
IBLB = os.path.join(ROOT, "oracle", "_ref", "IBLB")
APP = os.path.join(ROOT, "app")
WIN_ROOT = "C:/Users/phq16ja/Documents/Data"


def run_reference(args, workdir):
    c_fraction, c_num = args[0], args[1]
    base = os.path.join(workdir, WIN_ROOT)
    for d in ("Raw/%s/%s" % (c_num, c_fraction), "Cilia/%s/%s" % (c_num, c_fraction), "Flux"):
        os.makedirs(os.path.join(base, d), exist_ok=True)
    t0 = time.time()
    r = subprocess.run([IBLB] + args, cwd=workdir, capture_output=True, text=True, timeout=3000)
    return base, time.time() - t0, r


def run_ours(args, workdir, env_extra=None):
    base = os.path.join(workdir, "Data")
    env = dict(os.environ, IBLB_ROOT=base + "/")
    env.update(env_extra or {})
    t0 = time.time()
    r = subprocess.run([APP] + args, cwd=workdir, capture_output=True, text=True, timeout=3000,
                       env=env)
    return base, time.time() - t0, r


def rel_l2(a, b):
    num = math.sqrt(sum((x - y) ** 2 for x, y in zip(a, b)))
    den = math.sqrt(sum(y * y for y in b))
    return num / den if den > 0 else num


def read_table(path):
    rows = []
    with open(path) as f:
        for line in f:
            if line.strip():
                rows.append([float(x) for x in line.split()])
    return rows


def compare_simlog(a, b):
    skip = ("Completion time", "Total runtime", "MLUPS", "Non-finite flux", "GPUs:")
    la = [l.rstrip("\n") for l in open(a)]
    lb = [l.rstrip("\n") for l in open(b)]
    la = [l for l in la[1:] if l.strip() and not l.startswith(skip)]
    lb = [l for l in lb[1:] if l.strip() and not l.startswith(skip)]
    # a completion-time block is followed by the asctime line itself
    strip_dates = lambda ls: [l for l in ls if not (len(l.split()) == 5 and l.split()[-1].isdigit()
                                                    and ":" in l)]
    la, lb = strip_dates(la), strip_dates(lb)
    return {"identical": la == lb, "lines": len(la), "reference": la, "ours": lb}


def compare_cilia(a, b, length=96):
    la = [l for l in open(a).read().split("\n")]
    lb = [l for l in open(b).read().split("\n")]
    # marker index = running count of non-blank lines
    def rows(ls):
        out, k = [], 0
        for i, l in enumerate(ls):
            if l.strip():
                out.append((k, l, i))
                k += 1
        return out
    ra, rb = rows(la), rows(lb)
    same_count = len(ra) == len(rb)
    diff_other, diff_32, blank_layout_same = 0, 0, True
    for (k, l1, i1), (_, l2, i2) in zip(ra, rb):
        if l1 != l2:
            if k % length == 32:
                diff_32 += 1
            else:
                diff_other += 1
        elif i1 != i2 and k % length != 32:
            pass
    # blank-line layout (main.cu:991) compared away from marker 32
    return {"markers": len(ra), "same_count": same_count, "rows_differing_marker32": diff_32,
            "rows_differing_elsewhere": diff_other}


def compare_fluid(a, b):
    ta, tb = read_table(a), read_table(b)
    if len(ta) != len(tb):
        return {"rows": [len(ta), len(tb)], "error": "row count differs"}
    text_a = [l for l in open(a) if l.strip()]
    text_b = [l for l in open(b) if l.strip()]
    same_text = sum(1 for x, y in zip(text_a, text_b) if x == y)
    finite = all(math.isfinite(v) for r in ta for v in r)
    out = {"rows": len(ta), "text_identical_rows": same_text, "reference_finite": finite}
    if finite and all(math.isfinite(v) for r in tb for v in r):
        names = ["x", "y", "ux", "uy", "abs_u", "rho"]
        for c, n in enumerate(names):
            out["rel_l2_" + n] = rel_l2([r[c] for r in tb], [r[c] for r in ta])
    else:
        out["ours_finite"] = all(math.isfinite(v) for r in tb for v in r)
    return out


def compare_flux(a, b):
    ta = [l.split() for l in open(a) if l.strip()]
    tb = [l.split() for l in open(b) if l.strip()]
    out = {"reference_lines": [" ".join(x) for x in ta], "ours_lines": [" ".join(x) for x in tb],
           "text_identical": ta == tb}
    nonfinite = [i for i, r in enumerate(ta) if not math.isfinite(float(r[1]))]
    out["reference_first_nonfinite_line"] = nonfinite[0] if nonfinite else None
    nf2 = [i for i, r in enumerate(tb) if not math.isfinite(float(r[1]))]
    out["ours_first_nonfinite_line"] = nf2[0] if nf2 else None
    pairs = [(float(x[1]), float(y[1])) for x, y in zip(ta, tb)
             if math.isfinite(float(x[1])) and math.isfinite(float(y[1]))]
    out["max_rel_diff_finite"] = max([abs(p - q) / max(abs(p), 1e-300) for p, q in pairs] or [0.0])
    # per line (None where either side is non-finite), so that a failure says WHICH line moved.
    # The last line is a different animal: the reference writes it after its loop from a host
    # variable whose last asynchronous copy it never waits for (main.cu:1026-1034; every in-loop
    # line waits on Q_done, main.cu:998), so it may show the flux of an earlier iteration.
    out["rel_diff_per_line"] = [
        abs(float(x[1]) - float(y[1])) / max(abs(float(x[1])), 1e-300)
        if math.isfinite(float(x[1])) and math.isfinite(float(y[1])) else None
        for x, y in zip(ta, tb)]
    inloop = [d for d in out["rel_diff_per_line"][:-1] if d is not None]
    out["max_rel_diff_in_loop_lines"] = max(inloop or [0.0])
    out["reference_last_value"] = float(ta[-1][1]) if ta else None
    out["ours_last_value"] = float(tb[-1][1]) if tb else None
    return out


def case_compare(name, args, out):
    wd_ref = os.path.join(out, name, "ref")
    wd_our = os.path.join(out, name, "ours")
    for d in (wd_ref, wd_our):
        shutil.rmtree(d, ignore_errors=True)
        os.makedirs(d)
    ref_root, t_ref, r1 = run_reference(args, wd_ref)
    our_root, t_our, r2 = run_ours(args, wd_our)
    c_fraction, c_num = args[0], args[1]
    res = {"args": " ".join(args), "reference_exit": r1.returncode, "ours_exit": r2.returncode,
           "reference_wall_s": round(t_ref, 3), "ours_wall_s": round(t_our, 3),
           "reference_stderr_tail": r1.stderr[-400:], "ours_stderr_tail": r2.stderr[-400:]}
    sub = "%s/%s" % (c_num, c_fraction)
    res["simlog"] = compare_simlog(os.path.join(ref_root, "Raw", sub, "SimLog.txt"),
                                   os.path.join(our_root, "Raw", sub, "SimLog.txt"))
    fa = sorted(glob.glob(os.path.join(ref_root, "Flux", "*-flux.dat")))
    fb = sorted(glob.glob(os.path.join(our_root, "Flux", "*-flux.dat")))
    res["flux_file_names"] = [os.path.basename(x) for x in fa], [os.path.basename(x) for x in fb]
    if fa and fb:
        res["flux"] = compare_flux(fa[0], fb[0])
        shutil.copy(fa[0], os.path.join(out, name + "_reference-flux.dat"))
        shutil.copy(fb[0], os.path.join(out, name + "_ours-flux.dat"))
    res["fluid"], res["cilia"] = {}, {}
    for f in sorted(glob.glob(os.path.join(ref_root, "Raw", sub, "*-fluid.dat")),
                    key=lambda p: int(os.path.basename(p).split("-")[0])):
        it = os.path.basename(f).split("-")[0]
        g = os.path.join(our_root, "Raw", sub, os.path.basename(f))
        res["fluid"][it] = compare_fluid(f, g) if os.path.exists(g) else {"error": "missing in ours"}
    for f in sorted(glob.glob(os.path.join(ref_root, "Cilia", sub, "*-cilia.dat")),
                    key=lambda p: int(os.path.basename(p).split("-")[0])):
        it = os.path.basename(f).split("-")[0]
        g = os.path.join(our_root, "Cilia", sub, os.path.basename(f))
        res["cilia"][it] = compare_cilia(f, g) if os.path.exists(g) else {"error": "missing in ours"}
    # the bulky snapshot trees are scratch: keep one pair for the record
    for it in ("0", "10"):
        for kind, folder, ext in (("ref", ref_root, "Raw"), ("ours", our_root, "Raw")):
            src = os.path.join(folder, ext, sub, it + "-fluid.dat")
            if os.path.exists(src) and name == "inphase" and it == "10":
                shutil.copy(src, os.path.join(out, "%s_%s_%s-fluid.dat" % (name, kind, it)))
    shutil.copy(os.path.join(ref_root, "Raw", sub, "SimLog.txt"), os.path.join(out, name + "_reference-SimLog.txt"))
    shutil.copy(os.path.join(our_root, "Raw", sub, "SimLog.txt"), os.path.join(out, name + "_ours-SimLog.txt"))
    shutil.rmtree(wd_ref, ignore_errors=True)
    shutil.rmtree(wd_our, ignore_errors=True)
    return res


def case_time(name, args, out, steps, ours_args=None, ours_steps=None):
    wd_ref = os.path.join(out, name, "ref")
    wd_our = os.path.join(out, name, "ours")
    for d in (wd_ref, wd_our):
        shutil.rmtree(d, ignore_errors=True)
        os.makedirs(d)
    # the product is ~50x faster per step: it gets a longer run so that the steps outweigh the
    # ~2 s of process start-up (CUDA context) both programs pay
    ours_args, ours_steps = ours_args or args, ours_steps or steps
    _, t_ref, r1 = run_reference(args, wd_ref)
    _, t_our, r2 = run_ours(ours_args, wd_our)
    # start-up (context, allocation, initial state) measured by a run of zero-ish length
    short = list(args)
    short[6], short[7] = "0.00002", "1"  # 2 iterations, INTERVAL 2 (P_num > ITERATIONS would
    #                                      make INTERVAL 0 and `it % INTERVAL` trap, main.cu:301,938)
    _, t_ref0, _ = run_reference(short, wd_ref)
    _, t_our0, _ = run_ours(short, wd_our)
    shutil.rmtree(os.path.join(out, name), ignore_errors=True)
    return {"args": " ".join(args), "steps": steps, "ours_args": " ".join(ours_args),
            "ours_steps": ours_steps,
            "reference_wall_s": round(t_ref, 3), "ours_wall_s": round(t_our, 3),
            "reference_startup_s": round(t_ref0, 3), "ours_startup_s": round(t_our0, 3),
            "reference_ms_per_step": round(1e3 * (t_ref - t_ref0) / steps, 4),
            "ours_ms_per_step": round(1e3 * (t_our - t_our0) / ours_steps, 5),
            "exit": [r1.returncode, r2.returncode]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "refprog"))
    ap.add_argument("--no-timing", action="store_true")
    a = ap.parse_args()
    a.out = os.path.abspath(a.out)
    if not os.path.exists(IBLB):
        print("oracle/_ref/IBLB not built (make -C oracle ref_bin)")
        return 1
    os.makedirs(a.out, exist_ok=True)
    summary = {"binary": "oracle/_ref/IBLB: nvcc -O3 sm_100a of the unmodified main.cu + "
                         "LatticeBoltzmann.cu + ImmersedBoundary.cu + seconds.cpp"}
    # cilia6.sh:7 (phase-lagged) and cilia6.sh:17 (in phase), 100 iterations, snapshots every 10
    summary["lagged"] = case_compare("lagged", "1 6 48 1.0 1.0 5 0.001 10 0 1".split(), a.out)
    summary["inphase"] = case_compare("inphase", "6 6 48 1.0 1.0 5 0.001 10 0 1".split(), a.out)
    if not a.no_timing:
        summary["time_c1"] = case_time("time_c1", "6 6 48 1.0 1.0 5 0.05 5 0 0".split(), a.out, 5000,
                                       "6 6 48 1.0 1.0 5 3.0 5 0 0".split(), 300000)
        summary["time_c3"] = case_time("time_c3", "12 12 48 1.0 1.0 5 0.05 5 0 0".split(), a.out, 5000,
                                       "12 12 48 1.0 1.0 5 3.0 5 0 0".split(), 300000)
    with open(os.path.join(a.out, "summary.json"), "w") as f:
        json.dump(summary, f, indent=1)
    # short markdown for profiles/
    L = ["# The unmodified reference program vs `./app` on the same argument lines", ""]
    for name in ("lagged", "inphase"):
        s = summary[name]
        L.append("## `%s`  (%s)" % (s["args"], name))
        L.append("")
        L.append("* exit codes reference/ours: %s/%s; wall %.2f s / %.2f s" % (
            s["reference_exit"], s["ours_exit"], s["reference_wall_s"], s["ours_wall_s"]))
        L.append("* SimLog.txt identical (date/runtime lines aside): **%s** (%d lines)" % (
            s["simlog"]["identical"], s["simlog"]["lines"]))
        if "flux" in s:
            fx = s["flux"]
            L.append("* flux file names: %s vs %s; text-identical: **%s**; first non-finite line "
                     "reference %s / ours %s; max rel diff of finite entries %.2e" % (
                         s["flux_file_names"][0], s["flux_file_names"][1], fx["text_identical"],
                         fx["reference_first_nonfinite_line"], fx["ours_first_nonfinite_line"],
                         fx["max_rel_diff_finite"]))
            L.append("")
            L.append("| line | reference | ours |")
            L.append("|---|---|---|")
            for i, (x, y) in enumerate(zip(fx["reference_lines"], fx["ours_lines"])):
                L.append("| %d | `%s` | `%s` |" % (i, x, y))
        L.append("")
        L.append("| it | cilia rows differing (marker 32 / elsewhere) | fluid rows text-identical | rel-L2 ux | uy | rho |")
        L.append("|---|---|---|---|---|---|")
        for it in s["fluid"]:
            fl, ci = s["fluid"][it], s["cilia"].get(it, {})
            L.append("| %s | %s / %s of %s | %s of %s | %s | %s | %s |" % (
                it, ci.get("rows_differing_marker32"), ci.get("rows_differing_elsewhere"),
                ci.get("markers"), fl.get("text_identical_rows"), fl.get("rows"),
                ("%.2e" % fl["rel_l2_ux"]) if "rel_l2_ux" in fl else "non-finite",
                ("%.2e" % fl["rel_l2_uy"]) if "rel_l2_uy" in fl else "-",
                ("%.2e" % fl["rel_l2_rho"]) if "rel_l2_rho" in fl else "-"))
        L.append("")
    for name in ("time_c1", "time_c3"):
        if name in summary:
            t = summary[name]
            L.append("* whole-binary timing, no snapshots: reference `%s` (%d steps) %.3f ms/step "
                     "(wall %.2f s, start-up %.2f s); ours `%s` (%d steps) %.5f ms/step (wall %.2f s, "
                     "start-up %.2f s)" % (
                         t["args"], t["steps"], t["reference_ms_per_step"], t["reference_wall_s"],
                         t["reference_startup_s"], t["ours_args"], t["ours_steps"],
                         t["ours_ms_per_step"], t["ours_wall_s"], t["ours_startup_s"]))
    with open(os.path.join(a.out, "summary.md"), "w") as f:
        f.write("\n".join(L) + "\n")
    print("\n".join(L))
    return 0


if __name__ == "__main__":
    sys.exit(main())
