"""The product's `app` against the UNMODIFIED reference program (oracle/_ref/IBLB: main() and all,
built from the reference's own four source files by oracle/Makefile ref_bin) on the same argument
line, file by file (main.cu:758-790, 958-969, 988-992, 1002, 1030-1034).  Needs a B200.

The reference program is not deterministic (marker 32's write race, main.cu:158-172; the final flux
line, main.cu:1026-1034), so this file is named to collect AFTER every deterministic GPU test: under
`pytest -x` a surprise from the reference's side must not hide them."""
import os
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))


@pytest.mark.parametrize("name,line", [
    ("inphase", "6 6 48 1.0 1.0 5 0.001 10 0 1"),   # cilia6.sh:17, 100 iterations, 10 snapshots
    ("lagged", "1 6 48 1.0 1.0 5 0.0003 3 0 1"),    # cilia6.sh:7, 30 iterations (it overflows at 42)
])
def test_data_tree_matches_the_reference_program(tmp_path, name, line):
    import run_reference_program as rp
    if not os.path.exists(rp.IBLB) or not os.path.exists(rp.APP):
        pytest.skip("oracle/_ref/IBLB or app not built")
    r = rp.case_compare(name, line.split(), str(tmp_path))
    assert r["reference_exit"] == 0 and r["ours_exit"] == 0, r
    assert r["simlog"]["identical"], (r["simlog"]["reference"], r["simlog"]["ours"])
    assert r["flux_file_names"][0] == r["flux_file_names"][1]
    assert len(r["fluid"]) >= 3 and len(r["cilia"]) == len(r["fluid"])
    for it, c in r["cilia"].items():
        # every marker row identical as text except the racing marker 32 of each cilium (Q12)
        assert c["same_count"] and c["rows_differing_elsewhere"] == 0, (it, c)
    f0 = r["fluid"]["0"]
    assert f0["text_identical_rows"] == f0["rows"]  # rest state: byte for byte
    for it, f in r["fluid"].items():
        assert f["reference_finite"], it
        # 6 significant digits in the file; marker 32 (1 in 96, 0.0116 l.u. off when the other
        # fine point wins) moves u by up to 5e-3 relative, rho by far less
        assert f["rel_l2_rho"] <= 1e-5 and f["rel_l2_ux"] <= 1e-2 and f["rel_l2_uy"] <= 1e-2, (it, f)
    fx = r["flux"]
    assert fx["reference_first_nonfinite_line"] is None and fx["ours_first_nonfinite_line"] is None
    assert len(fx["reference_lines"]) == len(fx["ours_lines"])
    table = list(zip(fx["reference_lines"], fx["ours_lines"], fx["rel_diff_per_line"]))
    # Lines written inside the loop (the reference waits for its copy of Q there, main.cu:998).  In
    # phase the two programs agreed to 2e-6 on every line where this was first measured (six printed
    # digits, racing marker 32 included); the phase-lagged line is the unstable one (the reference's
    # coupling amplifies any difference ~1.25x per step until it overflows near step 42, DESIGN.md
    # section 6): whichever fine point wins marker 32 in this particular run of the reference shows
    # in the flux after 20 steps.
    tol = 1e-4 if name == "inphase" else 0.3
    assert fx["max_rel_diff_in_loop_lines"] <= tol, table
    # The last line: the reference prints a host variable whose last asynchronous copy it never waits
    # for (main.cu:1026-1034 after cudaStreamDestroy; in-loop copies main.cu:925-932), so it shows the
    # flux after iteration ITERATIONS-1 or after one of the few before it, as the timing of that run
    # has it (one step back is 1.27 % here: GPUTEST_r02).  Ours must be the flux after ITERATIONS-1;
    # the reference's must be one of the last four.
    from cuda_iblb_11_b200 import Params, Solver
    prm = Params.from_args(line.split())
    s = Solver(prm)
    s.step(prm.ITERATIONS)
    last = s.flux_history(prm.ITERATIONS - 4, 4)[::-1] * prm.x_scale   # k = 0, 1, 2, 3 steps back
    s.close()
    rel = lambda a, b: abs(a - b) / max(abs(b), 1e-300)
    assert rel(fx["ours_last_value"], last[0]) <= 1e-5, (fx["ours_last_value"], last.tolist())
    stale = [rel(fx["reference_last_value"], q) for q in last]
    assert min(stale) <= tol, (fx["reference_last_value"], last.tolist(), stale)
