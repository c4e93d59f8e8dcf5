"""The product's `app` against the UNMODIFIED reference program (oracle/_ref/IBLB: main() and all,
built from the reference's own four source files by oracle/Makefile ref_bin) on the same argument
line, file by file (main.cu:758-790, 958-969, 988-992, 1002, 1030-1034).  Needs a B200."""
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
    # in phase the racing marker moves the flux by < 1 %; the phase-lagged line is the unstable one
    # (the reference's coupling amplifies any difference ~1.25x per step until it overflows near
    # step 42): whichever fine point wins marker 32 in this particular run of
    # the reference shows in the flux after 20 steps
    assert fx["max_rel_diff_finite"] <= (1e-2 if name == "inphase" else 0.3)
