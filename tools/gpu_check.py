"""Development check on a GPU box: product vs oracle port vs reference CUDA build. Not a test."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from checkers import REF_FLAG_PICK_LOWEST_K, Checker, available, make_params, rel_l2
from cuda_iblb_11_b200 import FLAG_NO_GRAPH, FLAG_PICK_HIGHEST_K, Params, Solver


def cmp_case(name, X, Y, c_num=0, steps=(1, 10), gx=0., u0=0., kinds=("oracle", "ref_cuda"),
             flags=0, c_fraction=1):
    print("==", name, flush=True)
    s = Solver(Params.channel(X, Y, c_num=c_num, gx=gx, u0_amp=u0, flags=flags, c_fraction=c_fraction))
    cks = {}
    for k in kinds:
        if available(k):
            fl = 0
            if k == "oracle" and not (flags & FLAG_PICK_HIGHEST_K):
                fl = REF_FLAG_PICK_LOWEST_K
            cks[k] = Checker(k, make_params(X, Y, c_num=c_num, gx=gx, u0_amp=u0, flags=fl, c_fraction=c_fraction))
    done = 0
    for n in steps:
        s.step(n - done)
        rho, u = s.fields()
        for k, c in cks.items():
            c.step(n - done)
            r2, u2 = c.fields()
            line = "%s it=%d vs %s: rho %.3e u %.3e f %.3e" % (
                name, n, k, rel_l2(rho, r2), rel_l2(u, u2), rel_l2(s.populations(), c.populations()))
            if c_num:
                m1, m2 = s.markers(), c.markers()
                line += " s %.3e(%d) us %.3e Fs %.3e eps %d Q %.6e/%.6e force %.3e" % (
                    np.abs(m1[0] - m2[0]).max(), int((m1[0] != m2[0]).sum()),
                    np.abs(m1[1] - m2[1]).max(), rel_l2(m1[2], m2[2]),
                    int((m1[3] != m2[3]).sum()), s.Q, c.Q, rel_l2(s.force(), c.force()))
            print(line, flush=True)
        done = n
    s.close()
    for c in cks.values():
        c.close()


if __name__ == "__main__":
    cmp_case("c1", 288, 192, c_num=6, steps=(1, 2, 10, 20, 30))
    cmp_case("c1s", 288, 192, c_num=6, steps=(1, 10, 100, 1000), c_fraction=6)
    cmp_case("c3s", 576, 192, c_num=12, steps=(1, 10, 100), c_fraction=12)
    a = Solver(Params.channel(288, 192, c_num=6, c_fraction=6))
    b = Solver(Params.channel(288, 192, c_num=6, c_fraction=6, flags=FLAG_NO_GRAPH))
    a.step(21)
    b.step(21)
    print("graph vs direct:", [bool(np.array_equal(x, y)) for x, y in zip(a.fields(), b.fields())],
          a.Q == b.Q, a.it, b.it)
    for sol, nm in ((a, "graph"), (b, "direct")):
        sol.step(2000)
        sol.sync()
        t = sol.last_step_seconds
        print("C1 %s: %.2f us/step, %.1f MLUPS" % (nm, t / 2000 * 1e6, 288 * 192 * 2000 / t / 1e6))
    print("C1 phases:", b.profile_phases(50))
    for (X, Y, cn, cs) in ((576, 192, 12, 48), (8192, 192, 64, 128), (8192, 8192, 64, 128)):
        c = Solver(Params.channel(X, Y, c_num=cn, c_space=cs, c_fraction=cn, flags=FLAG_NO_GRAPH))
        c.step(20)
        ph = c.profile_phases(20)
        tot = sum(ph.values())
        print("%dx%d %d cilia phases: %s  IB share %.1f%%" % (
            X, Y, cn, ph, 100 * (tot - ph["lb"]) / tot))
        c.close()
