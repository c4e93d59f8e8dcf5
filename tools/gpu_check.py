"""Development check on a GPU box: product vs oracle port vs reference CUDA build. Not a test."""
import os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from checkers import Checker, make_params, rel_l2, available, REF_FLAG_PICK_LOWEST_K
from cuda_iblb_11_b200 import Params, Solver, FLAG_PICK_LOWEST_K, FLAG_NO_GRAPH

def cmp_case(name, X, Y, c_num=0, steps=(1, 10, 100), gx=0., u0=0., kinds=("oracle", "ref_cuda"), flags=0):
    print("==", name, flush=True)
    p = Params.channel(X, Y, c_num=c_num, gx=gx, u0_amp=u0, flags=flags)
    s = Solver(p)
    cks = {}
    for k in kinds:
        if available(k):
            cks[k] = Checker(k, make_params(X, Y, c_num=c_num, gx=gx, u0_amp=u0,
                             flags=(REF_FLAG_PICK_LOWEST_K if (flags & FLAG_PICK_LOWEST_K) and k == "oracle" else 0)))
    done = 0
    for n in steps:
        s.step(n - done)
        rho, u = s.fields()
        for k, c in cks.items():
            c.step(n - done)
            r2, u2 = c.fields()
            line = "%s it=%d vs %s: rho %.3e u %.3e" % (name, n, k, rel_l2(rho, r2), rel_l2(u, u2))
            f1 = s.populations(); f2 = c.populations()
            line += " f %.3e" % rel_l2(f1, f2)
            if c_num:
                m1 = s.markers(); m2 = c.markers()
                line += " s %.3e(%d) us %.3e Fs %.3e eps %d Q %.6e/%.6e force %.3e" % (
                    np.abs(m1[0] - m2[0]).max(), int((m1[0] != m2[0]).sum()),
                    np.abs(m1[1] - m2[1]).max(), rel_l2(m1[2], m2[2]), int((m1[3] != m2[3]).sum()),
                    s.Q, c.Q, rel_l2(s.force(), c.force()))
            print(line, flush=True)
        done = n
    s.close()
    for c in cks.values(): c.close()

cmp_case("lb64", 64, 64, steps=(1, 2, 50), gx=1e-6, u0=1e-3)
cmp_case("lb_generic_96x40", 96, 40, steps=(1, 50), gx=1e-6, u0=1e-3, kinds=("oracle",))
cmp_case("c1_hi", 288, 192, c_num=6, steps=(1, 2, 10, 200))
cmp_case("c1_lo", 288, 192, c_num=6, steps=(1, 10), flags=FLAG_PICK_LOWEST_K, kinds=("oracle", "ref_cuda"))
cmp_case("c3", 576, 192, c_num=12, steps=(1, 50))

# graph vs direct
p = Params.channel(288, 192, c_num=6); a = Solver(p)
p2 = Params.channel(288, 192, c_num=6, flags=FLAG_NO_GRAPH); b = Solver(p2)
a.step(101); b.step(101)
print("graph vs direct:", [bool(np.array_equal(x, y)) for x, y in zip(a.fields(), b.fields())], a.Q == b.Q)
for sol, nm in ((a, "graph"), (b, "direct")):
    sol.step(2000); sol.sync(); t = sol.last_step_seconds
    print("C1 %s: %.2f us/step, %.1f MLUPS" % (nm, t / 2000 * 1e6, 288 * 192 * 2000 / t / 1e6))
print("C1 phases:", b.profile_phases(50))
a.close(); b.close()

# perf C2
for X in (2048, 8192):
    p = Params.channel(X, X, gx=1e-7)
    s = Solver(p)
    s.step(5); s.sync()
    s.step(20); s.sync()
    t = s.last_step_seconds
    print("C2 %dx%d: %.3f ms/step, %.1f MLUPS, %.1f GB/s" % (X, X, t / 20 * 1e3, X * X * 20 / t / 1e6, X * X * 20 * 144 / t / 1e9), flush=True)
    s.close()
if available("ref_cuda"):
    c = Checker("ref_cuda", make_params(8192, 8192, gx=1e-7))
    c.step(2); c.step(5)
    t = c.last_step_seconds
    print("ref_cuda C2 8192: %.3f ms/step %.1f MLUPS" % (t / 5 * 1e3, 8192 * 8192 * 5 / t / 1e6))
    c.close()
    c = Checker("ref_cuda", make_params(288, 192, c_num=6))
    c.step(10); c.step(200)
    t = c.last_step_seconds
    print("ref_cuda C1: %.2f us/step %.2f MLUPS" % (t / 200 * 1e6, 288 * 192 * 200 / t / 1e6))
