"""Development diagnostic: first step / nodes where the fused step parts from the 4-kernel chain."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from cuda_iblb_11_b200 import FLAG_NO_FUSE, FLAG_NO_GRAPH, Params, Solver

for (cn, cs, cf, Y) in ((6, 48, 1, 192), (6, 48, 6, 192), (12, 48, 5, 192), (16, 128, 16, 192)):
    X = cn * cs
    kw = dict(c_num=cn, c_space=cs, c_fraction=cf)
    a = Solver(Params.channel(X, Y, flags=FLAG_NO_GRAPH, **kw))
    b = Solver(Params.channel(X, Y, flags=FLAG_NO_FUSE | FLAG_NO_GRAPH, **kw))
    for it in range(1, 40):
        a.step(1)
        b.step(1)
        fa, fb = a.force(), b.force()
        ra, rb = a.fields()[0], b.fields()[0]
        if not np.array_equal(fa, fb) or not np.array_equal(ra, rb):
            d = np.nonzero(fa != fb)[0]
            print("case", cn, cs, cf, "first mismatch at it", it, "force diffs", len(d),
                  "rho diffs", int((ra != rb).sum()))
            size = X * Y
            pa, pb = a.populations().reshape(-1, 9), b.populations().reshape(-1, 9)
            for j in np.nonzero(ra != rb)[0][:4]:
                y, x = divmod(int(j), X)
                print("   rho node x %d y %d fused %.17g chain %.17g" % (x, y, ra[j], rb[j]))
            bad = np.nonzero((pa != pb).any(axis=1))[0]
            for j in bad[:6]:
                y, x = divmod(int(j), X)
                print("   pop node x %d y %d which %s diff %s" % (x, y, np.nonzero(pa[j] != pb[j])[0],
                      (pa[j] - pb[j])[pa[j] != pb[j]]))
            sm = b.markers()
            S = sm[0].reshape(-1, 2)
            for j in bad[:3]:
                y, x = divmod(int(j), X)
                near = np.nonzero((np.abs(S[:, 0] - x) <= 3.5) & (np.abs(S[:, 1] - y) <= 3.5))[0]
                print("   markers within 3.5 of it:", [(int(k), round(float(S[k, 0]), 3), round(float(S[k, 1]), 3)) for k in near])
            for j in d[:12]:
                c, jj = divmod(int(j), size)
                y, x = divmod(jj, X)
                print("   comp %d x %d y %d fused %.17g chain %.17g" % (c, x, y, fa[j], fb[j]))
            s = b.markers()[0].reshape(-1, 2)
            if len(d):
                c, jj = divmod(int(d[0]), size)
                y, x = divmod(jj, X)
                near = np.nonzero((np.abs(s[:, 0] - x) <= 1.5) & (np.abs(s[:, 1] - y) <= 1.5))[0]
                print("   markers near first node:", [(int(k), float(s[k, 0]), float(s[k, 1])) for k in near],
                      "eps", b.markers()[3][near])
            break
    else:
        print("case", cn, cs, cf, "identical for 39 steps")
    a.close()
    b.close()
