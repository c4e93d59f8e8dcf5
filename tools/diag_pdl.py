"""Development diagnostic: which launch mode parts from the unfused chain, and when."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from cuda_iblb_11_b200 import FLAG_NO_FUSE, FLAG_NO_GRAPH, FLAG_NO_PDL, Params, Solver

cn, cs, cf, Y = 6, 48, 1, 192
X = cn * cs
kw = dict(c_num=cn, c_space=cs, c_fraction=cf)
modes = {"graph+pdl": 0, "direct+pdl": FLAG_NO_GRAPH, "graph": FLAG_NO_PDL,
         "direct": FLAG_NO_GRAPH | FLAG_NO_PDL}
b = Solver(Params.channel(X, Y, flags=FLAG_NO_FUSE | FLAG_NO_GRAPH, **kw))
S = {k: Solver(Params.channel(X, Y, flags=f, **kw)) for k, f in modes.items()}
for n in (1, 2, 37, 1, 1, 1, 1, 1, 1, 1, 1, 1):
    b.step(n)
    rb = b.populations()
    for k, s in S.items():
        s.step(n)
        ra = s.populations()
        nd = int((ra != rb).sum())
        print("after +%d (it %d) %-11s pop diffs %d nan fused %d nan chain %d" % (n, b.it, k, nd, int(np.isnan(ra).sum()), int(np.isnan(rb).sum())))
