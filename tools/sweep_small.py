"""Development (GPU box): C1 / C3 fused step time over tile width and block order."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_iblb_11_b200 import FLAG_DEV_KNOBS, Params, Solver


def t_step(p, n=4000):
    s = Solver(p)
    s.step(200)
    s.sync()
    best = 1e9
    for _ in range(3):
        s.step(n)
        s.sync()
        best = min(best, s.last_step_seconds / n)
    s.close()
    return best * 1e6


for (X, cn) in ((288, 6), (576, 12), (1152, 24)):
    for tw in ("6", "8"):
        for bf in ("0", "1"):
            os.environ["IBLB_TILE_W"], os.environ["IBLB_BAND_FIRST"] = tw, bf
            t = t_step(Params.channel(X, 192, c_num=cn, c_space=48, c_fraction=cn, flags=FLAG_DEV_KNOBS))
            print("%4dx192 %2d cilia  tile width %s  band first %s : %.2f us" % (X, cn, tw, bf, t), flush=True)
