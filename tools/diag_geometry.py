"""Development diagnostic: pure-LB rate against the channel's aspect ratio at a fixed node count."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_iblb_11_b200 import Params, Solver

for X, Y, gx in ((8192, 8192, 1e-7), (8192, 8192, 0.0)):
    s = Solver(Params.channel(X, Y, gx=gx))
    s.step(10); s.sync()
    best = 1e9
    for _ in range(3):
        s.step(50); s.sync()
        best = min(best, s.last_step_seconds / 50)
    print("gx %g" % gx, "%6d x %6d: %.4f ms/step  %.0f MLUPS  %.1f GB/s" % (X, Y, 1e3 * best, X * Y / best / 1e6, 144 * X * Y / best / 1e9))
    s.close()
