"""Development diagnostic: per-step time of the C3 case against the number of steps per call."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_iblb_11_b200 import FLAG_NO_PDL, Params, Solver

for name, fl in (("pdl", 0), ("no_pdl", FLAG_NO_PDL)):
    s = Solver(Params.channel(576, 192, c_num=12, c_space=48, c_fraction=12, flags=fl))
    s.step(64); s.sync()
    for n in (32, 64, 64, 256, 1024, 4096, 64, 32):
        t0 = time.perf_counter()
        s.step(n); s.sync()
        w = time.perf_counter() - t0
        print("%-7s step(%4d): device %.2f us/step, wall %.2f us/step" % (name, n, s.last_step_seconds * 1e6 / n, w * 1e6 / n))
    s.close()
