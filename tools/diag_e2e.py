"""Development diagnostic: where the end-to-end time of the C2 case goes (upload, steps, download)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from cuda_iblb_11_b200 import Params, Solver

X = Y = 8192
n = X * Y
rho = torch.empty(n, dtype=torch.float64, pin_memory=True).numpy()
u = torch.empty(2 * n, dtype=torch.float64, pin_memory=True).numpy()
rho[:] = 1.0
u[:] = 0.0
s = Solver(Params.channel(X, Y, gx=1e-7))
s.step(5); s.sync()
for rep in range(3):
    t0 = time.perf_counter(); s.set_fields(rho, u); s.sync(); t1 = time.perf_counter()
    s.step(20); s.sync(); t2 = time.perf_counter()
    for _ in range(20):
        s.step(1); q = s.Q
    t3 = time.perf_counter()
    s.fields(rho, u); t4 = time.perf_counter()
    print("set_fields %.1f ms (%.1f GB/s)  step(20) %.2f ms/step  20x[step(1)+Q] %.2f ms/step  fields %.1f ms (%.1f GB/s)" % (
        1e3 * (t1 - t0), 24 * n / (t1 - t0) / 1e9, 1e3 * (t2 - t1) / 20, 1e3 * (t3 - t2) / 20, 1e3 * (t4 - t3), 24 * n / (t4 - t3) / 1e9))
