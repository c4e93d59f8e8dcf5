"""Development (GPU box): per-step time of the C1/C3 fused step with two builds of the library,
interleaved on the same box:  python tools/ab_lib.py libA.so libB.so"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_iblb_11_b200.solver import Params


def load(path):
    L = ctypes.CDLL(path)
    L.iblb_last_error.restype = ctypes.c_char_p
    return L


def t_step(L, X, Y, c_num, c_space, n, gx=0.0):
    p = Params()
    L.iblb_params_default(ctypes.byref(p))
    p.c_fraction, p.c_num, p.c_space = max(1, c_num), c_num, c_space
    p.Re, p.T_num, p.T_pow, p.I_pow, p.P_num = 1.0, 1.0, 5, 1.0, 100
    p.XDIM, p.YDIM = X, Y
    assert L.iblb_params_derive(ctypes.byref(p)) == 0
    p.gx, p.device = gx, 0
    h = ctypes.c_void_p()
    assert L.iblb_create(ctypes.byref(p), ctypes.byref(h)) == 0, L.iblb_last_error()
    L.iblb_step(h, 200)
    L.iblb_sync(h)
    best = 1e9
    for _ in range(3):
        L.iblb_step(h, n)
        L.iblb_sync(h)
        t = ctypes.c_double()
        L.iblb_last_step_seconds(h, ctypes.byref(t))
        best = min(best, t.value / n)
    L.iblb_destroy(h)
    return best * 1e6


libs = [(os.path.basename(a), load(os.path.abspath(a))) for a in sys.argv[1:]]
for (X, Y, cn, cs) in ((288, 192, 6, 48), (576, 192, 12, 48), (576, 192, 0, 48)):
    for rep in range(2):
        for name, L in libs:
            print("%dx%d %2d cilia  %-24s %.2f us/step" % (
                X, Y, cn, name, t_step(L, X, Y, cn, cs, 4000, gx=0.0 if cn else 1e-7)), flush=True)
