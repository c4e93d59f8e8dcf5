"""Development: 100 fused C3 steps with the library given as argv[1] (for an ncu launch list)."""
import ctypes
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_iblb_11_b200.solver import Params
L = ctypes.CDLL(sys.argv[1])
p = Params()
L.iblb_params_default(ctypes.byref(p))
p.c_fraction, p.c_num, p.c_space = 12, 12, 48
p.Re, p.T_num, p.T_pow, p.I_pow, p.P_num = 1.0, 1.0, 5, 1.0, 100
p.XDIM, p.YDIM = 576, 192
assert L.iblb_params_derive(ctypes.byref(p)) == 0
p.device = 0
h = ctypes.c_void_p()
assert L.iblb_create(ctypes.byref(p), ctypes.byref(h)) == 0
L.iblb_step(h, 100)
L.iblb_sync(h)
L.iblb_destroy(h)
