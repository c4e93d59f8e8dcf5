"""Development (GPU box): same-box A/B of the pure-LB kernels on C2 (8192x8192, body force):
lb_fused_v2 (IBLB_V4=0) vs lb_fused_v4, interleaved three times, device time of 100 steps."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_iblb_11_b200 import FLAG_DEV_KNOBS, Params, Solver

X = Y = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
res = {}
for rep in range(3):
    for v in ("0", "1"):
        os.environ["IBLB_V4"] = v
        s = Solver(Params.channel(X, Y, gx=1e-7, flags=FLAG_DEV_KNOBS))
        s.step(20)
        s.sync()
        s.step(100)
        s.sync()
        t = s.last_step_seconds / 100
        s.close()
        res.setdefault(v, []).append(t)
        print("IBLB_V4=%s: %.4f ms/step, %.1f MLUPS, %.1f GB/s" % (
            v, 1e3 * t, X * Y / t / 1e6, 144 * X * Y / t / 1e9), flush=True)
for v, ts in res.items():
    print("best IBLB_V4=%s: %.4f ms" % (v, 1e3 * min(ts)))
