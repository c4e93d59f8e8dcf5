"""Development timing (GPU box): per-step time of the cilia cases, spreading fused into the LB kernel (2 launches) vs the
unfused chain (3 launches) vs pure LB on the same grid.  IB share = (t_ib_run - t_pure_lb) / t_ib_run."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_iblb_11_b200 import FLAG_DEV_KNOBS, FLAG_NO_FUSE, FLAG_NO_GRAPH, Params, Solver


def t_step(p, n, warm):
    s = Solver(p)
    s.step(warm)
    s.sync()
    best = 1e9
    for _ in range(3):
        s.step(n)
        s.sync()
        best = min(best, s.last_step_seconds / n)
    s.close()
    return best * 1e6


if __name__ == "__main__":
    cases = [(288, 192, 6, 48, 4000), (576, 192, 12, 48, 4000), (2304, 192, 48, 48, 2000),
             (8192, 192, 64, 128, 1000), (1152, 192, 24, 48, 3000), (4608, 192, 96, 48, 1500),
             (8192, 2048, 64, 128, 200), (8192, 8192, 64, 128, 40)]
    if len(sys.argv) > 1:
        cases = cases[:int(sys.argv[1])]
    for (X, Y, cn, cs, n) in cases:
        kw = dict(c_num=cn, c_space=cs, c_fraction=cn)
        # both block orders (iblb_create picks one by the grid's size; IBLB_BAND_FIRST forces)
        tf = []
        for bf in ("0", "1"):
            os.environ["IBLB_BAND_FIRST"] = bf
            tf.append(t_step(Params.channel(X, Y, flags=FLAG_DEV_KNOBS, **kw), n, 50))
        del os.environ["IBLB_BAND_FIRST"]
        print("      block order of lb_ib_tile_step: row tile fastest %.2f us, band tiles first %.2f us" %
              tuple(tf), flush=True)
        fused = t_step(Params.channel(X, Y, **kw), n, 50)
        fused_direct = t_step(Params.channel(X, Y, flags=FLAG_NO_GRAPH, **kw), n, 50)
        chain = t_step(Params.channel(X, Y, flags=FLAG_NO_FUSE, **kw), n, 50)
        lb = t_step(Params.channel(X, Y, gx=1e-7), n, 50)
        ph = Solver(Params.channel(X, Y, flags=FLAG_NO_GRAPH, **kw))
        ph.step(20)
        phases = ph.profile_phases(50)
        ph.close()
        print("      event-timed phases of the unfused chain (incl. ~2 us launch gap each): "
              "markers %.2f spread %.2f lb %.2f us" % (
                  1e6 * (phases["kinematics"] + phases["interpolate"]), 1e6 * phases["spread"],
                  1e6 * phases["lb"]), flush=True)
        print("%5dx%-5d %3d cilia: fused %.2f us (direct launches %.2f), unfused chain %.2f, "
              "pure LB %.2f us/step | %.0f MLUPS | IB share fused %.1f%% chain %.1f%%" % (
                  X, Y, cn, fused, fused_direct, chain, lb, X * Y / fused,
                  100 * (fused - lb) / fused, 100 * (chain - lb) / chain), flush=True)
