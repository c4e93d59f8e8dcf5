"""Development diagnostic (GPU box): follow the reference's CUDA build step by step against the
oracle port (marker 32 replayed from the reference) and against the product, and report where
they first part.  Not a test."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from checkers import REF_FLAG_PICK_LOWEST_K, Checker, make_params, rel_l2
from cuda_iblb_11_b200 import Params, Solver


def follow(c_num, c_fraction, nsteps, with_product=True):
    X, Y = 48 * c_num, 192
    print("== c_num %d c_fraction %d" % (c_num, c_fraction), flush=True)
    ref = Checker("ref_cuda", make_params(X, Y, c_num=c_num, c_fraction=c_fraction))
    orc = Checker("oracle", make_params(X, Y, c_num=c_num, c_fraction=c_fraction,
                                        flags=REF_FLAG_PICK_LOWEST_K))
    lo = Checker("oracle", make_params(X, Y, c_num=c_num, c_fraction=c_fraction,
                                       flags=REF_FLAG_PICK_LOWEST_K))
    hi = Checker("oracle", make_params(X, Y, c_num=c_num, c_fraction=c_fraction))
    prod = Solver(Params.channel(X, Y, c_num=c_num, c_fraction=c_fraction)) if with_product else None
    replay = np.zeros((nsteps, c_num, 4), np.float32)
    for it in range(nsteps):
        ref.step(1)
        lo.step(1)
        hi.step(1)
        mr = ref.markers()
        s, us = mr[0].reshape(-1, 2), mr[1].reshape(-1, 2)
        for m in range(c_num):
            replay[it, m, :2] = s[m * 96 + 32]
            replay[it, m, 2:] = us[m * 96 + 32]
        orc.set_marker32_replay(replay[:it + 1])
        orc.step(1)
        mo = orc.markers()
        tags = []
        mlo, mhi = lo.markers(), hi.markers()
        for m in range(c_num):
            k = m * 96 + 32
            is_lo = all(np.array_equal(mr[i].reshape(-1, 2)[k], mlo[i].reshape(-1, 2)[k]) for i in (0, 1))
            is_hi = all(np.array_equal(mr[i].reshape(-1, 2)[k], mhi[i].reshape(-1, 2)[k]) for i in (0, 1))
            tags.append("L" if is_lo else ("H" if is_hi else "T"))
        ds = int((mr[0] != mo[0]).sum())
        du = int((mr[1] != mo[1]).sum())
        de = int((mr[3] != mo[3]).sum())
        rr, ur = ref.fields()
        ro, uo = orc.fields()
        line = "it %3d m32 %s | oracle: s!= %d us!= %d eps!= %d Fs %.2e rho %.2e u %.2e Q %.3e" % (
            it + 1, "".join(tags), ds, du, de, rel_l2(mo[2], mr[2]), rel_l2(ro, rr), rel_l2(uo, ur),
            abs(orc.Q - ref.Q) / max(abs(ref.Q), 1e-300))
        if ds or du:
            bad = np.nonzero((mr[0] != mo[0]) | (mr[1] != mo[1]))[0][:6] // 2
            line += " first bad markers %s" % sorted(set(int(b) for b in bad))
        if prod is not None:
            prod.step(1)
            mp = prod.markers()
            rp, up = prod.fields()
            mask = np.ones(96 * c_num * 2, bool)
            for m in range(c_num):
                mask[2 * (m * 96 + 32):2 * (m * 96 + 32) + 2] = False
            line += " | product: s!= %d us!= %d (excl m32) Fs %.2e rho %.2e u %.2e" % (
                int((mp[0] != mr[0])[mask].sum()), int((mp[1] != mr[1])[mask].sum()),
                rel_l2(mp[2], mr[2]), rel_l2(rp, rr), rel_l2(up, ur))
        print(line, flush=True)
        if not np.isfinite(rr).all():
            print("reference fields are no longer finite at it", it + 1)
            break
    for x in (ref, orc, lo, hi):
        x.close()
    if prod is not None:
        prod.close()


if __name__ == "__main__":
    follow(6, 1, 45)
    follow(6, 6, 60)
    follow(12, 5, 25)
    follow(12, 12, 40)
