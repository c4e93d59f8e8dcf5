"""Generate tests/golden/*.npz from the REFERENCE'S OWN KERNELS (test infrastructure).

Two sources, both built from the unmodified sources under /root/reference by oracle/Makefile:
  --kind ref_cuda : reference kernels compiled by nvcc for sm_100a, run on a B200 (`gpurun`);
                    this is the authoritative fixture set (`golden_refcuda_*.npz`).
  --kind ref_cpu  : the same kernels compiled as host C++ through oracle/shim (CPU container).

Usage (GPU box):  python tests/golden/make_golden.py --kind ref_cuda --out gpurun_out/golden
then copy the .npz files into tests/golden/.

Cilia cases record, for EVERY step:
  replay32[it, m, 0:4]  marker 32 of cilium m as the reference produced it (s_x, s_y, u_s_x,
                        u_s_y).  Its two candidate fine points sit in different warps (SURVEY
                        Q12), so the winner — and, the four stores being separate instructions,
                        possibly a mix of both — is timing dependent.  The oracle replays it.
  trace32[it, m]        0: equals the lower candidate, 1: the higher one, 2: a torn mix
  chk[it, 0:5]          exact float64 sums of s, u_s (marker 32 excluded), F_s, eps and Q: lets the
                        test follow the reference step by step without storing every array
and at the checkpoint steps the marker arrays plus a fixed subset of the Eulerian fields.
"""
import argparse
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from checkers import (REF_FLAG_PICK_LOWEST_K, Checker, make_params)  # noqa: E402

BAND = 128  # rows of the fields kept for the cilia cases (the IB touches rows < 96)


def lb_case(kind, X, Y, steps, gx, u0):
    c = Checker(kind, make_params(X, Y, gx=gx, u0_amp=u0))
    c.step(steps)
    rho, u = c.fields()
    out = dict(XDIM=X, YDIM=Y, steps=steps, gx=gx, u0_amp=u0, rho=rho, u=u, f=c.populations())
    c.close()
    return out


def step_checksums(markers, Q, c_num):
    s, u_s, F_s, eps = markers
    mask = np.ones(96 * c_num, bool)
    mask[32::96] = False
    s2, u2 = s.reshape(-1, 2)[mask], u_s.reshape(-1, 2)[mask]
    return np.array([s2.astype(np.float64).sum(), u2.astype(np.float64).sum(),
                     np.abs(F_s.astype(np.float64)).sum(), float(eps.sum()), Q])


def cilia_case(kind, c_num, c_fraction, checkpoints):
    X, Y = 48 * c_num, 192
    c = Checker(kind, make_params(X, Y, c_num=c_num, c_fraction=c_fraction))
    # kinematics do not depend on the fluid: the two oracle runs give both candidates of every
    # racing marker at every step
    lo = Checker("oracle", make_params(X, Y, c_num=c_num, c_fraction=c_fraction,
                                       flags=REF_FLAG_PICK_LOWEST_K))
    hi = Checker("oracle", make_params(X, Y, c_num=c_num, c_fraction=c_fraction))
    nsteps = max(checkpoints)
    trace = np.zeros((nsteps, c_num), np.uint8)
    replay = np.zeros((nsteps, c_num, 4), np.float32)
    chk = np.zeros((nsteps, 5))
    lowest_elsewhere = True
    n_elsewhere_mismatch = 0
    out = dict(XDIM=X, YDIM=Y, c_num=c_num, c_fraction=c_fraction,
               checkpoints=np.array(checkpoints))
    sel = np.random.default_rng(1234).choice(X * BAND, 8192, replace=False)
    sel.sort()
    out["sel"] = sel
    for it in range(nsteps):
        c.step(1)
        lo.step(1)
        hi.step(1)
        mk = c.markers()
        s, us = mk[0].reshape(-1, 2), mk[1].reshape(-1, 2)
        mlo, mhi = lo.markers(), hi.markers()
        slo, ulo = mlo[0].reshape(-1, 2), mlo[1].reshape(-1, 2)
        shi, uhi = mhi[0].reshape(-1, 2), mhi[1].reshape(-1, 2)
        for m in range(c_num):
            k = m * 96 + 32
            replay[it, m, 0:2] = s[k]
            replay[it, m, 2:4] = us[k]
            is_lo = np.array_equal(s[k], slo[k]) and np.array_equal(us[k], ulo[k])
            is_hi = np.array_equal(s[k], shi[k]) and np.array_equal(us[k], uhi[k])
            trace[it, m] = 0 if is_lo else (1 if is_hi else 2)
        mask = np.ones(len(s), bool)
        mask[32::96] = False
        bad = int((s[mask] != slo[mask]).any(axis=1).sum() + (us[mask] != ulo[mask]).any(axis=1).sum())
        n_elsewhere_mismatch += bad
        if np.abs(s[mask] - slo[mask]).max() > 1e-4:
            lowest_elsewhere = False
        chk[it] = step_checksums(mk, c.Q, c_num)
        n = it + 1
        if n in checkpoints:
            rho, u = c.fields()
            size = X * Y
            out["rho_%d" % n] = rho[sel]
            out["ux_%d" % n] = u[sel]
            out["uy_%d" % n] = u[size + sel]
            out["ux_rowsum_%d" % n] = u[:size].reshape(Y, X).sum(axis=1)
            out["ux_colsum_%d" % n] = u[:size].reshape(Y, X).sum(axis=0)
            out["uy_colsum_%d" % n] = u[size:].reshape(Y, X).sum(axis=0)
            out["s_%d" % n], out["u_s_%d" % n], out["F_s_%d" % n], out["eps_%d" % n] = mk
            out["Q_%d" % n] = c.Q
            out["rho_sum_%d" % n] = rho.sum()
    out["trace32"] = trace
    out["replay32"] = replay
    out["chk"] = chk
    out["lowest_k_wins_elsewhere"] = lowest_elsewhere
    out["n_elsewhere_mismatch"] = n_elsewhere_mismatch
    for x in (c, lo, hi):
        x.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--kind", default="ref_cuda", choices=["ref_cuda", "ref_cpu"])
    ap.add_argument("--out", default=HERE)
    a = ap.parse_args()
    os.makedirs(a.out, exist_ok=True)
    tag = "refcuda" if a.kind == "ref_cuda" else "refcpu"
    np.savez_compressed(os.path.join(a.out, "golden_%s_lb64.npz" % tag),
                        **lb_case(a.kind, 64, 64, 50, 1e-6, 1e-3))
    # cilia6.sh first line (c_fraction 1): the reference's coupling is unstable for phase-lagged
    # cilia and overflows after ~40 steps -> short horizon
    np.savez_compressed(os.path.join(a.out, "golden_%s_c1.npz" % tag),
                        **cilia_case(a.kind, 6, 1, (1, 2, 5, 10, 15, 20, 30)))
    # cilia6.sh last line (c_fraction 6, all cilia in phase): stable -> long horizon
    np.savez_compressed(os.path.join(a.out, "golden_%s_c1s.npz" % tag),
                        **cilia_case(a.kind, 6, 6, (1, 10, 100, 300, 600)))
    np.savez_compressed(os.path.join(a.out, "golden_%s_c3.npz" % tag),
                        **cilia_case(a.kind, 12, 5, (1, 5, 10, 20)))
    # cilia12-size grid, in phase (c_fraction 12): multi-block boundary_check (Q13), stable
    np.savez_compressed(os.path.join(a.out, "golden_%s_c3s.npz" % tag),
                        **cilia_case(a.kind, 12, 12, (1, 10, 100, 300)))
    for f in sorted(os.listdir(a.out)):
        print(f, os.path.getsize(os.path.join(a.out, f)))


if __name__ == "__main__":
    main()
