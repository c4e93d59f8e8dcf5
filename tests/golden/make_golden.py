"""Generate tests/golden/*.npz from the REFERENCE'S OWN KERNELS (test infrastructure).

Two sources, both built from the unmodified sources under /root/reference by oracle/Makefile:
  --kind ref_cuda : reference kernels compiled by nvcc for sm_100a, run on a B200 (`gpurun`);
                    this is the authoritative fixture set (`golden_refcuda_*.npz`).
  --kind ref_cpu  : the same kernels compiled as host C++ through oracle/shim (CPU container).

Usage (GPU box):  python tests/golden/make_golden.py --kind ref_cuda --out gpurun_out/golden
then copy the .npz files into tests/golden/.

For the cilia cases the script also records, per step and cilium, which of the two candidate fine
points won the reference's write race on marker 32 (they sit in different warps, the
winner is timing dependent), by comparing the reference's marker positions with the oracle port
run both ways.  The oracle replays that trace, so oracle-vs-golden is an exact pin of the whole
IB path against the reference as it ran on the B200.
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


def cilia_case(kind, c_num, c_fraction, checkpoints):
    X, Y = 48 * c_num, 192
    c = Checker(kind, make_params(X, Y, c_num=c_num, c_fraction=c_fraction))
    lo = Checker("oracle", make_params(X, Y, c_num=c_num, c_fraction=c_fraction,
                                       flags=REF_FLAG_PICK_LOWEST_K))
    hi = Checker("oracle", make_params(X, Y, c_num=c_num, c_fraction=c_fraction))
    nsteps = max(checkpoints)
    trace = np.zeros((nsteps, c_num), np.uint8)
    lowest_elsewhere = True
    out = dict(XDIM=X, YDIM=Y, c_num=c_num, c_fraction=c_fraction,
               checkpoints=np.array(checkpoints))
    for it in range(nsteps):
        c.step(1)
        lo.step(1)
        hi.step(1)
        s = c.markers()[0].reshape(-1, 2)
        slo = lo.markers()[0].reshape(-1, 2)
        shi = hi.markers()[0].reshape(-1, 2)
        for m in range(c_num):
            k = m * 96 + 32
            if np.array_equal(s[k], shi[k]) and not np.array_equal(s[k], slo[k]):
                trace[it, m] = 1
        mask = np.ones(len(s), bool)
        mask[32::96] = False
        # positions differ from the oracle only through libm/libdevice rounding (rare, 1 ulp)
        if np.abs(s[mask] - slo[mask]).max() > 1e-4:
            lowest_elsewhere = False
        n = it + 1
        if n in checkpoints:
            rho, u = c.fields()
            size = X * Y
            mk = c.markers()
            # a fixed subset of band nodes (full fields would be MBs) + the first moments
            sel = np.random.default_rng(1234).choice(X * BAND, 8192, replace=False)
            sel.sort()
            out["sel"] = sel
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
    out["lowest_k_wins_elsewhere"] = lowest_elsewhere
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
    np.savez_compressed(os.path.join(a.out, "golden_%s_c1.npz" % tag),
                        **cilia_case(a.kind, 6, 1, (1, 2, 5, 10, 20)))
    np.savez_compressed(os.path.join(a.out, "golden_%s_c3.npz" % tag),
                        **cilia_case(a.kind, 12, 5, (1, 5, 10)))
    for f in sorted(os.listdir(a.out)):
        print(f, os.path.getsize(os.path.join(a.out, f)))


if __name__ == "__main__":
    main()
