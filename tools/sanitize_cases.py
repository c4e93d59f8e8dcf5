#!/usr/bin/env python
"""Small runs of every launch mode against the range-checked build of the library (GPU box):

    python -m cuda_iblb_11_b200.build --checked
    IBLB_LIB=cuda_iblb_11_b200/libiblb_checked.so python tools/sanitize_cases.py c1_fused split1024 ...

(compute-sanitizer, for which these cases were written, is closed on the GPU pool:
profiles/r02_sanitizer_unavailable.md.)  The checked build counts, on the device, every index into
the population buffers, force band, marker arrays and shared-memory tables that leaves its range;
each case prints that count, which must be 0.

Cases: c1_fused (CUDA graph + programmatic dependent launch, 70 iterations), c1_nofuse (separate
kinematics / interpolate / spread / LB kernels), split1024 (tall grid: band tiles + lb_fused_v2 above),
slabs_local (two x-slabs on one GPU coupled through iblb_connect_local, cilia across the seam),
pure_lb (vector and generic kernels), getters (fields / force / populations / markers / checkpoint).
Each case ends by comparing against the unfused chain where that is cheap, so a sanitizer-clean
run is also a correct one."""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from cuda_iblb_11_b200 import (FLAG_NO_FUSE, FLAG_NO_GRAPH, Params, Solver, debug_violations,  # noqa: E402
                               slabs)


def same(a, b):
    for x, y in zip(a.fields() + a.markers(), b.fields() + b.markers()):
        assert np.array_equal(x, y)
    assert a.Q == b.Q


def c1_fused():
    a = Solver(Params.channel(288, 192, c_num=6, c_fraction=6))
    b = Solver(Params.channel(288, 192, c_num=6, c_fraction=6, flags=FLAG_NO_FUSE | FLAG_NO_GRAPH))
    for n in (3, 67):
        a.step(n)
        b.step(n)
    same(a, b)


def c1_nofuse():
    b = Solver(Params.channel(288, 192, c_num=6, c_fraction=1, flags=FLAG_NO_FUSE | FLAG_NO_GRAPH))
    b.step(20)
    assert np.isfinite(b.fields()[0]).all()


def split1024():
    kw = dict(c_num=4, c_space=128, c_fraction=4)
    a = Solver(Params.channel(512, 1024, **kw))
    b = Solver(Params.channel(512, 1024, flags=FLAG_NO_FUSE | FLAG_NO_GRAPH, **kw))
    for n in (2, 35):
        a.step(n)
        b.step(n)
    same(a, b)


def slabs_local():
    X, Y, kw = 288, 192, dict(c_num=6, c_fraction=6)
    whole = Solver(Params.channel(X, Y, **kw))
    parts = []
    for r in range(2):
        x0, w = slabs.slab_range(r, 2, X, 48)
        parts.append((x0, w, Solver(Params.channel(X, Y, slab_x0=x0, slab_w=w, **kw))))
    for r in range(2):
        parts[r][2].connect_local(parts[r - 1][2], parts[(r + 1) % 2][2])
    for _ in range(12):
        for (_, _, s) in parts:
            s.step(1)
    whole.step(12)
    rho = np.empty((Y, X))
    for (x0, w, s) in parts:
        rho[:, x0:x0 + w] = s.fields()[0].reshape(Y, w)
    assert np.array_equal(rho.ravel(), whole.fields()[0])


def pure_lb():
    for (X, Y) in ((256, 256), (96, 40)):
        s = Solver(Params.channel(X, Y, gx=1e-6, u0_amp=1e-3))
        s.step(10)
        assert np.isfinite(s.fields()[1]).all()
        s.close()


def getters():
    s = Solver(Params.channel(288, 192, c_num=6, c_fraction=6))
    s.step(5)
    s.fields(); s.force(); s.populations(); s.markers()
    with tempfile.TemporaryDirectory() as d:
        s.save_checkpoint(os.path.join(d, "a.ckp"))
        s.step(3)
        s.load_checkpoint(os.path.join(d, "a.ckp"))
    s.step(40)
    assert np.isfinite(s.fields()[0]).all()


if __name__ == "__main__":
    for name in sys.argv[1:]:
        globals()[name]()
        v = debug_violations()
        print("case %s ok, index violations so far: %s" % (
            name, "n/a (product build)" if v < 0 else v), flush=True)
        assert v <= 0, "range check failed"
