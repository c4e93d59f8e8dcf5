"""Two (or more) ranks, one GPU each: x-slabs coupled through CUDA IPC peer stores and device-side
step flags (iblb_export_halo / iblb_connect), checked against the whole channel on rank 0.
Launched by tests/test_gpu_slabs_fullsize.py via torch.distributed.run.  Development/test tool."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch.distributed as dist

from cuda_iblb_11_b200 import Params, Solver, slabs


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    dist.init_process_group("gloo")
    ok = True
    for (X, Y, kw, steps) in ((256 * world, 256, dict(gx=1e-6, u0_amp=1e-3), 40),
                              (288 * world, 192, dict(c_num=6 * world, c_fraction=6 * world), 25)):
        align = 48 if kw.get("c_num") else 1
        x0, w = slabs.slab_range(rank, world, X, align)
        s = Solver(Params.channel(X, Y, slab_x0=x0, slab_w=w, device=local, **kw))
        slabs.connect_ring(s, dist, rank, world)
        s.step(steps)  # ranks run free: ordering is by the device-side flags
        s.sync()
        rho, u = s.fields()
        out = [None] * world
        dist.gather_object((x0, w, rho, u, s.Q), out if rank == 0 else None, dst=0)
        dist.barrier()
        s.close()
        if rank == 0:
            whole = Solver(Params.channel(X, Y, device=local, **kw))
            whole.step(steps)
            r2, u2 = whole.fields()
            r2 = r2.reshape(Y, X)
            ux2, uy2 = u2[:X * Y].reshape(Y, X), u2[X * Y:].reshape(Y, X)
            q = 0.0
            for (a, b, r, uu, qq) in out:
                ok &= np.array_equal(r.reshape(Y, b), r2[:, a:a + b])
                ok &= np.array_equal(uu[:b * Y].reshape(Y, b), ux2[:, a:a + b])
                ok &= np.array_equal(uu[b * Y:].reshape(Y, b), uy2[:, a:a + b])
                q += qq
            ok &= (q == whole.Q)
            whole.close()
            print("case %dx%d %s: %s" % (X, Y, kw, "ok" if ok else "MISMATCH"), flush=True)
    if rank == 0:
        print("IPC SLABS OK" if ok else "IPC SLABS FAILED", flush=True)
    dist.barrier()
    dist.destroy_process_group()
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
