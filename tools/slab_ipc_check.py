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


def run_ahead_case(rank, world, local):
    """A rank that stops to read its fields while its neighbours run ahead: the getters pull from
    the previous source buffer, ghost columns included, which is exactly what a neighbour one step
    ahead overwrites.  The release counters of slab_wait hold that neighbour back until this rank
    steps again, so what is read is the state of the rank's own iteration — no barrier needed."""
    X, Y, kw = 192 * world, 192, dict(c_num=4 * world, c_fraction=4 * world)
    x0, w = slabs.slab_range(rank, world, X, 48)
    s = Solver(Params.channel(X, Y, slab_x0=x0, slab_w=w, device=local, **kw))
    slabs.connect_ring(s, dist, rank, world)
    n1, n2 = 9, 16
    if rank == 0:
        s.step(n1)
        got = [s.fields() for _ in range(3)][-1]   # read (three times: widen the window) ...
        pops = s.populations()
        s.step(n2 - n1)                            # ... then catch up
    else:
        s.step(n2)                                 # everybody else enqueues all steps at once
        got = pops = None
    s.sync()
    end = s.fields()
    out = [None] * world
    dist.gather_object((x0, w, end[0]), out if rank == 0 else None, dst=0)
    dist.barrier()
    s.close()
    ok = True
    if rank == 0:
        whole = Solver(Params.channel(X, Y, device=local, **kw))
        whole.step(n1)
        r1, u1 = whole.fields()
        p1 = whole.populations().reshape(Y, X, 9)[:, x0:x0 + w]
        ok &= np.array_equal(got[0].reshape(Y, w), r1.reshape(Y, X)[:, x0:x0 + w])
        ok &= np.array_equal(got[1][:w * Y].reshape(Y, w), u1[:X * Y].reshape(Y, X)[:, x0:x0 + w])
        ok &= np.array_equal(pops.reshape(Y, w, 9), p1)
        whole.step(n2 - n1)
        r2 = whole.fields()[0].reshape(Y, X)
        for (a, b, r) in out:
            ok &= np.array_equal(r.reshape(Y, b), r2[:, a:a + b])
        whole.close()
        print("case run-ahead (rank 0 reads at it=%d while the others run to %d): %s" % (
            n1, n2, "ok" if ok else "MISMATCH"), flush=True)
    return bool(ok)


def lost_neighbour_case(rank, world, local):
    """Ranks asked for different step counts: the surviving slab_wait gives up after
    IBLB_SLAB_TIMEOUT_S and the next getter reports it instead of a GPU that spins for ever."""
    os.environ["IBLB_SLAB_TIMEOUT_S"] = "2"
    X, Y = 128 * world, 128
    x0, w = slabs.slab_range(rank, world, X, 1)
    s = Solver(Params.channel(X, Y, slab_x0=x0, slab_w=w, device=local, gx=1e-6))
    slabs.connect_ring(s, dist, rank, world)
    s.step(4 if rank == 0 else 7)  # rank 0 never does steps 5..7
    try:
        s.fields()
        reported = False
    except Exception as e:
        reported = "timed out" in str(e)
    if rank == 0:
        ok = not reported                 # it did all it was asked for
    else:
        # its neighbour never arrived -- or, further away, a neighbour said that ITS neighbour never
        # arrived: a slab that gives up poisons its neighbours' flags, so every rank reports
        ok = reported
    dist.barrier()
    s.close()
    del os.environ["IBLB_SLAB_TIMEOUT_S"]
    allr = [None] * world
    dist.all_gather_object(allr, (rank, ok, reported))
    if rank == 0:
        print("case lost neighbour (timeout instead of a hang): %s %s" % (
            "ok" if all(x[1] for x in allr) else "FAILED", allr), flush=True)
    return bool(ok)


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
    ok &= run_ahead_case(rank, world, local)
    ok &= lost_neighbour_case(rank, world, local)
    allok = [None] * world
    dist.all_gather_object(allok, ok)
    ok = all(allok)
    if rank == 0:
        print("IPC SLABS OK" if ok else "IPC SLABS FAILED", flush=True)
    dist.barrier()
    dist.destroy_process_group()
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
