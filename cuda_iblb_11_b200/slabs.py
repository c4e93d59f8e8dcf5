"""Host-side plumbing for multi-GPU runs: x-slab partition, ring neighbours, IPC-handle exchange
over torch.distributed (NCCL on the GPU box, gloo in CPU tests), and the one-run-per-GPU
assignment for parameter sweeps (multiapp.sh / cilia6.sh / cilia12.sh as an ensemble).

No data-path collective: after `connect_ring` the fused LB kernel stores boundary columns into
the neighbours' ghost columns itself (include/iblb.h, iblb_connect)."""
import re


def slab_range(rank, world, XDIM, align=1):
    """Columns [x0, x0+w) owned by `rank`: equal split, remainder to the low ranks, widths kept
    multiples of `align` (e.g. c_space so that cilia bases never straddle) when possible."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    units = XDIM // align
    if units * align != XDIM or units < world:
        raise ValueError("XDIM=%d cannot be split in %d slabs of multiples of %d" % (XDIM, world, align))
    base, rem = divmod(units, world)
    w = (base + (1 if rank < rem else 0)) * align
    x0 = (rank * base + min(rank, rem)) * align
    return x0, w


def ring_neighbours(rank, world):
    """(left, right) ranks on the periodic ring; a single rank is its own neighbour."""
    return (rank - 1) % world, (rank + 1) % world


def exchange_handles(dist, handle_bytes, world):
    """all_gather of the opaque iblb_halo_handle bytes; returns the list indexed by rank."""
    if dist is None or world == 1:
        return [handle_bytes]
    out = [None] * world
    dist.all_gather_object(out, handle_bytes)
    return out


def connect_ring(solver, dist, rank, world):
    """Export this slab's handle, gather everybody's, connect to the ring neighbours."""
    from .solver import HaloHandle
    mine = bytes(solver.export_halo())
    hs = [HaloHandle.from_buffer_copy(b) for b in exchange_handles(dist, mine, world)]
    left, right = ring_neighbours(rank, world)
    solver.connect(hs[left], hs[right])
    if dist is not None and world > 1:
        dist.barrier()


_APP_LINE = re.compile(r"^\s*\./app\s+(.*\S)\s*$")


def parse_sweep_script(text):
    """The `./app ...` argument lines of a reference batch script (cilia6.sh:7-17 etc.);
    `module load` and other site-specific lines are ignored."""
    runs = []
    for line in text.splitlines():
        m = _APP_LINE.match(line)
        if m:
            runs.append(m.group(1).split())
    return runs


def assign_runs(n_runs, world):
    """Run i goes to GPU i mod world; returns, per rank, the list of run indices in order."""
    return [[i for i in range(n_runs) if i % world == r] for r in range(world)]
