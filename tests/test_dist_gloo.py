"""World-size-2 gloo test (CPU) of the multi-GPU host plumbing: IPC-handle exchange, ring wiring
and sweep assignment, with a stand-in solver object (no GPU here; the real peer stores are
exercised by tests/test_gpu_slabs_fullsize.py on the box)."""
import ctypes
import os
import socket

import torch.distributed as dist
import torch.multiprocessing as mp

from cuda_iblb_11_b200 import HaloHandle, slabs


class FakeSlab:
    """mimics Solver.export_halo / connect; records who it was wired to"""

    def __init__(self, rank, x0, w):
        self.rank, self.x0, self.w = rank, x0, w
        self.left = self.right = None

    def export_halo(self):
        h = HaloHandle()
        h.device, h.pid, h.width, h.ghost, h.ld = self.rank, 1000 + self.rank, self.w, 1, 64
        h.plane = (self.w + 2) * 64
        h.mem[0][0] = 0xA0 + self.rank
        return h

    def connect(self, left, right):
        self.left, self.right = left, right


def _worker(rank, world, port, X, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    x0, w = slabs.slab_range(rank, world, X, 48)
    s = FakeSlab(rank, x0, w)
    slabs.connect_ring(s, dist, rank, world)
    runs = slabs.assign_runs(6, world)[rank]
    q.put((rank, x0, w, s.left.device, s.right.device, s.left.mem[0][0], s.right.width, runs))
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_ring_connect_over_gloo_world_size_2():
    world, X = 2, 288
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, world, port, X, q)) for r in range(world)]
    for p in ps:
        p.start()
    out = sorted(q.get(timeout=120) for _ in range(world))
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    (r0, x00, w0, l0, rr0, m0, rw0, runs0), (r1, x01, w1, l1, rr1, m1, rw1, runs1) = out
    assert (x00, w0, x01, w1) == (0, 144, 144, 144)
    # two slabs: each one's left AND right neighbour is the other one
    assert (l0, rr0, l1, rr1) == (1, 1, 0, 0)
    assert m0 == 0xA1 and m1 == 0xA0 and rw0 == 144 and rw1 == 144
    assert runs0 == [0, 2, 4] and runs1 == [1, 3, 5]


def test_handle_is_plain_bytes():
    h = HaloHandle()
    assert ctypes.sizeof(h) == len(bytes(h)) == 240  # sizeof(iblb_halo_handle), include/iblb.h
    assert bytes(HaloHandle.from_buffer_copy(bytes(h))) == bytes(h)
