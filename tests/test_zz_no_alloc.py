"""iblb_get_fields_static: the field getter for host programs that cannot allocate device memory.

iblb_get_fields reserves a 128 MB device staging buffer with cudaMalloc the first time it is called.
A host program that gave the whole GPU to its own arena at start-up cannot afford that (the late
cudaMalloc fails somewhere in the middle of its run), and neither can one in which another host
thread keeps a stream capture open in the global capture mode (the runtime then refuses cudaMalloc
from every thread).  iblb_get_fields_static stages through a fixed 6 MiB array that is part of the
library image instead, for grids up to IBLB_STATIC_FIELDS_MAX_NODES nodes.

These tests call it through the C ABI (ctypes, cuda_iblb_11_b200/solver.py) and hold it to the
existing getter bit for bit -- both run the same kernel on the same populations, so the tolerance is
zero.  All calls share the one array, hence the back-to-back calls on different grids and values: a
value left behind by the previous call would show up as a mismatch.

The file name makes pytest collect it last (`pytest --collect-only`): under `-x` a failure here can
hide no other test.  Needs a B200: `pytest -m gpu`.
"""
import ctypes

import numpy as np
import pytest

from cuda_iblb_11_b200 import IblbError, Params, Solver

pytestmark = pytest.mark.gpu

MAX_NODES = 262144  # IBLB_STATIC_FIELDS_MAX_NODES, include/iblb.h


def free_device_bytes():
    cudart = ctypes.CDLL("libcudart.so")
    free, total = ctypes.c_size_t(0), ctypes.c_size_t(0)
    assert cudart.cudaMemGetInfo(ctypes.byref(free), ctypes.byref(total)) == 0
    return free.value


def same_as_allocating_getter(s, what):
    rho_s, u_s = s.fields_static()
    rho, u = s.fields()
    assert np.isfinite(rho).all() and np.isfinite(u).all(), what
    assert np.array_equal(rho_s, rho), "rho, " + what
    assert np.array_equal(u_s, u), "u, " + what
    return rho, u


def test_static_getter_equals_the_allocating_getter_on_calls_in_a_row():
    # C3 (576x192, 12 cilia, fused single-launch path: the getter has to rebuild the force band) and
    # a small pure-LB channel with quite different values: the second and third call find the
    # first one's data in the array
    big = Solver(Params.channel(576, 192, c_num=12, c_fraction=12))
    small = Solver(Params.channel(128, 64, gx=1e-6, u0_amp=1e-3))
    big.step(40)
    small.step(7)
    # the static getter goes first: at this point neither solver has a staging buffer of its own
    rho_b, u_b = big.fields_static()
    rho_s, u_s = small.fields_static()
    rho_b2, u_b2 = big.fields_static()
    assert np.array_equal(rho_b, rho_b2) and np.array_equal(u_b, u_b2)
    r, u = big.fields()
    assert np.array_equal(rho_b, r) and np.array_equal(u_b, u)
    r, u = small.fields()
    assert np.array_equal(rho_s, r) and np.array_equal(u_s, u)
    assert np.abs(u_b).max() > 0 and np.abs(u_s).max() > 0 and not np.array_equal(u_b[:u_s.size], u_s)
    # and again after more steps, the two getters alternating
    for n in (1, 32):
        big.step(n)
        small.step(n)
        same_as_allocating_getter(big, "C3 after +%d" % n)
        same_as_allocating_getter(small, "channel after +%d" % n)
    big.close()
    small.close()


def test_static_getter_before_the_first_step_and_with_one_output_only():
    s = Solver(Params.channel(288, 192, c_num=6, c_fraction=1))  # C1
    same_as_allocating_getter(s, "C1 at iteration 0")
    s.step(5)
    rho, u = same_as_allocating_getter(s, "C1 after 5")
    from cuda_iblb_11_b200 import lib
    only_rho = np.full(s.size, np.nan)
    only_u = np.full(2 * s.size, np.nan)
    assert lib().iblb_get_fields_static(s._h, only_rho.ctypes.data, None) == 0
    assert lib().iblb_get_fields_static(s._h, None, only_u.ctypes.data) == 0
    assert np.array_equal(only_rho, rho) and np.array_equal(only_u, u)
    assert lib().iblb_get_fields_static(s._h, None, None) != 0
    s.close()


def test_static_getter_takes_no_device_memory():
    s = Solver(Params.channel(576, 192, c_num=12, c_fraction=12))
    s.step(3)
    s.fields_static()  # the first call on a device looks up the array's address
    before = free_device_bytes()
    for _ in range(3):
        s.step(2)
        s.fields_static()
    assert free_device_bytes() == before
    s.close()


def test_static_getter_refuses_grids_beyond_its_array():
    # 2048 x 192 = 393 216 nodes > 262 144: a clear error, and the solver stays usable
    s = Solver(Params.channel(2048, 192, gx=1e-6))
    assert s.size > MAX_NODES
    s.step(2)
    with pytest.raises(IblbError) as err:
        s.fields_static()
    assert "exceed" in str(err.value) and str(MAX_NODES) in str(err.value)
    rho, u = s.fields()
    assert np.isfinite(rho).all() and np.isfinite(u).all()
    s.close()
    # the largest grid that fits: exactly MAX_NODES nodes
    s = Solver(Params.channel(2048, 128, gx=1e-6, u0_amp=1e-3))
    assert s.size == MAX_NODES
    s.step(3)
    same_as_allocating_getter(s, "2048x128")
    s.close()
