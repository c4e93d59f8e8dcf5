"""Host-side logic and the C-ABI surface, no GPU needed.

 * libiblb.so loads and exports every symbol include/iblb.h declares;
 * iblb_params_from_args / iblb_params_derive reproduce main.cu:284-321,336 (compared with the
   oracle's restatement of the same lines) including the reference's error behaviour;
 * slab partition / ring / sweep-script helpers;
 * the product package never imports the oracle.
"""
import ctypes
import os
import re

import numpy as np
import pytest

from checkers import oracle_derive
from cuda_iblb_11_b200 import SYMBOLS, IblbError, Params, lib
from cuda_iblb_11_b200 import slabs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    header = open(os.path.join(ROOT, "include", "iblb.h")).read()
    declared = set(re.findall(r"\b(iblb_[a-z_0-9]+)\s*\(", header))
    L = lib()
    assert declared, "no declarations parsed"
    for name in sorted(declared):
        assert hasattr(L, name), "libiblb.so does not export %s" % name
    assert set(SYMBOLS) == declared
    assert b"sm_100a" in L.iblb_version()


def test_params_struct_layout_matches_header():
    # the ctypes mirror and the C struct agree on size: default() writes the whole struct
    p = Params.default()
    assert (p.c_fraction, p.c_num, p.c_space, p.YDIM, p.LENGTH) == (1, 6, 48, 192, 96)
    assert p.device == -1 and p.flags == 0 and p.slab_w == 0


@pytest.mark.parametrize("line", [
    "1 6 48 1.0 1.0 5 1.0 100 0 0",
    "5 12 48 1.0 1.0 5 1.0 100 1 0",
    "2 6 48 0.5 2.5 4 0.3 10 0 1",
    "1 256 128 1.0 1.0 5 0.002 2 0 0",
    "3 7 50 2.0 9.99 3 1.5 7 0 0",
])
def test_ten_argument_line_matches_reference_derivation(line):
    w = line.split()
    p = Params.from_args(w)
    d = oracle_derive(int(w[0]), int(w[1]), int(w[2]), float(w[3]), float(w[4]), int(w[5]),
                      float(w[6]), int(w[7]))
    assert (p.XDIM, p.T, p.ITERATIONS, p.INTERVAL, p.p_step) == (
        d.XDIM, d.T, d.ITERATIONS, d.INTERVAL, d.p_step)
    for k in ("SPEED", "t_scale", "x_scale", "s_scale", "TAU", "TAU2", "Ma"):
        assert getattr(p, k) == getattr(d, k), k  # bit-for-bit
    assert p.YDIM == 192 and p.LENGTH == 96
    assert (p.ShARC, p.BigData) == (int(w[8]), int(w[9]))


def test_cilia6_reference_values():
    p = Params.from_args("1 6 48 1.0 1.0 5 1.0 100 0 0".split())
    assert (p.XDIM, p.T, p.ITERATIONS, p.INTERVAL, p.p_step) == (288, 100000, 100000, 1000, 16666)
    tau = 0.008 * 96 / (0.577 * 0.577) + 0.5
    assert abs(p.TAU - tau) < 1e-14 and abs(p.TAU2 - (1 / (12 * (tau - 0.5)) + 0.5)) < 1e-14
    assert abs(p.TAU - 2.806798) < 1e-6 and abs(p.TAU2 - 0.536125) < 1e-6
    assert abs(p.x_scale - 0.0625) < 1e-15 and abs(p.t_scale - 6.7e-4) < 1e-15


@pytest.mark.parametrize("script_line,expect", [
    ("1 6 1.0 100000 100000 100 1 0", "1 6 48 1.0 1.0 5 1.0 100 1 0"),     # app.sh:7
    ("11 12 1.0 100000 100000 100 1 0", "11 12 48 1.0 1.0 5 1.0 100 1 0"),  # cilia12.sh:13
    ("2 6 1.0 100000 50000 100 0 1", "2 6 48 1.0 1.0 5 0.5 100 0 1"),
])
def test_legacy_eight_argument_script_line(script_line, expect):
    a = Params.from_args(script_line.split())
    b = Params.from_args(expect.split())
    for name, _ in Params._fields_:
        assert getattr(a, name) == getattr(b, name), name


def test_too_few_arguments_message_and_code():
    # main.cu:284-289: "Too few arguments! <n> entered of 10 required. "
    with pytest.raises(IblbError) as e:
        Params.from_args("1 6 48 1.0".split())
    assert "libiblb error 3" in str(e.value)
    assert "Too few arguments! 4 entered of 10 required." in str(e.value)


def test_not_enough_cilia_message_and_code():
    # main.cu:303-308: XDIM = 3*48 = 144 < 2*LENGTH = 192
    with pytest.raises(IblbError) as e:
        Params.from_args("1 3 48 1.0 1.0 5 1.0 100 0 0".split())
    assert "libiblb error 4" in str(e.value)
    assert "not enough cilia in simulation! Cilia spacing of 48 requires at least 4 cilia" in str(e.value)


def test_create_without_gpu_fails_loudly_not_silently():
    """No CPU fallback: without a CUDA device iblb_create returns IBLB_ECUDA."""
    from conftest import has_gpu
    if has_gpu():
        pytest.skip("GPU present")
    p = Params.channel(64, 64)
    h = ctypes.c_void_p()
    rc = lib().iblb_create(ctypes.byref(p), ctypes.byref(h))
    assert rc == 2 and not h.value
    assert lib().iblb_last_error()


def test_create_rejects_bad_geometry_before_touching_cuda():
    for kw in (dict(XDIM=2, YDIM=64), dict(XDIM=64, YDIM=63)):
        p = Params.channel(**kw)
        h = ctypes.c_void_p()
        assert lib().iblb_create(ctypes.byref(p), ctypes.byref(h)) == 1
    p = Params.channel(288, 192, c_num=6)
    p.XDIM = 300  # with cilia XDIM must equal c_num*c_space (main.cu:298)
    h = ctypes.c_void_p()
    assert lib().iblb_create(ctypes.byref(p), ctypes.byref(h)) == 1


# ------------------------------------------------------------------------------ slabs / sweeps
def test_slab_ranges_tile_the_channel():
    for X, world, align in ((32768, 8, 128), (8192, 4, 1), (288, 2, 48), (100, 3, 1), (576, 8, 1)):
        xs = [slabs.slab_range(r, world, X, align) for r in range(world)]
        assert xs[0][0] == 0 and sum(w for _, w in xs) == X
        for (a, wa), (b, _) in zip(xs, xs[1:]):
            assert a + wa == b
        assert all(w % align == 0 and x0 % align == 0 for x0, w in xs)
        assert max(w for _, w in xs) - min(w for _, w in xs) <= align
    with pytest.raises(ValueError):
        slabs.slab_range(0, 8, 288, 48)  # 6 cilia cannot be split over 8 slabs at cilium granularity


def test_ring_neighbours():
    assert slabs.ring_neighbours(0, 1) == (0, 0)
    assert slabs.ring_neighbours(0, 8) == (7, 1)
    assert slabs.ring_neighbours(7, 8) == (6, 0)


CILIA6_SH = """#! /bin/bash -l

module load libs/CUDA/8.0.44/binary

module load dev/gcc/4.9.4

./app 1 6 1.0 100000 100000 100 1 0

./app 2 6 1.0 100000 100000 100 1 0

./app 3 6 1.0 100000 100000 100 1 0
"""


def test_sweep_script_parsing_and_assignment():
    runs = slabs.parse_sweep_script(CILIA6_SH)
    assert runs == [["1", "6", "1.0", "100000", "100000", "100", "1", "0"],
                    ["2", "6", "1.0", "100000", "100000", "100", "1", "0"],
                    ["3", "6", "1.0", "100000", "100000", "100", "1", "0"]]
    ps = [Params.from_args(r) for r in runs]
    assert [p.c_fraction for p in ps] == [1, 2, 3] and all(p.XDIM == 288 for p in ps)
    assert slabs.assign_runs(6, 8) == [[0], [1], [2], [3], [4], [5], [], []]
    assert slabs.assign_runs(6, 4) == [[0, 4], [1, 5], [2], [3]]


def test_c_driver_partition_equals_the_python_one():
    """iblb_slab_range (the x-slab partition of iblb_run_gpus / IBLB_GPUS) against
    cuda_iblb_11_b200.slabs.slab_range, which bench.py and the IPC tests use."""
    import ctypes
    L = lib()
    a, b = ctypes.c_uint(), ctypes.c_uint()
    for X, n, align in ((32768, 8, 128), (288, 2, 48), (288, 4, 48), (576, 5, 48), (102, 4, 1),
                        (8192, 3, 1), (4096, 8, 128)):
        for r in range(n):
            assert L.iblb_slab_range(r, n, X, align, ctypes.byref(a), ctypes.byref(b)) == 0
            assert (a.value, b.value) == slabs.slab_range(r, n, X, align)
    assert L.iblb_slab_range(0, 7, 288, 48, ctypes.byref(a), ctypes.byref(b)) != 0  # 6 units, 7 slabs
    assert L.iblb_slab_range(0, 2, 100, 48, ctypes.byref(a), ctypes.byref(b)) != 0  # not whole units
    with pytest.raises(ValueError):
        slabs.slab_range(0, 7, 288, 48)


def test_sweep_launcher_plan_names_and_summary(tmp_path, capsys):
    """cuda_iblb_11_b200.sweep: assignment of a script's lines to GPUs, flux-file names
    (main.cu:612), flux-curve reader and the summary table; no GPU involved."""
    from cuda_iblb_11_b200 import sweep
    runs, order = sweep.plan(CILIA6_SH, 2)
    assert len(runs) == 3
    assert [(g, i) for g, i, _ in order] == [(0, 0), (0, 2), (1, 1)]
    p = Params.from_args(runs[1])
    assert sweep.flux_file_name(p) == "2_6_48_1_1x5-flux.dat"
    p = Params.from_args("6 12 48 0.5 2.5 4 1.0 100 0 0".split())
    assert sweep.flux_file_name(p) == "6_12_48_0.5_2.5x4-flux.dat"
    f = tmp_path / "x-flux.dat"
    f.write_text("0\t0\n0.0134\t1.25e-05\n\n0.0268\t2.5e-05\n")
    assert sweep.read_flux(str(f)) == [(0.0, 0.0), (0.0134, 1.25e-05), (0.0268, 2.5e-05)]
    rows = [dict(run=0, gpu=1, args="1 6", error=None, final_t=0.02, final_Q=1e-5, points=3,
                 seconds=1.5, mlups=10.0, nonfinite_from_t=None),
            dict(run=1, gpu=None, args="2 6", error="worker died", final_t=None, final_Q=None,
                 points=0, seconds=None, mlups=None, nonfinite_from_t=0.0268)]
    out = tmp_path / "Flux" / "sweep-summary.tsv"
    sweep.write_summary(rows, str(out))
    lines = out.read_text().splitlines()
    assert lines[0].split("\t") == ["run", "gpu", "args", "final_t", "final_Q", "points", "seconds",
                                    "MLUPS", "nonfinite_from_t", "error"]
    assert lines[1].split("\t") == ["0", "1", "1 6", "0.02", "1e-05", "3", "1.5", "10.0", "", ""]
    assert lines[2].split("\t")[8:] == ["0.0268", "worker died"]
    script = tmp_path / "s.sh"
    script.write_text(CILIA6_SH)
    assert sweep.main([str(script), "--gpus", "8", "--dry-run"]) == 0
    assert "gpu 2: run 2: ./app 3 6 1.0 100000 100000 100 1 0" in capsys.readouterr().out


def test_product_package_never_touches_the_oracle():
    pkg = os.path.join(ROOT, "cuda_iblb_11_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle/" not in text.replace("the oracle", "") or f == "__init__.py", f
                assert "libiblb_oracle" not in text and "libiblb_ref" not in text, f
                assert "checkers" not in text, f
