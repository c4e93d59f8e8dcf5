"""x-slab coupling, full-size (BASELINE.json configs[1]) properties, and the driver-level run.
Needs a B200: `pytest -m gpu`."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

from checkers import (REF_FLAG_PICK_LOWEST_K, Checker, make_params, reference_tau, rel_l2)
from cuda_iblb_11_b200 import FLAG_HALF_FORCE_BODY, Params, Solver, lib, run
from cuda_iblb_11_b200 import slabs

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def n_gpus():
    cudart = ctypes.CDLL("libcudart.so")
    n = ctypes.c_int(0)
    cudart.cudaGetDeviceCount(ctypes.byref(n))
    return n.value


def gather_slabs(parts, X, Y):
    """slab-local (rho,u) in reference layout -> global arrays"""
    rho = np.empty((Y, X))
    ux = np.empty((Y, X))
    uy = np.empty((Y, X))
    for (x0, w, s) in parts:
        r, u = s.fields()
        rho[:, x0:x0 + w] = r.reshape(Y, w)
        ux[:, x0:x0 + w] = u[:w * Y].reshape(Y, w)
        uy[:, x0:x0 + w] = u[w * Y:].reshape(Y, w)
    return rho.ravel(), np.concatenate([ux.ravel(), uy.ravel()])


def make_local_ring(X, Y, n, devices, **kw):
    align = kw.get("c_space", 48) if kw.get("c_num", 0) else 1
    parts = []
    for r in range(n):
        x0, w = slabs.slab_range(r, n, X, align)
        parts.append((x0, w, Solver(Params.channel(X, Y, slab_x0=x0, slab_w=w,
                                                   device=devices[r % len(devices)], **kw))))
    for r in range(n):
        l, rr = slabs.ring_neighbours(r, n)
        parts[r][2].connect_local(parts[l][2], parts[rr][2])
    return parts


@pytest.mark.parametrize("n_slabs,X,Y,kw", [
    (2, 128, 64, dict(gx=1e-6, u0_amp=1e-3)),
    (3, 96, 128, dict(gx=0.0, u0_amp=4e-3)),
    (4, 102, 40, dict(gx=1e-6, u0_amp=1e-3)),           # generic path, uneven widths
    (2, 288, 192, dict(c_num=6, c_fraction=6)),          # cilia straddle the slab seam
    (3, 288, 192, dict(c_num=6, c_fraction=1)),
    (4, 576, 192, dict(c_num=12, c_fraction=12)),
])
def test_slabs_on_one_gpu_are_bit_identical_to_the_whole_channel(n_slabs, X, Y, kw):
    whole = Solver(Params.channel(X, Y, **kw))
    parts = make_local_ring(X, Y, n_slabs, [0], **kw)
    done = 0
    for n in (1, 2, 9, 30):
        for _ in range(n - done):
            for (_, _, s) in parts:  # lockstep
                s.step(1)
        whole.step(n - done)
        done = n
        rho, u = gather_slabs(parts, X, Y)
        r2, u2 = whole.fields()
        assert np.array_equal(rho, r2), "rho it=%d" % n
        assert np.array_equal(u, u2), "u it=%d" % n
    if kw.get("c_num"):
        q = sum(s.Q for (_, _, s) in parts)  # the flux plane lives in exactly one slab
        assert q == whole.Q
    for (_, _, s) in parts:
        s.close()
    whole.close()


@pytest.mark.skipif("n_gpus() < 2")
def test_slabs_across_two_gpus_same_process():
    X, Y, kw = 256, 256, dict(gx=1e-6, u0_amp=1e-3)
    whole = Solver(Params.channel(X, Y, device=0, **kw))
    parts = make_local_ring(X, Y, 2, [0, 1], **kw)
    for _ in range(20):
        for (_, _, s) in parts:
            s.step(1)
    whole.step(20)
    rho, u = gather_slabs(parts, X, Y)
    r2, u2 = whole.fields()
    assert np.array_equal(rho, r2) and np.array_equal(u, u2)


@pytest.mark.skipif("n_gpus() < 2")
def test_slabs_across_processes_ipc():
    """one process per GPU, CUDA IPC peer stores + device flags (the bench's N>1 path)"""
    cmd = ["python", "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", "29671",
           os.path.join(ROOT, "tools", "slab_ipc_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "IPC SLABS OK" in r.stdout


# ------------------------------------------------------------------------------ full size (C2)
def test_c2_full_size_equals_narrow_oracle_channel():
    """configs[1]: 8192x8192 forced channel.  The problem is x-invariant, so every column must
    equal the oracle's 4x8192 channel — the full-size run checked exactly, at oracle cost 4/8192."""
    Y, gx, u0, steps = 8192, 1e-7, 1e-3, 30
    s = Solver(Params.channel(8192, Y, gx=gx, u0_amp=u0))
    o = Checker("oracle", make_params(4, Y, gx=gx, u0_amp=u0, threads=4))
    s.step(steps)
    o.step(steps)
    rho, u = s.fields()
    ro, uo = o.fields()
    rho = rho.reshape(Y, 8192)
    ux = u[:8192 * Y].reshape(Y, 8192)
    uy = u[8192 * Y:].reshape(Y, 8192)
    # x-invariance is exact (same arithmetic in every column)
    assert np.array_equal(rho, np.broadcast_to(rho[:, :1], rho.shape))
    assert np.array_equal(ux, np.broadcast_to(ux[:, :1], ux.shape))
    assert rel_l2(rho[:, 0], ro.reshape(Y, 4)[:, 0]) <= 1e-10
    assert rel_l2(ux[:, 0], uo[:4 * Y].reshape(Y, 4)[:, 0]) <= 1e-10
    assert np.abs(uy).max() <= 1e-10 * np.abs(ux).max() + np.abs(uo[4 * Y:]).max() * 2
    s.close()


def test_c2_full_size_mass_conservation_and_start_up_flow():
    """no forcing: total mass conserved to the C_S-truncation drift (~1e-12/step); uniform forcing
    from rest: away from the walls u_x grows by exactly the momentum the reference's scheme
    injects per step, (1-1/2TAU) + 1/(2 TAU2) times g (TRT odd moments relax with 1/TAU2 while
    the Guo prefactor uses TAU: LatticeBoltzmann.cu:56,80-81)."""
    Y = 8192
    s = Solver(Params.channel(8192, Y, u0_amp=2e-3))
    m0 = s.fields()[0].sum()
    s.step(20)
    m1 = s.fields()[0].sum()
    assert abs(m1 - m0) <= 1e-10 * m0
    s.close()
    g, n = 1e-7, 40
    s = Solver(Params.channel(8192, Y, gx=g))
    s.step(n)
    rho, u = s.fields()
    ux = u[:8192 * Y].reshape(Y, 8192)[Y // 2]
    tau, tau2 = reference_tau()
    # the core of the channel never feels the walls in 40 steps: plug flow, closed form below
    o = Checker("oracle", make_params(4, 256, gx=g, threads=1))
    o.step(n)
    core = o.fields()[1][:4 * 256].reshape(256, 4)[128, 0]
    # u = 3e-6 is a difference of O(0.1) populations: 40 steps of 1e-17 absolute rounding
    # (FMA vs none) are 1e-10 relative to it; the absolute error is what is bounded
    assert np.abs(ux - core).max() <= 1e-14
    assert np.abs(ux / core - 1).max() <= 1e-9
    assert core > 0.5 * n * g
    s.close()


# ------------------------------------------------------------------------------ driver-level run
def test_run_writes_the_reference_data_tree(tmp_path):
    """iblb_run == the reference's main(): SimLog.txt, flux file, fluid/cilia snapshots
    (main.cu:589-631, 758-790, 936-1005, 1030-1060), values checked against the oracle through the
    reference's 6-significant-digit text format."""
    p = Params.from_args("6 6 48 1.0 1.0 5 1.0 100 0 1".split())  # BigData on
    p.ITERATIONS, p.INTERVAL, p.device = 41, 20, 0
    root = str(tmp_path / "Data")
    run(p, root)
    raw = os.path.join(root, "Raw", "6", "6")
    cil = os.path.join(root, "Cilia", "6", "6")
    flux = os.path.join(root, "Flux", "6_6_48_1_1x5-flux.dat")
    assert sorted(os.listdir(raw)) == ["0-fluid.dat", "20-fluid.dat", "40-fluid.dat", "SimLog.txt"]
    assert sorted(os.listdir(cil)) == ["0-cilia.dat", "20-cilia.dat", "40-cilia.dat"]
    log = open(os.path.join(raw, "SimLog.txt")).read()
    for needle in ("Size: 288x192", "Iterations: 41", "Reynolds Number: 1",
                   "Relaxation times: 2.8068, 0.536125", "Spatial step: 6.25e-08m",
                   "Time step: 6.7e-07s", "Mach number: 0.0138648", "Phase Step: 6/6",
                   "Big Data is ON", "Running on local GPU", "Total runtime: "):
        assert needle in log, needle
    o = Checker("oracle", make_params(288, 192, c_num=6, c_fraction=6,
                                      flags=REF_FLAG_PICK_LOWEST_K, threads=8))
    rows = [l.split("\t") for l in open(flux).read().splitlines()]
    assert [r[0] for r in rows] == ["0", "0.0134", "0.0268", "0.02747"]  # it*t_scale, %g
    qs = []
    for n in (1, 21, 41):
        o.step(n - o.it)
        qs.append(o.Q)
    for r, q in zip(rows, qs + [qs[-1]]):
        assert float(r[1]) == pytest.approx(q * 0.0625, rel=2e-5, abs=1e-12)
    # fluid snapshot of iteration 40: one line per node, x fastest, blank line after each row
    lines = open(os.path.join(raw, "40-fluid.dat")).read().split("\n")
    data = [l for l in lines if l]
    assert len(data) == 288 * 192 and len(lines) == 288 * 192 + 192 + 1
    rho, u = o.fields()
    j = 37 * 288 + 101
    x, y, ux, uy, ab, r = [float(v) for v in data[j].split("\t")]
    s_scale = p.s_scale
    assert x == pytest.approx(101 * 0.0625) and y == pytest.approx(37 * 0.0625)
    assert ux == pytest.approx(u[j] * s_scale, rel=2e-5, abs=1e-9)
    assert uy == pytest.approx(u[288 * 192 + j] * s_scale, rel=2e-5, abs=1e-9)
    assert r == pytest.approx(rho[j], rel=2e-6)
    cl = open(os.path.join(cil, "40-cilia.dat")).read().split("\n")
    cdata = [l for l in cl if l]
    assert len(cdata) == 576
    sm, us, _, eps = o.markers()
    k = 200
    v = [float(t) for t in cdata[k].split("\t")]
    assert v[0] == pytest.approx(sm[2 * k] * 0.0625, rel=2e-5)
    assert v[1] == pytest.approx(sm[2 * k + 1] * 0.0625, rel=2e-5)
    assert v[2] == pytest.approx(us[2 * k] * s_scale, rel=2e-5, abs=1e-9)
    assert int(v[4]) == eps[k]


def test_run_restarts_from_a_checkpoint(tmp_path, monkeypatch):
    """Driver level: IBLB_CHECKPOINT / IBLB_RESTART around iblb_run.  A 41-iteration run that
    restarts from the checkpoint a 21-iteration run left writes the same snapshots and the same
    final flux as the uninterrupted 41-iteration run."""
    def params(iterations):
        p = Params.from_args("6 6 48 1.0 1.0 5 1.0 100 0 1".split())
        p.ITERATIONS, p.INTERVAL, p.device = iterations, 20, 0
        return p
    full, part = str(tmp_path / "full"), str(tmp_path / "part")
    run(params(41), full)
    ck = str(tmp_path / "run.ckp")
    monkeypatch.setenv("IBLB_CHECKPOINT", ck)
    run(params(21), part)
    monkeypatch.delenv("IBLB_CHECKPOINT")
    monkeypatch.setenv("IBLB_RESTART", ck)
    run(params(41), part)
    monkeypatch.delenv("IBLB_RESTART")
    for sub, name in (("Raw/6/6", "40-fluid.dat"), ("Cilia/6/6", "40-cilia.dat"),
                      ("Raw/6/6", "20-fluid.dat")):
        assert open(os.path.join(full, sub, name)).read() == open(os.path.join(part, sub, name)).read()
    fa = open(os.path.join(full, "Flux", "6_6_48_1_1x5-flux.dat")).read().splitlines()
    fb = open(os.path.join(part, "Flux", "6_6_48_1_1x5-flux.dat")).read().splitlines()
    assert fb[:2] == fa[:2] and fb[-2:] == fa[-2:]  # it 0, 20 ... 40, end-of-run line



@pytest.mark.parametrize("n_slabs", [2, 3])
def test_run_on_several_gpus_writes_the_same_tree(tmp_path, n_slabs):
    """iblb_run_gpus / `IBLB_GPUS=N ./app ...`: the C++ driver cuts the channel into N x-slabs on
    the GPUs of this process (wrapping onto the GPUs there are), couples them with
    iblb_connect_local, steps them in lockstep and gathers the snapshots: every file of the Data/
    tree is byte-identical to the one-GPU run's (SimLog aside: timings).  Checkpoints are one file
    per slab and restart the same way."""
    def params(iterations):
        p = Params.from_args("6 6 48 1.0 1.0 5 1.0 100 0 1".split())
        p.ITERATIONS, p.INTERVAL, p.device = iterations, 20, 0
        return p
    one, many, part = str(tmp_path / "one"), str(tmp_path / "many"), str(tmp_path / "part")
    run(params(41), one, gpus=1)
    run(params(41), many, gpus=n_slabs)
    names = [("Raw/6/6", "0-fluid.dat"), ("Raw/6/6", "20-fluid.dat"), ("Raw/6/6", "40-fluid.dat"),
             ("Cilia/6/6", "20-cilia.dat"), ("Cilia/6/6", "40-cilia.dat"),
             ("Flux", "6_6_48_1_1x5-flux.dat")]
    for sub, name in names:
        assert open(os.path.join(one, sub, name)).read() == open(os.path.join(many, sub, name)).read(), name
    assert "GPUs: %d x-slabs" % n_slabs in open(os.path.join(many, "Raw/6/6", "SimLog.txt")).read()
    # checkpoint after 21 iterations on N slabs, restart on N slabs, same final files
    ck = str(tmp_path / "run.ckp")
    os.environ["IBLB_CHECKPOINT"] = ck
    try:
        run(params(21), part, gpus=n_slabs)
    finally:
        del os.environ["IBLB_CHECKPOINT"]
    assert all(os.path.exists("%s.%d" % (ck, r)) for r in range(n_slabs))
    os.environ["IBLB_RESTART"] = ck
    try:
        run(params(41), part, gpus=n_slabs)
    finally:
        del os.environ["IBLB_RESTART"]
    for sub, name in (("Raw/6/6", "40-fluid.dat"), ("Cilia/6/6", "40-cilia.dat")):
        assert open(os.path.join(one, sub, name)).read() == open(os.path.join(part, sub, name)).read()
    # A run killed while it renames its per-slab files leaves files of two iterations side by side:
    # such a set must be refused, not stepped.  Build one from two honest sets (iteration 20 / 40).
    import shutil
    from cuda_iblb_11_b200 import IblbError
    assert not [f for f in os.listdir(str(tmp_path)) if f.endswith(".new")]
    late = str(tmp_path / "late.ckp")
    os.environ["IBLB_CHECKPOINT"] = late
    try:
        run(params(41), str(tmp_path / "late"), gpus=n_slabs)
    finally:
        del os.environ["IBLB_CHECKPOINT"]
    shutil.copy("%s.%d" % (late, n_slabs - 1), "%s.%d" % (ck, n_slabs - 1))
    os.environ["IBLB_RESTART"] = ck
    try:
        with pytest.raises(IblbError) as err:
            run(params(41), str(tmp_path / "mixed"), gpus=n_slabs)
    finally:
        del os.environ["IBLB_RESTART"]
    assert "not one set" in str(err.value)


def test_app_runs_on_several_gpus_from_the_environment(tmp_path):
    app = os.path.join(ROOT, "app")
    if not os.path.exists(app):
        pytest.skip("app not built")
    # 30 iterations with P_num = 100: INTERVAL = 30/100 = 0 (main.cu:301), so there is no periodic
    # output and the flux file is the end-of-run line alone (main.cu:1030-1034); the reference
    # itself cannot run this line (`it % INTERVAL`, main.cu:938, divides by zero).  200 iterations:
    # INTERVAL = 2, a line at it = 0, 2, ... 198 (main.cu:996-1004) and the end-of-run line.
    for iterations, lines in (("30", 1), ("200", 101)):
        outs = []
        for gpus in ("1", "2"):
            root = tmp_path / ("D%s_%s" % (gpus, iterations))
            env = dict(os.environ, IBLB_ROOT=str(root), IBLB_ITERATIONS=iterations, IBLB_GPUS=gpus)
            r = subprocess.run([app] + "6 6 48 1.0 1.0 5 1.0 100 0 0".split(), capture_output=True,
                               text=True, env=env)
            assert r.returncode == 0, r.stderr
            outs.append(open(str(root / "Flux" / "6_6_48_1_1x5-flux.dat")).read())
        assert outs[0] == outs[1] and len(outs[0].splitlines()) == lines
        if lines > 1:
            assert outs[0].splitlines()[0].split() == ["0", "0"]  # it = 0: the fluid is still at rest


def test_app_executable_surface(tmp_path):
    app = os.path.join(ROOT, "app")
    if not os.path.exists(app):
        pytest.skip("app not built")
    r = subprocess.run([app, "1", "6"], capture_output=True, text=True)
    assert r.returncode == 1 and "Too few arguments! 2 entered of 10 required." in r.stdout
    r = subprocess.run([app] + "1 3 48 1.0 1.0 5 1.0 100 0 0".split(), capture_output=True, text=True)
    assert r.returncode == 1 and "not enough cilia in simulation!" in r.stdout
    env = dict(os.environ, IBLB_ROOT=str(tmp_path / "D"), IBLB_ITERATIONS="30")
    r = subprocess.run([app] + "1 6 1.0 100000 100000 100 0 0".split(), capture_output=True,
                       text=True, env=env)  # legacy 8-value script line (app.sh:7)
    assert r.returncode == 0, r.stderr
    assert "Running Simulation..." in r.stdout
    assert os.path.exists(str(tmp_path / "D" / "Flux" / "1_6_48_1_1x5-flux.dat"))


def test_sweep_launcher_runs_a_script_and_gathers_the_flux_curves(tmp_path):
    """cuda_iblb_11_b200.sweep (config C5): the `./app` lines of a multiapp.sh-style script run as
    independent replicas (worker process per GPU) and give the flux of stand-alone runs."""
    from cuda_iblb_11_b200 import sweep
    script = ("#! /bin/bash -l\nmodule load libs/CUDA/8.0.44/binary\n\n"
              "./app 6 6 1.0 100000 100000 100 0 0\n\n./app 3 6 1.0 100000 100000 100 0 0\n")
    root = str(tmp_path / "Data")
    rows = sweep.run_sweep(script, n_gpus=1, root=root, iterations=41, interval=20)
    assert [r["error"] for r in rows] == [None, None]
    assert [r["points"] for r in rows] == [4, 4] and all(r["gpu"] == 0 for r in rows)
    for r, cf in zip(rows, (6, 3)):
        s = Solver(Params.channel(288, 192, c_num=6, c_fraction=cf))
        s.step(41)
        assert r["final_Q"] == pytest.approx(s.Q * 0.0625, rel=2e-5, abs=1e-12)
        s.close()
    table = open(os.path.join(root, "Flux", "sweep-summary.tsv")).read().splitlines()
    assert len(table) == 3 and table[1].split("\t")[2] == "6 6 1.0 100000 100000 100 0 0"

