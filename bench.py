#!/usr/bin/env python
"""bench.py — MLUPS of the IB-LBM time-step path on B200 (BASELINE.json metric).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c4|c1|c3] [--impl reference]

N = 1 : configs[1] of BASELINE.json — pure-LB channel 8192x8192, uniform body force, FP64.
N > 1 : launched by torch.distributed.run, one rank per GPU.  The channel is N*8192 wide and 8192
        high, x-slab partitioned (8192 columns per GPU, weak scaling); boundary columns are stored
        straight into the neighbours' ghost columns over NVLink by the fused kernel (CUDA IPC peer
        mappings), steps ordered by device-side flags; torch.distributed only exchanges the IPC
        handles and provides the barrier / max-over-ranks.
A "step" is one lattice update of the whole channel.  value = nodes * K / (max over ranks of the
device time of the K timed steps, CUDA events on the solver's stream) / 1e6.
`--impl reference` runs the reference's own kernels (nvcc sm_100a build of the unmodified sources,
oracle/_ref/libiblb_ref_cuda.so — the reference has no CPU path) on the same configuration.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BYTES_PER_LUP = 144  # 9 populations x 8 B x (1 read + 1 write), SURVEY.md §8(d)

WORKLOADS = {
    # name: (XDIM per GPU or total, YDIM, c_num total or per gpu, c_space, scaling)
    "c2": dict(desc="pure-LB channel 8192x8192 per GPU, body force gx=1e-7, no IB (configs[1])",
               X=8192, Y=8192, c_num=0, c_space=48, scaling="weak", gx=1e-7),
    "c4": dict(desc="32768x32768 channel, 256-cilia carpet, slab-partitioned (configs[3])",
               X=32768, Y=32768, c_num=256, c_space=128, scaling="strong", gx=0.0),
    # the reference's coupling overflows within ~40 steps for phase-lagged cilia, so
    # the timed cilia cases use the in-phase line of the sweep (cilia6.sh:17); work per step is
    # the same for every c_fraction
    "c1": dict(desc="cilia6.sh case 288x192, 6 cilia, c_fraction 6 (configs[0])",
               X=288, Y=192, c_num=6, c_space=48, scaling="replicas", gx=0.0, c_fraction=6),
    "c3": dict(desc="cilia12.sh grid 576x192, 12 cilia, c_fraction 12 (configs[2])",
               X=576, Y=192, c_num=12, c_space=48, scaling="replicas", gx=0.0, c_fraction=12),
}


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic():
    """DRAM bytes per launch of the dominant kernel from the committed ncu capture, or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "roofline_traffic.json")) as f:
            return json.load(f)
    except Exception:
        return None


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled DURING the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "100", "-i", str(self.gpu)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        rows = [l for (t, l) in self.lines if t0 - 0.05 <= t <= t1 + 0.15] or \
               [l for (_, l) in self.lines]
        for l in rows:
            c = [x.strip() for x in l.split(",")]
            try:
                sm.append(float(c[1]))
                mx = float(c[2])
                for n, v in zip(names, c[4:8]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx,
                "reasons": sorted(reasons), "samples": len(sm)}


def dist_env():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    return world, rank, local


class Plumbing:
    """torch.distributed (NCCL) for barrier / max / handle exchange; a no-op at world size 1."""

    def __init__(self, n_gpus):
        self.world, self.rank, self.local = dist_env()
        self.dist = None
        self.torch = None
        if self.world > 1:
            import torch
            import torch.distributed as dist
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            torch.cuda.set_device(self.local)
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
            self.dist, self.torch = dist, torch

    def barrier(self):
        if self.dist:
            self.dist.barrier()

    def max(self, v):
        if not self.dist:
            return v
        t = self.torch.tensor([v], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def sum(self, v):
        if not self.dist:
            return v
        t = self.torch.tensor([v], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return float(t.item())

    def min(self, v):
        if not self.dist:
            return v
        t = self.torch.tensor([v], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MIN)
        return float(t.item())

    def broadcast_object(self, obj):
        if not self.dist:
            return obj
        box = [obj]
        self.dist.broadcast_object_list(box, src=0)
        return box[0]

    def all_gather_bytes(self, b):
        if not self.dist:
            return [b]
        out = [None] * self.world
        self.dist.all_gather_object(out, b)
        return out

    def close(self):
        if self.dist:
            self.dist.barrier()
            self.dist.destroy_process_group()


def pinned(n):
    """float64 host buffer in pinned memory (torch is the allocator; plumbing only)."""
    try:
        import torch
        return torch.empty(n, dtype=torch.float64, pin_memory=True).numpy()
    except Exception:
        return np.empty(n)


def make_solver(pl, wl, n_gpus):
    import ctypes
    from cuda_iblb_11_b200 import HaloHandle, Params, Solver
    w = WORKLOADS[wl]
    if w["scaling"] == "weak":
        X_total, Wl = w["X"] * n_gpus, w["X"]
    elif w["scaling"] == "strong":
        X_total, Wl = w["X"], w["X"] // n_gpus
    else:  # replicas: every rank runs the whole (small) case
        X_total, Wl = w["X"], w["X"]
    slabbed = n_gpus > 1 and w["scaling"] != "replicas"
    p = Params.channel(X_total, w["Y"], c_num=w["c_num"], c_space=w["c_space"], gx=w["gx"],
                       c_fraction=w.get("c_fraction", max(1, w["c_num"])), device=pl.local, slab_x0=pl.rank * Wl if slabbed else 0,
                       slab_w=Wl if slabbed else 0)
    s = Solver(p)
    if slabbed:
        mine = bytes(s.export_halo())
        hs = [HaloHandle.from_buffer_copy(b) for b in pl.all_gather_bytes(mine)]
        s.connect(hs[(pl.rank - 1) % n_gpus], hs[(pl.rank + 1) % n_gpus])
        pl.barrier()
    nodes_total = X_total * w["Y"] * (n_gpus if w["scaling"] == "replicas" else 1)
    return s, p, nodes_total, Wl * w["Y"]


def config_block(workload, X, Y, n, nodes_local):
    """`config` of the JSON line, the same keys (and at N = 1 the same values) in both arms."""
    w = WORKLOADS[workload]
    return {"workload": "%s: %s" % (workload, w["desc"]),
            "grid": [int(X), int(Y)], "columns_per_gpu": int(nodes_local // int(Y)),
            "parallelism": ("1 GPU" if n == 1 else "%d independent replicas, one per GPU" % n
                            if w["scaling"] == "replicas"
                            else "x-slabs x%d, NVLink peer-store halos" % n),
            "l2_policy": "working set %.1f GB per GPU >> 126 MB L2" % (nodes_local * 144 / 1e9)
            if nodes_local * 144 > 1e9 else "L2-resident (latency-bound case)"}


def _oracle():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import checkers
    return checkers


def parity_c2(pl, s, p, steps_done, w):
    """Self-check of the timed C2 run, every rank on its own columns: the forced channel is
    x-invariant, so all columns must be EXACTLY equal, and the column profile must equal the
    oracle's 4-column channel of the same height after the same number of steps (what
    tests/test_gpu_slabs_fullsize.py::test_c2_full_size_equals_narrow_oracle_channel does on one
    GPU).  The oracle is the checker here, never the thing measured."""
    Y, W = int(p.YDIM), s.width
    rho, u = s.fields()
    rho = rho.reshape(Y, W)
    ux = u[:W * Y].reshape(Y, W)
    x_inv = bool(np.array_equal(rho, np.broadcast_to(rho[:, :1], rho.shape)) and
                 np.array_equal(ux, np.broadcast_to(ux[:, :1], ux.shape)))
    prof = None
    if pl.rank == 0:
        ck = _oracle()
        o = ck.Checker("oracle", ck.make_params(4, Y, gx=w["gx"], threads=os.cpu_count() or 1))
        o.step(steps_done)
        ro, uo = o.fields()
        o.close()
        prof = (ro.reshape(Y, 4)[:, 0].copy(), uo[:4 * Y].reshape(Y, 4)[:, 0].copy())
    prof = pl.broadcast_object(prof)
    e_rho = float(np.linalg.norm(rho[:, 0] - prof[0]) / np.linalg.norm(prof[0]))
    e_ux = float(np.linalg.norm(ux[:, 0] - prof[1]) / max(np.linalg.norm(prof[1]), 1e-300))
    a_ux = float(np.abs(ux[:, 0] - prof[1]).max())
    # u starts from rest and is still ~1e-5 after 100 steps, a difference of O(0.1) populations:
    # the rounding of those (1e-17 absolute per step, FMA vs none) is what bounds the error, so
    # the relative tolerance is backed by an absolute one (as tests/test_gpu_slabs_fullsize.py does)
    ok = x_inv and e_rho <= 1e-10 and (e_ux <= 1e-10 or a_ux <= max(1e-16 * steps_done, 2e-14))
    ok_all = pl.min(1.0 if ok else 0.0) == 1.0
    return {"ok": ok_all, "what": "every rank: exact x-invariance of its columns + column profile vs "
                                  "the oracle's 4x%d channel after %d steps (tol 1e-10)" % (Y, steps_done),
            "x_invariant": bool(pl.min(1.0 if x_inv else 0.0) == 1.0),
            "rel_l2_rho": pl.max(e_rho), "rel_l2_ux": pl.max(e_ux), "max_abs_ux": pl.max(a_ux),
            "ux_rule": "rel L2 <= 1e-10 or max abs <= max(1e-16 x steps, 2e-14) (u is a difference of "
                       "O(0.1) populations)"}


def parity_c2_moving(pl, w, steps=25, X=128):
    """The C2 kernel on a channel that is NOT at rest (u0_amp = 1e-3, the profile of
    tests/test_gpu_slabs_fullsize.py): here the relative bound on u means something, no absolute
    floor.  Rank 0, a narrow channel of the full height (the forced channel is x-invariant)."""
    if pl.rank != 0:
        return None
    from cuda_iblb_11_b200 import Params, Solver
    Y = w["Y"]
    s = Solver(Params.channel(X, Y, gx=w["gx"], u0_amp=1e-3, device=pl.local))
    s.step(steps)
    rho, u = s.fields()
    s.close()
    ck = _oracle()
    o = ck.Checker("oracle", ck.make_params(4, Y, gx=w["gx"], u0_amp=1e-3, threads=os.cpu_count() or 1))
    o.step(steps)
    ro, uo = o.fields()
    o.close()
    rho, ux = rho.reshape(Y, X), u[:X * Y].reshape(Y, X)
    x_inv = bool(np.array_equal(rho, np.broadcast_to(rho[:, :1], rho.shape)) and
                 np.array_equal(ux, np.broadcast_to(ux[:, :1], ux.shape)))
    pr, pu = ro.reshape(Y, 4)[:, 0], uo[:4 * Y].reshape(Y, 4)[:, 0]
    e_rho = float(np.linalg.norm(rho[:, 0] - pr) / np.linalg.norm(pr))
    e_ux = float(np.linalg.norm(ux[:, 0] - pu) / np.linalg.norm(pu))
    return {"ok": bool(x_inv and e_rho <= 1e-10 and e_ux <= 1e-10),
            "what": "%dx%d channel, u0_amp 1e-3, %d steps: exact x-invariance + column profile vs the "
                    "oracle's 4x%d channel, rel L2 <= 1e-10 on rho and u_x, no absolute floor" % (X, Y, steps, Y),
            "x_invariant": x_inv, "rel_l2_rho": e_rho, "rel_l2_ux": e_ux, "max_ux": float(np.abs(pu).max())}


C4S = dict(X=4096, Y=1024, c_num=32, c_space=128, steps=40)  # down-scaled C4 for the self-check


def parity_c4s(pl, n):
    """Down-scaled C4 (4096x1024, 32 cilia 128 apart): every rank steps its x-slab of the ring
    (CUDA IPC peer stores + device flags, the bench's N>1 path) AND the whole channel on its own
    GPU (single-launch IB+LB path, split for tall grids) and compares its columns bit for bit;
    rank 0 also holds the whole channel against the oracle.  N = 1: fused vs unfused chain."""
    from cuda_iblb_11_b200 import FLAG_NO_FUSE, FLAG_NO_GRAPH, HaloHandle, Params, Solver
    c = C4S
    X, Y, steps = c["X"], c["Y"], c["steps"]
    kw = dict(c_num=c["c_num"], c_space=c["c_space"], c_fraction=c["c_num"], device=pl.local)
    whole = Solver(Params.channel(X, Y, **kw))
    whole.step(steps)
    r2, u2 = whole.fields()
    if n > 1:
        Wl = X // n
        x0 = pl.rank * Wl
        part = Solver(Params.channel(X, Y, slab_x0=x0, slab_w=Wl, **kw))
        mine = bytes(part.export_halo())
        hs = [HaloHandle.from_buffer_copy(b) for b in pl.all_gather_bytes(mine)]
        part.connect(hs[(pl.rank - 1) % n], hs[(pl.rank + 1) % n])
        pl.barrier()
        part.step(steps)
        part.sync()
        pl.barrier()
        how = "x-slabs x%d (IPC) == whole channel, bitwise" % n
    else:
        Wl, x0 = X, 0
        part = Solver(Params.channel(X, Y, flags=FLAG_NO_FUSE | FLAG_NO_GRAPH, **kw))
        part.step(steps)
        how = "single-launch IB+LB path == unfused kernel chain, bitwise"
    r1, u1 = part.fields()
    sl = slice(x0, x0 + Wl)
    same = bool(np.array_equal(r1.reshape(Y, Wl), r2.reshape(Y, X)[:, sl]) and
                np.array_equal(u1[:Wl * Y].reshape(Y, Wl), u2[:X * Y].reshape(Y, X)[:, sl]) and
                np.array_equal(u1[Wl * Y:].reshape(Y, Wl), u2[X * Y:].reshape(Y, X)[:, sl]))
    q_part = pl.sum(part.Q) if n > 1 else part.Q
    same_q = bool(q_part == whole.Q)
    pl.barrier()
    part.close()
    out = {"grid": [X, Y], "cilia": c["c_num"], "steps": steps, "how": how,
           "bitwise_equal": bool(pl.min(1.0 if same else 0.0) == 1.0), "flux_equal": same_q}
    if pl.rank == 0:
        ck = _oracle()
        o = ck.Checker("oracle", ck.make_params(X, Y, c_num=c["c_num"], c_space=c["c_space"],
                                               c_fraction=c["c_num"], flags=ck.REF_FLAG_PICK_LOWEST_K,
                                               threads=os.cpu_count() or 1))
        o.step(steps)
        ro, uo = o.fields()
        m1, m2 = whole.markers(), o.markers()
        out["vs_oracle"] = {"rel_l2_rho": ck.rel_l2(r2, ro), "rel_l2_u": ck.rel_l2(u2, uo),
                            "markers_bit_equal": bool(np.array_equal(m1[0], m2[0]) and
                                                      np.array_equal(m1[1], m2[1]) and
                                                      np.array_equal(m1[3], m2[3])),
                            "tol": "rho 1e-10, u 1e-8, markers bit-equal"}
        v = out["vs_oracle"]
        out["oracle_ok"] = bool(v["rel_l2_rho"] <= 1e-10 and v["rel_l2_u"] <= 1e-8 and
                                v["markers_bit_equal"])
        o.close()
    whole.close()
    ok = out["bitwise_equal"] and out["flux_equal"] and out.get("oracle_ok", True)
    out["ok"] = bool(pl.min(1.0 if ok else 0.0) == 1.0)
    return out


def c4_leg(pl, n, steps=20, warmup=3):
    """configs[3]: 32768x32768 + 256 cilia, x-slabs over the N GPUs (strong scaling; one GPU holds
    all of it: 154.6 GB).  Same timing rules as the headline (device events, max over ranks)."""
    s, p, nodes_total, nodes_local = make_solver(pl, "c4", n)
    s.step(warmup)
    s.sync()
    pl.barrier()
    s.step(steps)
    s.sync()
    t = pl.max(s.last_step_seconds)
    launches = int(pl.sum(float(s.last_step_launches)))
    q = pl.sum(s.Q) if n > 1 else s.Q
    pl.barrier()
    s.close()
    return {"workload": WORKLOADS["c4"]["desc"], "grid": [int(p.XDIM), int(p.YDIM)], "n_gpus": n,
            "scaling": "strong", "value": round(nodes_total * steps / t / 1e6, 1), "unit": "MLUPS",
            "ms_per_step": round(1e3 * t / steps, 4), "steps": steps, "warmup": warmup,
            "gpu_launches": launches, "flux_finite": bool(np.isfinite(q)),
            "frac_of_measured_hbm_peak": round(BYTES_PER_LUP * nodes_total / n * steps / t / 1e9 /
                                               measured_peak()[0], 4)}


def ib_share_leg(device, steps=4000):
    """IB share of the step at 12 cilia (north_star target <= 10 %): the cilia12.sh grid with its
    cilia (single-launch IB+LB step) against the same grid as pure LB, both CUDA-graph replayed,
    device time by events; share = (fused - pure) / fused."""
    from cuda_iblb_11_b200 import Params, Solver
    w = WORKLOADS["c3"]

    def one(X, Y, c_num, c_space, c_fraction):
        res = {}
        for name, kw in (("fused", dict(c_num=c_num, c_space=c_space, c_fraction=c_fraction)),
                         ("pure", dict(gx=1e-7))):
            s = Solver(Params.channel(X, Y, device=device, **kw))
            s.step(200)
            s.sync()
            best = 1e9
            for _ in range(3):
                s.step(steps)
                s.sync()
                best = min(best, s.last_step_seconds / steps)
            s.close()
            res[name] = best * 1e6
        return {"grid": [X, Y], "cilia": c_num, "fused_us": round(res["fused"], 3),
                "pure_lb_us": round(res["pure"], 3),
                "share": round((res["fused"] - res["pure"]) / res["fused"], 4)}

    c3 = one(w["X"], w["Y"], w["c_num"], w["c_space"], w["c_fraction"])
    w1 = WORKLOADS["c1"]
    others = [one(w1["X"], w1["Y"], w1["c_num"], w1["c_space"], w1["c_fraction"]),  # cilia6.sh grid
              one(2304, 192, 48, 48, 48)]                                             # 4 x the cilia12.sh width
    return {"workload": "c3: " + w["desc"], "fused_us": c3["fused_us"],
            "pure_lb_us": c3["pure_lb_us"], "steps": steps, "share": c3["share"],
            "other_grids": others,
            "how": "iblb_last_step_seconds of %d graph-replayed steps, best of 3, on two solvers" % steps}


def c5_leg(pl, n, steps=4000, warmup=200):
    """configs[4] (multiapp.sh:7-9, cilia6.sh:7-17 as one run per GPU): N independent replicas of the
    cilia6.sh case, one per rank, no coupling; value = nodes of all replicas x steps / the slowest
    rank's device time.  In-phase line of the sweep (the phase-lagged ones overflow within ~40
    steps in the reference too; work per step is the same)."""
    from cuda_iblb_11_b200 import Params, Solver
    w = WORKLOADS["c1"]
    s = Solver(Params.channel(w["X"], w["Y"], c_num=w["c_num"], c_space=w["c_space"],
                              c_fraction=w["c_fraction"], device=pl.local))
    s.step(warmup)
    s.sync()
    pl.barrier()
    s.step(steps)
    s.sync()
    t = pl.max(s.last_step_seconds)
    launches = int(pl.sum(float(s.last_step_launches)))
    q = s.Q
    finite = pl.min(1.0 if np.isfinite(q) else 0.0) == 1.0
    pl.barrier()
    s.close()
    nodes = w["X"] * w["Y"] * n
    return {"workload": "c5: %d replicas of %s" % (n, w["desc"]), "grid": [w["X"], w["Y"]],
            "n_gpus": n, "scaling": "replicas", "value": round(nodes * steps / t / 1e6, 1),
            "unit": "MLUPS", "us_per_step": round(1e6 * t / steps, 3), "steps": steps,
            "warmup": warmup, "gpu_launches": launches, "flux_finite": bool(finite)}


def run_ours(a):
    pl = Plumbing(a.gpus)
    n = max(1, a.gpus)
    s, p, nodes_total, nodes_local = make_solver(pl, a.workload, n)
    sampler = ClockSampler(pl.local)
    if pl.rank == 0:
        sampler.start()
        time.sleep(0.25)
    # warm-up right before the timed region (an idle GPU drops its clocks within a fraction of a
    # second, which would otherwise be billed to the first timed steps)
    s.step(a.warmup)
    s.sync()
    pl.barrier()
    # ---- timed region: K steps, device time by CUDA events on the solver's stream ----
    pl.barrier()
    s.sync()
    t0 = time.time()
    s.step(a.steps)
    s.sync()
    t_dev = s.last_step_seconds
    launches = s.last_step_launches
    pl.barrier()
    t1 = time.time()
    clocks = sampler.stop(t0, t1) if pl.rank == 0 else None
    t = pl.max(t_dev)
    launches_total = int(pl.sum(float(launches)))
    value = nodes_total * a.steps / t / 1e6

    # ---- self-check of the state the timed steps produced (before the e2e leg resets it) ----
    w_cfg = WORKLOADS[a.workload]
    parity = None
    if not a.no_parity:
        try:
            if a.workload == "c2":
                parity = {"c2": parity_c2(pl, s, p, a.warmup + a.steps, w_cfg)}
        except Exception as e:  # reported, never hidden
            parity = {"c2": {"ok": False, "error": "%s: %s" % (type(e).__name__, e)}}

    # ---- e2e: same metric through the C ABI with HOST buffers: upload (rho,u), K steps with the
    # reference's per-step flux readback (main.cu:928), download (rho,u) (main.cu:944-952) ----
    q = 0.0
    if nodes_local <= (1 << 27) or a.force_e2e:
        rho = pinned(nodes_local)
        u = pinned(2 * nodes_local)
        rho[:] = 1.0
        u[:] = 0.0
        pl.barrier()
        s.sync()
        e0 = time.perf_counter()
        s.set_fields(rho, u)
        pl.barrier()  # every slab re-initialised before any neighbour pulls from its ghosts
        if nodes_local <= (1 << 22):
            # launch-bound grids: the reference's "Q every step" (main.cu:928) comes from the
            # device-side flux history, one read per batch, so the CUDA-graph path stays on
            done = 0
            while done < a.steps:
                k = min(2048, a.steps - done)
                s.step(k)
                q = float(s.flux_history(done, k)[-1])
                done += k
        else:
            for _ in range(a.steps):
                s.step(1)
                q = s.Q
        s.fields(rho, u)
        pl.barrier()
        e1 = pl.max(time.perf_counter() - e0)
        e2e_value = round(nodes_total * a.steps / e1 / 1e6, 1)
        h2d = 24 * nodes_local * n / a.steps
        d2h = (24 * nodes_local * n + 8 * max(1, p.YDIM // 64) * n * a.steps) / a.steps
    else:
        # 24 B/node of pinned host memory per slab (>3 GB): not attempted unless asked for
        e2e_value, h2d, d2h = None, 0, 0

    peak, peak_src = measured_peak()
    # dominant kernel = lb_fused_v4 (v2 on the 2G boundary columns of a slab): in the pure-LB workload
    # the timed region is K launches of it
    lb_launches_per_rank = a.steps * (1 if n == 1 or WORKLOADS[a.workload]["scaling"] == "replicas" else 2)
    achieved = BYTES_PER_LUP * (nodes_total / n) * a.steps / t_dev / 1e9
    tr = ncu_traffic()
    # (whole channel with cilia on one GPU, short grid: the single-launch IB+LB kernel instead)
    w_ = WORKLOADS[a.workload]
    whole = n == 1 or w_["scaling"] == "replicas"
    kernel = ("lb_ib_tile_step" if (w_["c_num"] > 0 and whole and w_["Y"] <= 512)
              else "lb_fused_v4" if (w_["c_num"] == 0 and w_["Y"] % 128 == 0) else "lb_fused_v2")
    roof = {"bound": "hbm", "kernel": kernel, "achieved": round(achieved, 1),
            "peak": peak, "peak_source": peak_src, "unit": "GB/s",
            "frac": round(achieved / peak, 4),
            # the measured peak is a torch copy_, which this kernel can exceed; HGX B200 nominal
            # HBM3e bandwidth for reference (B200_PROFILING.md)
            "nominal_peak": 7700.0, "frac_of_nominal": round(achieved / 7700.0, 4),
            "algorithmic_bytes_per_launch": BYTES_PER_LUP * nodes_local,
            "avg_launch_ms": round(1e3 * t_dev / a.steps, 4),
            "launches_of_kernel": lb_launches_per_rank,
            # ncu capture of this kernel on one GPU over the whole channel; not measured on slabs
            "traffic": tr.get("dram_bytes_per_launch") if (tr and a.workload == "c2" and n == 1) else None,
            # not measured in this run: read from a committed builder-run ncu capture of the same kernel
            "traffic_source": ("profiles/roofline_traffic.json (builder-run ncu --set full capture, "
                               "replayed here)") if (tr and a.workload == "c2" and n == 1) else None}
    out = None
    if pl.rank == 0:
        out = {
            "metric": "MLUPS", "value": round(value, 1), "unit": "MLUPS", "n_gpus": n,
            "steps": a.steps, "warmup": a.warmup, "ms_per_step": round(1e3 * t / a.steps, 4),
            "higher_is_better": True, "scaling": WORKLOADS[a.workload]["scaling"],
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_block(a.workload, p.XDIM, p.YDIM, n, nodes_local),
            "gpu_launches": launches_total,
            "e2e": {"value": e2e_value, "unit": "MLUPS",
                    "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "what": "iblb_set_fields(host rho,u) + K x [iblb_step(1) + iblb_get_flux] (small grids: "
                            "iblb_step(batch) + iblb_get_flux_history(batch)) + iblb_get_fields(host), "
                            "wall clock, max over ranks"},
            "roofline": roof, "clocks": clocks, "flux_check": q,
        }
    s.close()
    # ---- further legs on the same box, reported inside the same JSON line ----
    extras = {}
    if not a.no_parity:
        try:
            c4s = parity_c4s(pl, n)
        except Exception as e:
            c4s = {"ok": False, "error": "%s: %s" % (type(e).__name__, e)}
        parity = dict(parity or {}, c4s=c4s)
        if a.workload == "c2":
            try:
                mv = parity_c2_moving(pl, WORKLOADS["c2"])
            except Exception as e:
                mv = {"ok": False, "error": "%s: %s" % (type(e).__name__, e)}
            if mv is not None:
                parity["c2_moving"] = mv
        parity["ok"] = all(v.get("ok", False) for v in parity.values() if isinstance(v, dict))
    if a.workload == "c2" and not a.no_extras:
        try:
            extras["c4"] = c4_leg(pl, n)
        except Exception as e:
            extras["c4"] = {"error": "%s: %s" % (type(e).__name__, e), "n_gpus": n}
        try:
            extras["c5"] = c5_leg(pl, n)
        except Exception as e:
            extras["c5"] = {"error": "%s: %s" % (type(e).__name__, e), "n_gpus": n}
        if n == 1:
            try:
                extras["ib_share"] = ib_share_leg(pl.local)
            except Exception as e:
                extras["ib_share"] = {"error": "%s: %s" % (type(e).__name__, e)}
    if out is not None:
        out["parity"] = parity
        out.update(extras)
    return pl, out


def cpu_baseline(a, target_seconds=12.0):
    """Oracle port (plain C + OpenMP, all host cores) on a bounded sample of the same workload:
    a narrower channel of the same height; MLUPS is per-node so the figure carries over."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from checkers import Checker, make_params
    w = WORKLOADS[a.workload]
    cores = os.cpu_count() or 1
    if w["c_num"] == 0:
        X, Y, c_num = 256, w["Y"], 0
    else:
        X, Y, c_num = 48 * 6, 192, 6
    c = Checker("oracle", make_params(X, Y, c_num=c_num, c_fraction=max(1, c_num), gx=w["gx"],
                                      threads=cores))
    c.step(1)
    c.step(2)
    per = max(c.last_step_seconds / 2, 1e-6)
    n = int(max(2, min(2000, target_seconds / per)))
    c.step(n)
    t = c.last_step_seconds
    c.close()
    return {"value": round(X * Y * n / t / 1e6, 2), "unit": "MLUPS", "cores": cores,
            "kind": "port",
            "sample": "oracle port (C+OpenMP), %dx%d%s, %d steps, %.1f s" % (
                X, Y, " with %d cilia" % c_num if c_num else " pure LB", n, t)}


def run_reference(a):
    """Reference arm: the reference's own kernels.  The reference has no CPU path (BASELINE.json),
    so this is its CUDA build for sm_100a (oracle/_ref/libiblb_ref_cuda.so, unmodified sources)
    on one B200; if that library is absent, the same kernels compiled as host C++
    (libiblb_ref_cpu.so) or the oracle port on a bounded sample."""
    world, rank, local = dist_env()
    if rank != 0:
        return None
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from checkers import Checker, available, make_params
    w = WORKLOADS[a.workload]
    cores = os.cpu_count() or 1
    X, Y, c_num = w["X"], w["Y"], w["c_num"]
    kind, sample = "ref_cuda", None
    if w["scaling"] == "strong" or not available("ref_cuda"):
        # the reference cannot address 32768^2 (32-bit 9*j index, 352 GB): largest comparable grid
        if available("ref_cuda"):
            X, Y, c_num = 8192, 192, 64
            sample = "reference CUDA build on 8192x192 with 64 cilia (cannot address 32768^2)"
        else:
            kind = "ref_cpu" if available("ref_cpu") else "oracle"
            X, Y, c_num = (256, w["Y"], 0) if w["c_num"] == 0 else (288, 192, 6)
            sample = "%s on %dx%d (no GPU build of the reference available)" % (kind, X, Y)
    c = Checker(kind, make_params(X, Y, c_num=c_num, c_space=w["c_space"] if c_num else 48,
                                  c_fraction=max(1, c_num), gx=w["gx"], threads=cores))
    steps, warm = a.steps, a.warmup
    if kind != "ref_cuda":
        steps, warm = min(steps, 3), 1
    sampler = ClockSampler(local)
    sampler.start()
    time.sleep(0.25)
    c.step(warm)
    t0 = time.time()
    c.step(steps)
    t = c.last_step_seconds
    t1 = time.time()
    clocks = sampler.stop(t0, t1)
    c.close()
    value = X * Y * steps / t / 1e6
    launches = steps * (4 + (4 if c_num else 0)) if kind == "ref_cuda" else 0
    if sample is None:
        sample = "reference CUDA build (nvcc sm_100a, unmodified sources) on the full %dx%d" % (X, Y)
    return {
        "impl": "reference", "metric": "MLUPS", "value": round(value, 2), "unit": "MLUPS",
        "n_gpus": 1, "steps": steps, "warmup": warm, "ms_per_step": round(1e3 * t / steps, 4),
        "higher_is_better": True, "scaling": w["scaling"], "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        # the reference is single-GPU whatever N the driver launches: its line describes one GPU
        "config": config_block(a.workload, X, Y, 1, X * Y),
        "gpu_launches": launches,
        "cpu_baseline": {"value": round(value, 2), "unit": "MLUPS", "cores": cores,
                         "kind": "reference", "sample": sample + "; the reference has no CPU path"},
        "e2e": {"value": round(value, 2), "unit": "MLUPS", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "clocks": clocks,
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--force-e2e", action="store_true", help="run the host-buffer leg at any size")
    ap.add_argument("--no-parity", action="store_true", help="skip the self-check legs")
    ap.add_argument("--no-extras", action="store_true", help="skip the C4 and IB-share legs")
    a = ap.parse_args()
    a.warmup = max(a.warmup, 3)
    if a.impl == "reference":
        out = run_reference(a)
        if out is not None:
            print(json.dumps(out), flush=True)
        return 0
    pl, out = run_ours(a)
    if out is not None:
        if max(1, a.gpus) == 1 and not a.no_cpu_baseline:
            out["cpu_baseline"] = cpu_baseline(a)
        print(json.dumps(out), flush=True)
    pl.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
