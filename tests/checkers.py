"""TEST INFRASTRUCTURE: ctypes bindings of the three checkers behind oracle/ref_api.h.

  Checker("oracle")    oracle/libiblb_oracle.so         plain-C restatement (always available)
  Checker("ref_cpu")   oracle/_ref/libiblb_ref_cpu.so   reference kernels as host C++ (shim)
  Checker("ref_cuda")  oracle/_ref/libiblb_ref_cuda.so  reference kernels, nvcc sm_100a (GPU box)

Nothing in the product package imports this module.
"""
import ctypes
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

REF_FLAG_HALF_FORCE_BODY = 1
REF_FLAG_PICK_LOWEST_K = 2
REF_FLAG_MUCUS = 4  # oracle port; the reference side is libiblb_ref_cpu_mucus.so ("ref_cpu_mucus")


class RefParams(ctypes.Structure):
    _fields_ = [
        ("XDIM", ctypes.c_int), ("YDIM", ctypes.c_int), ("c_num", ctypes.c_int),
        ("c_space", ctypes.c_int), ("T", ctypes.c_int), ("p_step", ctypes.c_int),
        ("TAU", ctypes.c_double), ("TAU2", ctypes.c_double),
        ("gx", ctypes.c_double), ("gy", ctypes.c_double), ("u0_amp", ctypes.c_double),
        ("threads", ctypes.c_int), ("flags", ctypes.c_int),
        ("length", ctypes.c_int), ("fine_points", ctypes.c_int),
        ("shape_scale", ctypes.c_double), ("flux_back", ctypes.c_int),
    ]


class OracleDerived(ctypes.Structure):
    _fields_ = [
        ("XDIM", ctypes.c_uint), ("T", ctypes.c_uint), ("ITERATIONS", ctypes.c_uint),
        ("INTERVAL", ctypes.c_uint), ("p_step", ctypes.c_int),
        ("SPEED", ctypes.c_double), ("t_scale", ctypes.c_double), ("x_scale", ctypes.c_double),
        ("s_scale", ctypes.c_double), ("TAU", ctypes.c_double), ("TAU2", ctypes.c_double),
        ("Ma", ctypes.c_double),
    ]


_LIBS = {
    "oracle": ("oracle/libiblb_oracle.so", "oracle_"),
    "ref_cpu": ("oracle/_ref/libiblb_ref_cpu.so", "ref_"),
    "ref_cuda": ("oracle/_ref/libiblb_ref_cuda.so", "ref_"),
    # the reference with its commented-out "WITH MUCUS PRESENT" coefficient block swapped in
    "ref_cpu_mucus": ("oracle/_ref/libiblb_ref_cpu_mucus.so", "ref_"),
}


def lib_path(kind):
    return os.path.join(ROOT, _LIBS[kind][0])


def available(kind):
    return os.path.exists(lib_path(kind))


def reference_tau(T=100000, Re=1.0, LENGTH=96):
    """TAU, TAU2 of main.cu:314-321 (host C_S = 0.577, LENGTH = 96)."""
    speed = 0.8 * 1000 / T
    tau = (speed * LENGTH) / (Re * 0.577 * 0.577) + 1. / 2.
    tau2 = 1. / (12. * (tau - (1. / 2.))) + (1. / 2.)
    return tau, tau2


def make_params(XDIM, YDIM=192, c_num=0, c_space=48, T=100000, c_fraction=1, Re=1.0,
                gx=0.0, gy=0.0, u0_amp=0.0, threads=0, flags=0, TAU=None, TAU2=None,
                length=0, fine_points=0, shape_scale=0.0, flux_back=0):
    tau, tau2 = reference_tau(T, Re, length or 96)
    p_step = (T * c_fraction // c_num) if c_num else 0
    return RefParams(XDIM, YDIM, c_num, c_space, T, p_step,
                     tau if TAU is None else TAU, tau2 if TAU2 is None else TAU2,
                     gx, gy, u0_amp, threads, flags, length, fine_points, shape_scale, flux_back)


class Checker:
    def __init__(self, kind, params):
        path, prefix = _LIBS[kind]
        self.kind = kind
        self.lib = ctypes.CDLL(os.path.join(ROOT, path))
        self.p = params
        g = lambda name: getattr(self.lib, prefix + name)
        self._create, self._step, self._destroy = g("create"), g("step"), g("destroy")
        self._get_fields, self._get_force, self._get_f = g("get_fields"), g("get_force"), g("get_f")
        self._get_markers, self._get_Q, self._get_it = g("get_markers"), g("get_Q"), g("get_it")
        self._secs = g("last_step_seconds")
        self._create.restype = ctypes.c_void_p
        self._create.argtypes = [ctypes.POINTER(RefParams)]
        self._step.argtypes = [ctypes.c_void_p, ctypes.c_int]
        self._destroy.argtypes = [ctypes.c_void_p]
        for fn, n in ((self._get_fields, 2), (self._get_force, 1), (self._get_f, 1),
                      (self._get_markers, 4)):
            fn.argtypes = [ctypes.c_void_p] + [ctypes.c_void_p] * n
        self._get_Q.restype = ctypes.c_double
        self._get_Q.argtypes = [ctypes.c_void_p]
        self._get_it.argtypes = [ctypes.c_void_p]
        self._secs.restype = ctypes.c_double
        self._secs.argtypes = [ctypes.c_void_p]
        self.h = self._create(ctypes.byref(params))
        if not self.h:
            raise RuntimeError("checker %s: create failed" % kind)
        self.size = params.XDIM * params.YDIM
        self.Ns = (params.length or 96) * params.c_num

    def set_marker32_trace(self, trace):
        """oracle port only: trace[it, m] != 0 -> the higher fine point wins marker 32 at step it."""
        t = np.ascontiguousarray(trace, np.uint8)
        self.lib.oracle_set_marker32_trace.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
        self.lib.oracle_set_marker32_trace(self.h, t.ctypes.data, t.shape[0])

    def set_marker32_replay(self, vals):
        """oracle port only: vals[it, m, 0:4] = (s_x, s_y, u_s_x, u_s_y) of marker 32 of cilium m
        at step it, as the reference's CUDA build produced them (covers torn racing stores)."""
        t = np.ascontiguousarray(vals, np.float32)
        self.lib.oracle_set_marker32_replay.argtypes = [ctypes.c_void_p, ctypes.c_void_p,
                                                        ctypes.c_int]
        self.lib.oracle_set_marker32_replay(self.h, t.ctypes.data, t.shape[0])

    def step(self, n=1):
        rc = self._step(self.h, n)
        if rc != 0:
            raise RuntimeError("checker %s: step failed rc=%d" % (self.kind, rc))

    def fields(self):
        rho = np.empty(self.size)
        u = np.empty(2 * self.size)
        self._get_fields(self.h, rho.ctypes.data, u.ctypes.data)
        return rho, u

    def force(self):
        f = np.empty(2 * self.size)
        self._get_force(self.h, f.ctypes.data)
        return f

    def populations(self):
        f = np.empty(9 * self.size)
        self._get_f(self.h, f.ctypes.data)
        return f

    def markers(self):
        n = max(self.Ns, 1)
        s = np.zeros(2 * n, np.float32)
        u_s = np.zeros(2 * n, np.float32)
        F_s = np.zeros(2 * n, np.float32)
        eps = np.zeros(n, np.int32)
        if self.Ns:
            self._get_markers(self.h, s.ctypes.data, u_s.ctypes.data, F_s.ctypes.data,
                              eps.ctypes.data)
        return s[:2 * self.Ns], u_s[:2 * self.Ns], F_s[:2 * self.Ns], eps[:self.Ns]

    @property
    def Q(self):
        return self._get_Q(self.h)

    @property
    def it(self):
        return self._get_it(self.h)

    @property
    def last_step_seconds(self):
        return self._secs(self.h)

    def close(self):
        if self.h:
            self._destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def oracle_derive(c_fraction, c_num, c_space, Re, T_num, T_pow, I_pow, P_num):
    lib = ctypes.CDLL(lib_path("oracle"))
    lib.oracle_derive.argtypes = [ctypes.c_uint, ctypes.c_uint, ctypes.c_uint, ctypes.c_double,
                                  ctypes.c_float, ctypes.c_uint, ctypes.c_float, ctypes.c_uint,
                                  ctypes.POINTER(OracleDerived)]
    d = OracleDerived()
    lib.oracle_derive(c_fraction, c_num, c_space, Re, T_num, T_pow, I_pow, P_num, ctypes.byref(d))
    return d


def rel_l2(a, b):
    """relative L2 difference ||a-b|| / ||b|| (b = the reference side)."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    den = np.linalg.norm(b)
    num = np.linalg.norm(a - b)
    return num / den if den > 0 else num
