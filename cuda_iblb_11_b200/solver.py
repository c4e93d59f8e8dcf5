"""ctypes host mirror of include/iblb.h — the Python face of libiblb.so.

Names and argument meaning follow the reference's driver (CUDA_IBLB_11/main.cu:263-321): a run is
described by the ten positional arguments `c_fraction c_num c_space Re T_num T_pow I_pow P_num
ShARC BigData`; `Params.from_args` parses the same line `./app` takes.  There is no CPU fallback:
if libiblb.so is missing or a CUDA call fails, an exception is raised.
"""
import ctypes
import os

import numpy as np

_PKG = os.path.dirname(os.path.abspath(__file__))
# IBLB_LIB: development only — another build of the same library (libiblb_checked.so)
LIB_PATH = os.environ.get("IBLB_LIB") or os.path.join(_PKG, "libiblb.so")

FLAG_HALF_FORCE_BODY = 1
FLAG_PICK_HIGHEST_K = 2
FLAG_NO_GRAPH = 4
FLAG_NO_FUSE = 8
FLAG_NO_PDL = 16
FLAG_DEV_KNOBS = 64   # honour IBLB_V4 / IBLB_TILE_W / IBLB_BAND_FIRST (development A/B)
FLAG_MUCUS = 32

IPC_HANDLE_BYTES = 64


class IblbError(RuntimeError):
    pass


class Params(ctypes.Structure):
    """struct iblb_params (include/iblb.h)."""
    _fields_ = [
        ("c_fraction", ctypes.c_uint), ("c_num", ctypes.c_uint), ("c_space", ctypes.c_uint),
        ("Re", ctypes.c_double), ("T_num", ctypes.c_float), ("T_pow", ctypes.c_uint),
        ("I_pow", ctypes.c_float), ("P_num", ctypes.c_uint),
        ("ShARC", ctypes.c_int), ("BigData", ctypes.c_int),
        ("XDIM", ctypes.c_uint), ("YDIM", ctypes.c_uint), ("T", ctypes.c_uint),
        ("ITERATIONS", ctypes.c_uint), ("INTERVAL", ctypes.c_uint), ("LENGTH", ctypes.c_uint),
        ("p_step", ctypes.c_int),
        ("SPEED", ctypes.c_double), ("t_scale", ctypes.c_double), ("x_scale", ctypes.c_double),
        ("s_scale", ctypes.c_double), ("TAU", ctypes.c_double), ("TAU2", ctypes.c_double),
        ("Ma", ctypes.c_double),
        ("gx", ctypes.c_double), ("gy", ctypes.c_double), ("u0_amp", ctypes.c_double),
        ("flags", ctypes.c_uint), ("device", ctypes.c_int),
        ("slab_x0", ctypes.c_uint), ("slab_w", ctypes.c_uint),
        ("fine_points", ctypes.c_uint), ("shape_scale", ctypes.c_double),
        ("flux_back", ctypes.c_uint),
    ]

    @classmethod
    def default(cls):
        p = cls()
        lib().iblb_params_default(ctypes.byref(p))
        return p

    @classmethod
    def from_args(cls, args):
        """`args`: the words after `./app` (10-value form, or the scripts' legacy 8-value form)."""
        argv = [b"app"] + [str(a).encode() for a in args]
        arr = (ctypes.c_char_p * len(argv))(*argv)
        p = cls()
        _check(lib().iblb_params_from_args(len(argv), arr, ctypes.byref(p)))
        return p

    @classmethod
    def channel(cls, XDIM, YDIM, c_num=0, c_space=48, c_fraction=1, Re=1.0, T_num=1.0, T_pow=5,
                gx=0.0, gy=0.0, u0_amp=0.0, flags=0, device=-1, slab_x0=0, slab_w=0,
                LENGTH=0, fine_points=0, shape_scale=0.0, flux_back=0):
        """Parameter block for an arbitrary channel (C2/C4 style) with the reference's scaling.
        LENGTH / fine_points / shape_scale / flux_back: cilium geometry the reference hard-wires
        (96 / 9600 / 111 / 5); 0 keeps its value."""
        p = cls.default()
        if LENGTH:
            p.LENGTH = LENGTH
        p.fine_points, p.shape_scale, p.flux_back = fine_points, shape_scale, flux_back
        p.c_fraction, p.c_num, p.c_space = c_fraction, c_num, c_space
        p.Re, p.T_num, p.T_pow, p.I_pow, p.P_num = Re, T_num, T_pow, 1.0, 100
        p.XDIM, p.YDIM = XDIM, YDIM
        _check(lib().iblb_params_derive(ctypes.byref(p)))
        p.gx, p.gy, p.u0_amp, p.flags, p.device = gx, gy, u0_amp, flags, device
        p.slab_x0, p.slab_w = slab_x0, slab_w
        return p

    def derive(self):
        _check(lib().iblb_params_derive(ctypes.byref(self)))
        return self


class HaloHandle(ctypes.Structure):
    _fields_ = [
        ("mem", (ctypes.c_ubyte * IPC_HANDLE_BYTES) * 2),
        ("flag", ctypes.c_ubyte * IPC_HANDLE_BYTES),
        ("plane", ctypes.c_ulonglong),
        ("ld", ctypes.c_uint), ("ghost", ctypes.c_uint), ("width", ctypes.c_uint),
        ("device", ctypes.c_int), ("pid", ctypes.c_int),
        ("uuid", ctypes.c_ubyte * 16),
    ]


_lib = None

# every symbol include/iblb.h declares (tests check the library exports all of them)
SYMBOLS = [
    "iblb_params_default", "iblb_params_from_args", "iblb_params_derive", "iblb_create",
    "iblb_destroy", "iblb_step", "iblb_sync", "iblb_iteration", "iblb_get_fields",
    "iblb_get_force", "iblb_get_populations", "iblb_get_markers", "iblb_get_flux",
    "iblb_set_fields", "iblb_marker_fine_index", "iblb_last_step_seconds",
    "iblb_last_step_launches", "iblb_profile_phases", "iblb_export_halo", "iblb_connect",
    "iblb_connect_local", "iblb_set_marker_replay", "iblb_run", "iblb_last_error", "iblb_version",
    "iblb_save_checkpoint", "iblb_load_checkpoint", "iblb_debug_violations", "iblb_run_gpus",
    "iblb_device_count", "iblb_get_flux_history", "iblb_slab_range", "iblb_get_fields_static",
]


def lib():
    """Load libiblb.so (built in-tree by cuda_iblb_11_b200/build.py); no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise IblbError("libiblb.so not built: run `python -m cuda_iblb_11_b200.build` "
                        "(or __graft_entry__.build()); there is no CPU fallback")
    L = ctypes.CDLL(LIB_PATH)
    vp, dp = ctypes.c_void_p, ctypes.POINTER(ctypes.c_double)
    L.iblb_last_error.restype = ctypes.c_char_p
    L.iblb_version.restype = ctypes.c_char_p
    L.iblb_params_default.argtypes = [ctypes.POINTER(Params)]
    L.iblb_params_default.restype = None
    L.iblb_params_derive.argtypes = [ctypes.POINTER(Params)]
    L.iblb_params_from_args.argtypes = [ctypes.c_int, ctypes.POINTER(ctypes.c_char_p),
                                        ctypes.POINTER(Params)]
    L.iblb_create.argtypes = [ctypes.POINTER(Params), ctypes.POINTER(vp)]
    L.iblb_destroy.argtypes = [vp]
    L.iblb_step.argtypes = [vp, ctypes.c_uint]
    L.iblb_sync.argtypes = [vp]
    L.iblb_iteration.argtypes = [vp]
    L.iblb_iteration.restype = ctypes.c_uint
    L.iblb_get_fields.argtypes = [vp, vp, vp]
    L.iblb_get_fields_static.argtypes = [vp, vp, vp]
    L.iblb_get_force.argtypes = [vp, vp]
    L.iblb_get_populations.argtypes = [vp, vp]
    L.iblb_get_markers.argtypes = [vp, vp, vp, vp, vp]
    L.iblb_get_flux.argtypes = [vp, dp]
    L.iblb_set_fields.argtypes = [vp, vp, vp]
    L.iblb_get_flux_history.argtypes = [vp, ctypes.c_uint, ctypes.c_uint, vp]
    L.iblb_save_checkpoint.argtypes = [vp, ctypes.c_char_p]
    L.iblb_load_checkpoint.argtypes = [vp, ctypes.c_char_p]
    L.iblb_marker_fine_index.argtypes = [vp, ctypes.c_uint]
    L.iblb_last_step_seconds.argtypes = [vp, dp]
    L.iblb_last_step_launches.argtypes = [vp, ctypes.POINTER(ctypes.c_ulonglong)]
    L.iblb_profile_phases.argtypes = [vp, ctypes.c_uint, dp, dp, dp, dp]
    L.iblb_export_halo.argtypes = [vp, ctypes.POINTER(HaloHandle)]
    L.iblb_connect.argtypes = [vp, ctypes.POINTER(HaloHandle), ctypes.POINTER(HaloHandle)]
    L.iblb_connect_local.argtypes = [vp, vp, vp]
    L.iblb_set_marker_replay.argtypes = [vp, ctypes.c_uint, vp, ctypes.c_uint]
    L.iblb_run.argtypes = [ctypes.POINTER(Params), ctypes.c_char_p]
    L.iblb_debug_violations.restype = ctypes.c_longlong
    L.iblb_run_gpus.argtypes = [ctypes.POINTER(Params), ctypes.c_char_p, ctypes.c_int]
    _lib = L
    return L


def _check(rc):
    if rc != 0:
        raise IblbError("libiblb error %d: %s" % (rc, lib().iblb_last_error().decode()))


class Solver:
    """One IB-LBM run on one GPU (or one x-slab of a wider channel)."""

    def __init__(self, params):
        self.p = params
        self._h = ctypes.c_void_p()
        _check(lib().iblb_create(ctypes.byref(params), ctypes.byref(self._h)))
        self.width = params.slab_w if params.slab_w else params.XDIM
        self.size = self.width * params.YDIM
        self.Ns = (params.LENGTH or 96) * params.c_num

    # --- stepping -------------------------------------------------------------------------
    def step(self, n=1):
        _check(lib().iblb_step(self._h, n))

    def sync(self):
        _check(lib().iblb_sync(self._h))

    @property
    def it(self):
        return lib().iblb_iteration(self._h)

    # --- state ----------------------------------------------------------------------------
    def fields(self, out_rho=None, out_u=None):
        rho = np.empty(self.size) if out_rho is None else out_rho
        u = np.empty(2 * self.size) if out_u is None else out_u
        _check(lib().iblb_get_fields(self._h, rho.ctypes.data, u.ctypes.data))
        return rho, u

    def fields_static(self, out_rho=None, out_u=None):
        """fields() without any device allocation: staged through the library's fixed 6 MiB array
        (iblb_get_fields_static; grids up to IBLB_STATIC_FIELDS_MAX_NODES nodes, IblbError beyond)."""
        rho = np.empty(self.size) if out_rho is None else out_rho
        u = np.empty(2 * self.size) if out_u is None else out_u
        _check(lib().iblb_get_fields_static(self._h, rho.ctypes.data, u.ctypes.data))
        return rho, u

    def force(self):
        f = np.empty(2 * self.size)
        _check(lib().iblb_get_force(self._h, f.ctypes.data))
        return f

    def populations(self):
        f = np.empty(9 * self.size)
        _check(lib().iblb_get_populations(self._h, f.ctypes.data))
        return f

    def markers(self):
        n = max(self.Ns, 1)
        s = np.zeros(2 * n, np.float32)
        u_s = np.zeros(2 * n, np.float32)
        F_s = np.zeros(2 * n, np.float32)
        eps = np.zeros(n, np.int32)
        _check(lib().iblb_get_markers(self._h, s.ctypes.data, u_s.ctypes.data, F_s.ctypes.data,
                                      eps.ctypes.data))
        return s[:2 * self.Ns], u_s[:2 * self.Ns], F_s[:2 * self.Ns], eps[:self.Ns]

    @property
    def Q(self):
        q = ctypes.c_double()
        _check(lib().iblb_get_flux(self._h, ctypes.byref(q)))
        return q.value

    def flux_history(self, first_it, n):
        """Q after each of the iterations first_it .. first_it+n-1 (the reference's per-step Q,
        main.cu:928) from the device-side history: one copy, one synchronisation."""
        q = np.empty(n)
        _check(lib().iblb_get_flux_history(self._h, first_it, n, q.ctypes.data))
        return q

    def set_fields(self, rho, u):
        rho = np.ascontiguousarray(rho, np.float64)
        u = np.ascontiguousarray(u, np.float64)
        assert rho.size == self.size and u.size == 2 * self.size
        _check(lib().iblb_set_fields(self._h, rho.ctypes.data, u.ctypes.data))

    def save_checkpoint(self, path):
        """Complete time-stepping state to `path` (iblb_save_checkpoint)."""
        _check(lib().iblb_save_checkpoint(self._h, os.fsencode(path)))

    def load_checkpoint(self, path):
        """State of a solver with the same parameters from `path` (iblb_load_checkpoint)."""
        _check(lib().iblb_load_checkpoint(self._h, os.fsencode(path)))

    def marker_fine_index(self, j):
        return lib().iblb_marker_fine_index(self._h, j)

    def set_marker_replay(self, marker_j, vals):
        """Test hook (iblb_set_marker_replay): vals[it, m, 0:4] = (s_x, s_y, u_s_x, u_s_y) of
        marker `marker_j` of cilium m at iteration it, as a recorded reference run produced them."""
        if vals is None:
            _check(lib().iblb_set_marker_replay(self._h, marker_j, None, 0))
            return
        v = np.ascontiguousarray(vals, np.float32)
        assert v.ndim == 3 and v.shape[1] == self.p.c_num and v.shape[2] == 4
        _check(lib().iblb_set_marker_replay(self._h, marker_j, v.ctypes.data, v.shape[0]))

    # --- measurement ----------------------------------------------------------------------
    @property
    def last_step_seconds(self):
        t = ctypes.c_double()
        _check(lib().iblb_last_step_seconds(self._h, ctypes.byref(t)))
        return t.value

    @property
    def last_step_launches(self):
        n = ctypes.c_ulonglong()
        _check(lib().iblb_last_step_launches(self._h, ctypes.byref(n)))
        return n.value

    def profile_phases(self, reps=10):
        t = [ctypes.c_double() for _ in range(4)]
        _check(lib().iblb_profile_phases(self._h, reps, *[ctypes.byref(x) for x in t]))
        return dict(kinematics=t[0].value, interpolate=t[1].value, spread=t[2].value,
                    lb=t[3].value)

    # --- slab coupling --------------------------------------------------------------------
    def export_halo(self):
        h = HaloHandle()
        _check(lib().iblb_export_halo(self._h, ctypes.byref(h)))
        return h

    def connect(self, left, right):
        _check(lib().iblb_connect(self._h, ctypes.byref(left), ctypes.byref(right)))

    def connect_local(self, left, right):
        _check(lib().iblb_connect_local(self._h, left._h, right._h))

    def close(self):
        if self._h:
            lib().iblb_destroy(self._h)
            self._h = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def debug_violations():
    """Index violations counted by an IBLB_CHECKED build of libiblb.so; -1 for a product build."""
    return lib().iblb_debug_violations()


def run(params, root=None, gpus=None):
    """The whole reference program (main.cu) for `params`: steps + Data/ tree under `root`;
    `gpus`: x-slabs over that many GPUs of this process (None: IBLB_GPUS or 1)."""
    r = None if root is None else str(root).encode()
    if gpus is None:
        _check(lib().iblb_run(ctypes.byref(params), r))
    else:
        _check(lib().iblb_run_gpus(ctypes.byref(params), r, gpus))
