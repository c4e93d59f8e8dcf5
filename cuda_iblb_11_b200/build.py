"""In-tree build of libiblb.so (C ABI, include/iblb.h) and the `app` driver with nvcc for sm_100a.

No torch involved: the product is a plain C-ABI shared library.  The built files stay in the tree
(git-ignored) so that they travel to the GPU box with the snapshot.
"""
import os
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
LIB = os.path.join(PKG, "libiblb.so")
APP = os.path.join(ROOT, "app")

CUDA_SOURCES = ["lb_kernels.cu", "ib_kernels.cu", "fused_step.cu", "solver.cu", "step.cu",
                "slab.cu", "checkpoint.cu"]
HOST_SOURCES = ["params.cpp", "run.cpp"]
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _run(cmd):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("build failed: %s\n%s" % (" ".join(cmd), r.stdout))
    return r.stdout


def build_checked(verbose=False):
    """Development twin of the library with -DIBLB_CHECKED (device-side range checks on every index,
    iblb_debug_violations() reports the count): cuda_iblb_11_b200/libiblb_checked.so, loaded instead
    of libiblb.so when IBLB_LIB points at it (tools/sanitize_cases.py)."""
    out = os.path.join(PKG, "libiblb_checked.so")
    srcs = [os.path.join(CSRC, s) for s in CUDA_SOURCES + HOST_SOURCES]
    cmd = ["nvcc", "-DIBLB_CHECKED", *ARCH, "-O3", "-lineinfo", "-std=c++17", "-shared", "-Xcompiler",
           "-fPIC", "-I", os.path.join(ROOT, "include"), "-I", CSRC, "-o", out, *srcs, "-lcudart"]
    log = _run(cmd)
    if verbose:
        print(log)
    return out


def build(force=False, verbose=False):
    """Compile libiblb.so and app if any source is newer.  Returns the library path."""
    headers = [os.path.join(CSRC, h) for h in ("iblb_device.cuh", "iblb_internal.h",
                                               "lb_device.cuh", "ib_device.cuh")]
    headers.append(os.path.join(ROOT, "include", "iblb.h"))
    srcs = [os.path.join(CSRC, s) for s in CUDA_SOURCES + HOST_SOURCES]
    if force or _newer(LIB, srcs + headers):
        cmd = ["nvcc", *ARCH, "-O3", "-lineinfo", "-std=c++17", "-shared", "-Xcompiler", "-fPIC",
               "-Xptxas", "-v", "-I", os.path.join(ROOT, "include"), "-I", CSRC,
               "-o", LIB, *srcs, "-lcudart"]
        out = _run(cmd)
        if verbose:
            print(out)
    app_src = os.path.join(CSRC, "app_main.cpp")
    if force or _newer(APP, [app_src, LIB]):
        _run(["g++", "-O2", "-std=c++17", "-I", os.path.join(ROOT, "include"), app_src,
              "-o", APP, "-L", PKG, "-liblb", "-Wl,-rpath," + PKG])
    return LIB


if __name__ == "__main__":
    if "--checked" in sys.argv:
        print(build_checked(verbose=True))
    else:
        print(build(force="--force" in sys.argv, verbose=True))
