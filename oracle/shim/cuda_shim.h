/* TEST INFRASTRUCTURE — not part of the product path.
 *
 * Host-side stand-in for the handful of CUDA language extensions the reference kernels use
 * (CUDA_IBLB_11/LatticeBoltzmann.cu, ImmersedBoundary.cu, main.cu:56-252), so that those files
 * can be compiled UNMODIFIED, where they lie under /root/reference, as plain host C++ by g++.
 * The three header names the reference includes ("cuda_runtime.h",
 * "device_launch_parameters.h", "device_functions.h") forward to this file.
 *
 * Execution model: a kernel "launch" is a loop over blocks (OpenMP across blocks) and, inside a
 * block, either a plain loop over threads (kernels whose only barrier is the trailing one) or
 * one ucontext fiber per thread with round-robin scheduling, where __syncthreads() yields — which
 * gives true block-barrier semantics for `spread` (ImmersedBoundary.cu:180-233) and
 * `boundary_check` (main.cu:214).  See oracle/ref_cpu.cpp for the launcher.
 */
#ifndef IBLB_ORACLE_CUDA_SHIM_H
#define IBLB_ORACLE_CUDA_SHIM_H

#include <cmath>
#include <cstdlib>
#include <cstring>

#define __global__
#define __device__
#define __host__
#define __constant__
#define __shared__ static thread_local

struct shim_dim3 {
    unsigned x, y, z;
    shim_dim3(unsigned x_ = 1, unsigned y_ = 1, unsigned z_ = 1) : x(x_), y(y_), z(z_) {}
};
typedef shim_dim3 dim3;

extern thread_local shim_dim3 threadIdx, blockIdx, blockDim, gridDim;

/* yields to the block scheduler when the launcher runs the block as fibers, else a no-op */
void __syncthreads();

/* the reference files call the unqualified names with float/double arguments; in device code
 * CUDA resolves them to the float/double overloads, on the host that needs the std:: set. */
using std::abs;
using std::sqrt;
using std::nearbyint;
using std::pow;
using std::cos;
using std::sin;

static inline long long __double_as_longlong(double v) {
    long long r;
    std::memcpy(&r, &v, sizeof r);
    return r;
}
static inline double __longlong_as_double(long long v) {
    double r;
    std::memcpy(&r, &v, sizeof r);
    return r;
}
static inline unsigned long long atomicCAS(unsigned long long *addr, unsigned long long cmp,
                                           unsigned long long val) {
    return __sync_val_compare_and_swap(addr, cmp, val);
}

#endif
