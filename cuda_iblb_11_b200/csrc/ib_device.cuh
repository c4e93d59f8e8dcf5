// Device helpers of the immersed-boundary side, shared by the stand-alone IB kernels
// (ib_kernels.cu) and the fused IB+LB step (fused_step.cu).  Every function here reproduces a
// specific piece of the reference with its float/double rounding points (SURVEY Q3,Q6,Q14,Q15).
#pragma once
#include "iblb_device.cuh"

namespace iblb {

// beat-shape coefficient tables, main.cu:56-74 ("without mucus" set): data, n fastest
__device__ __constant__ const double kA[7 * 2 * 3] = {
    -0.654, 0.393,  -0.097, 0.079,  0.119,  0.119,  0.009,  1.895,  -0.018, 0.158,  0.010,
    0.003,  0.013,  0.040,  0.787,  -1.516, 0.032,  -0.302, -0.252, -0.015, 0.035,  -0.552,
    -0.126, -0.341, 0.035,  0.006,  -0.029, -0.068, 0.202,  0.716,  -0.118, 0.142,  0.110,
    -0.013, -0.043, 0.096,  0.263,  0.186,  -0.067, -0.032, -0.002, 0.015};
__device__ __constant__ const double kB[7 * 2 * 3] = {
    0.0,    0.284, 0.006,  -0.059, 0.018,  0.053,  0.009,  0.0,    0.192,  -0.050, 0.012,
    -0.007, -0.014, -0.017, 0.0,   1.045,  0.317,  0.226,  0.004,  -0.082, -0.040, 0.0,
    -0.499, 0.423, 0.138,  0.125,  0.075,  0.067,  0.0,    -1.017, -0.276, -0.196, -0.037,
    0.025,  0.023, 0.0,    0.339,  -0.327, -0.114, -0.105, -0.057, -0.055};

// main.cu:36-54, the "WITH MUCUS PRESENT" set the reference keeps commented out (IBLB_FLAG_MUCUS)
__device__ __constant__ const double kAmucus[7 * 2 * 3] = {
    -0.449, 0.130,  -0.169, 0.063,  -0.050, -0.040, -0.068, 2.076,  -0.003, 0.054,  0.007,
    0.026,  0.022,  0.010,  -0.072, -1.502, 0.260,  -0.123, 0.011,  -0.009, 0.196,  -1.074,
    -0.230, -0.305, -0.180, -0.069, 0.001,  -0.080, 0.658,  0.793,  -0.251, 0.049,  0.009,
    0.023,  -0.111, 0.381,  0.331,  0.193,  0.082,  0.029,  0.002,  0.048};
__device__ __constant__ const double kBmucus[7 * 2 * 3] = {
    0.0,   -0.030, -0.093, 0.037,  0.062,  0.016,  -0.065, 0.0,   0.080,  -0.044, -0.017,
    0.052, 0.007,  0.051,  0.0,    1.285,  -0.036, -0.244, -0.093, -0.137, 0.095,  0.0,
    -0.298, 0.513, 0.004,  -0.222, 0.035,  -0.128, 0.0,    -1.034, 0.050,  0.143,  0.043,
    0.098, -0.054, 0.0,    0.210,  -0.367, 0.009,  0.120,  -0.024, 0.102};

#define IBLB_PI_REF 3.14159  // main.cu:29

namespace {

// main.cu:102-103
__device__ __forceinline__ int beat_phase(int it, int m, int p_step, int T) {
    return (it + m * p_step == T) ? T : (it + m * p_step) % T;
}

// main.cu:137-145 for one marker; cs[n-1] = (cos, sin)(n*2*PI*phase/T).  Explicit intrinsics: the
// same bits in every kernel that inlines it (see lb_device.cuh).
__device__ __forceinline__ float2 marker_pos(const float *cf, const double2 *cs, float offset,
                                             double scale) {
    float sx = __dadd_rn(__dmul_rn(__dmul_rn(scale, (double)cf[0]), 0.5), (double)offset);
    float sy = __dmul_rn(__dmul_rn(scale, (double)cf[1]), 0.5);
#pragma unroll
    for (int n = 1; n < 7; n++) {
        const double tx = __fma_rn((double)cf[2 * n + 0], cs[n - 1].x,
                                   __dmul_rn((double)cf[14 + 2 * n + 0], cs[n - 1].y));
        const double ty = __fma_rn((double)cf[2 * n + 1], cs[n - 1].x,
                                   __dmul_rn((double)cf[14 + 2 * n + 1], cs[n - 1].y));
        sx = __dadd_rn((double)sx, __dmul_rn(scale, tx));
        sy = __dadd_rn((double)sy, __dmul_rn(scale, ty));
    }
    return make_float2(sx, sy);
}

// trig table of one cilium at iteration `it`: t in [0,12): n = t/2+1, even t -> cos, odd -> sin of
// n*2*PI*phase/T evaluated left to right in double as main.cu:143-144
__device__ __forceinline__ double beat_trig(int t, int it, int m, int p_step, int T) {
    const int n = t / 2 + 1;
    const int phase = beat_phase(it, m, p_step, T);
    const double arg = n * 2. * IBLB_PI_REF * phase / T;
    return (t & 1) ? sin(arg) : cos(arg);
}

// position (translated + wrapped, main.cu:193-198) and velocity (main.cu:147-151,200-209) of one
// marker of cilium m; cs0/cs1: trig tables of iterations it and it-1
__device__ __forceinline__ void marker_state(const float *cf, const double2 *cs0,
                                             const double2 *cs1, int m, int it, int c_space_i,
                                             int c_num_i, int XDIM, double scale, float2 &s,
                                             float2 &v) {
    const double c_space = c_space_i, c_num = c_num_i;
    const float offset = 1. * (m - (c_num - 1) / 2.) * c_space;  // main.cu:105
    const float2 b = marker_pos(cf, cs0, offset, scale);
    v = make_float2(0.f, 0.f);
    if (it > 0) {
        const float2 l = marker_pos(cf, cs1, offset, scale);
        v.x = b.x - l.x;
        v.y = b.y - l.y;
    }
    float sx = (c_space * c_num) / 2. + b.x;
    if (sx < 0) sx += XDIM;
    else if (sx > XDIM) sx -= XDIM;
    s = make_float2(sx, b.y + 1);
}

// one factor of the 3-point delta, ImmersedBoundary.cu:21-81 (truncated prefactors, float |r|,
// double body, float result); explicit intrinsics for kernel-independent bits
__device__ __forceinline__ float delta_1d(float r) {
    if (!(r <= 1.5)) return 0.f;
    const double rd = r;
    if (r <= 0.5) {  // 0.33333 * (1 + sqrt(1 - 3 r^2))
        const double q = sqrt(__fma_rn(__dmul_rn(-3., rd), rd, 1.));
        return __dmul_rn(0.33333, __dadd_rn(1., q));
    }
    // 0.16667 * (5 - 3 r - sqrt(1 - 3 (1 - r)^2))
    const double d = 1 - r;  // float subtraction, as ImmersedBoundary.cu:46 (`d = 1-dx`)
    const double q = sqrt(__fma_rn(__dmul_rn(-3., d), d, 1.));
    return __dmul_rn(0.16667, __dsub_rn(__fma_rn(-3., rd, 5.), q));
}

// rho and the uncorrected velocity u* (LatticeBoltzmann.cu:396-405, summation order i = 0..8) of
// the node the reference's interpolate addresses as (x, y): it indexes j = y*XDIM + x WITHOUT
// wrapping (ImmersedBoundary.cu:119-122), so x = -1 lands on (XDIM-1, y-1) and x = XDIM on
// (0, y+1).  The node's nine populations are pulled out of the post-collision buffer `src` (pull
// form of LatticeBoltzmann.cu:173-373 incl. the wall rules).  False when j lies outside the array
// or the node's source columns are not held by this slab: such a node contributes nothing.
// (offsets first, loads second: lb_ib_tile_step forms the offsets before it waits for the previous
// iteration — they do not depend on the fluid — and issues only the nine loads after)
__device__ __forceinline__ bool support_node_offsets(const Lattice &L, int x, int y,
                                                     long long (&o)[9]) {
    const long long j = (long long)y * L.XDIM + x;
    if (j < 0 || j >= (long long)L.XDIM * L.YDIM) return false;
    int ya = y, xa = x;
    if ((unsigned)x >= (unsigned)L.XDIM) {  // off the row: the division only here
        ya = (int)(j / L.XDIM);
        xa = (int)(j - (long long)ya * L.XDIM);
    }
    // global column -> a local column whose three source columns are present
    int xl = xa - L.x0;
    if (xl < -(L.G - 1)) xl += L.XDIM;
    else if (xl > L.W + L.G - 2) xl -= L.XDIM;
    if (xl < -(L.G - 1) || xl > L.W + L.G - 2) return false;
#pragma unroll
    for (int q = 0; q < 9; q++) {
        int pi = q, xsrc = xl - kCx[q], ysrc = ya - kCy[q];
        if (ysrc < 0) {
            pi = (q == 2) ? 4 : (q == 5) ? 7 : 8;
            xsrc = xl; ysrc = ya;
        } else if (ysrc >= L.YDIM) {
            pi = (q == 4) ? 2 : (q == 7) ? 6 : 5;
            xsrc = xl; ysrc = ya;
        }
        o[q] = pi * L.plane + L.off(xsrc, ysrc);
    }
    return true;
}

__device__ __forceinline__ bool support_node_pull(const Lattice &L, const double *__restrict__ src,
                                                  int x, int y, double (&f)[9]) {
    long long o[9];
    if (!support_node_offsets(L, x, y, o)) return false;
#pragma unroll
    for (int q = 0; q < 9; q++) f[q] = src[o[q]];
    return true;
}

__device__ __forceinline__ bool support_node_moments(const Lattice &L,
                                                     const double *__restrict__ src, int x, int y,
                                                     double &rho, double &ux, double &uy) {
    double f[9];
    if (!support_node_pull(L, src, x, y, f)) return false;
    rho = f[0] + f[1] + f[2] + f[3] + f[4] + f[5] + f[6] + f[7] + f[8];
    ux = (f[1] - f[3] + f[5] - f[6] - f[7] + f[8]) / rho;
    uy = (f[2] - f[4] + f[5] + f[6] - f[7] - f[8]) / rho;
    return true;
}

// 2*delta*rho*(u_s - u), ImmersedBoundary.cu:126-127 evaluated left to right in double; `del` is
// the float product of the two 1-D delta factors (ImmersedBoundary.cu:78).  Explicit intrinsics:
// never contracted with the accumulation that follows, whatever kernel inlines it.
__device__ __forceinline__ double interp_product(float del, double rho, float us, double u) {
    return __dmul_rn(__dmul_rn(__dmul_rn(2., (double)del), rho), __dsub_rn((double)us, u));
}

// Term i (0..8) of the interpolation sum of marker (sm, us): 2*delta*rho*(u_s - u) at support node
// i (ImmersedBoundary.cu:117-128).
__device__ __forceinline__ void interp_term(const Lattice &L, const double *__restrict__ src,
                                            float2 sm, float2 us, int i, double &tx, double &ty) {
    tx = 0.;
    ty = 0.;
    const double xs = sm.x, ys = sm.y;
    const int x0 = (int)nearbyint(xs), y0 = (int)nearbyint(ys);
    const int x = x0 + kCx[i], y = y0 + kCy[i];
    double rho, ux, uy;
    if (!support_node_moments(L, src, x, y, rho, ux, uy)) return;
    const float dlx = delta_1d(fabsf(x - sm.x));
    const float dly = delta_1d(fabsf(y - sm.y));
    const float del = dlx * dly;  // float product (ImmersedBoundary.cu:78)
    tx = interp_product(del, rho, us.x, ux);
    ty = interp_product(del, rho, us.y, uy);
}

}  // namespace
}  // namespace iblb
