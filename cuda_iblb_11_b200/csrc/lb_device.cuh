// Device helpers shared by the LB kernels (lb_kernels.cu) and the fused IB+LB step (fused_step.cu):
// moments, TRT collision with Guo forcing, vector load/shuffle helpers.  sm_100a only.
#pragma once
#include "iblb_device.cuh"

namespace iblb {
namespace {

constexpr double kIcs2 = 1.0 / (kCs * kCs);                     // 1/cs^2
constexpr double kIcs4 = 1.0 / (kCs * kCs * kCs * kCs);         // 1/cs^4
constexpr double kIcs4h = 1.0 / (2.0 * kCs * kCs * kCs * kCs);  // 1/(2 cs^4)
constexpr double kIcs2h = 1.0 / (2.0 * kCs * kCs);              // 1/(2 cs^2)
constexpr double kW0 = 4.0 / 9.0, kW1 = 1.0 / 9.0, kW2 = 1.0 / 36.0;

// reflection partners used by the wall rules (pull form)
//   y == 0      : f2 <- f4, f5 <- f7, f6 <- f8   (same node)
//   y == YDIM-1 : f4 <- f2, f7 <- f6, f8 <- f5   (same node)

__device__ __forceinline__ double shfl_up1(double v) { return __shfl_up_sync(0xffffffffu, v, 1); }
__device__ __forceinline__ double shfl_dn1(double v) {
    return __shfl_down_sync(0xffffffffu, v, 1);
}

__device__ __forceinline__ double2 ldg2(const double *p) {
    return __ldg(reinterpret_cast<const double2 *>(p));
}

// The arithmetic below is written with explicit round-to-nearest intrinsics so that every kernel
// that inlines it (lb_fused_v1/v2, the fused IB+LB step, the band kernel) produces the same bits:
// left to the compiler, which product is contracted into which sum depends on the inlining context.
#define DM(a, b) __dmul_rn((a), (b))
#define DA(a, b) __dadd_rn((a), (b))
#define DS(a, b) __dsub_rn((a), (b))
#define DF(a, b, c) __fma_rn((a), (b), (c))

// TRT relaxation of one antiparallel pair (a = +c, b = -c) plus the Guo term.
// LatticeBoltzmann.cu:88-134 with f0 split into its even/odd parts analytically.  The FP64 pipe is
// half busy at HBM speed (the forced C2 step is 3.7 % slower than the unforced one), so the
// operation count matters:
//   * 0.5*(fa +- fb) - f0 is one FMA (0.5*x is exact: same bits as multiply-then-subtract);
//   * the Guo term F_(+-c) = kw [ +-(c.F)/cs^2 - (u.F)/cs^2 + (c.u)(c.F)/cs^4 ] is split into its
//     even part S and odd part D, which join the even / odd relaxation terms before they meet the
//     populations: new_a = fa + (S - w+ d+) + (D - w- d-), new_b = fb + (S - w+ d+) - (D - w- d-).
// kI4 = kw/cs^4, kI2 = kw/cs^2 per weight class; kI2uF = kI2 (u.F) per node.
// sab = fa + fb and dab = fa - fb come from the caller, which needs them for the moments as well.
__device__ __forceinline__ void relax_pair(double &fa, double &fb, double sab, double dab,
                                           double rw, double cu, double base, double wp,
                                           double wm, bool forced, double cF, double kI4,
                                           double kI2, double kI2uF) {
    const double f0p = DM(rw, DF(DM(cu, cu), kIcs4h, base));
    const double f0m = DM(rw, DM(cu, kIcs2));
    const double dp = DF(0.5, sab, -f0p);
    const double dm = DF(0.5, dab, -f0m);
    if (forced) {
        const double A = DF(-wp, dp, DF(DM(cu, cF), kI4, -kI2uF));
        const double B = DF(-wm, dm, DM(cF, kI2));
        const double ta = DA(fa, A), tb = DA(fb, A);
        fa = DA(ta, B);
        fb = DS(tb, B);
    } else {
        const double na = DF(-wm, dm, DF(-wp, dp, fa));
        const double nb = DF(wm, dm, DF(-wp, dp, fb));
        fa = na;
        fb = nb;
    }
}

// moments (LatticeBoltzmann.cu:396-405, summation order i = 0..8), half-force velocity
// (ImmersedBoundary.cu:249-255), then equilibrium + forcing + TRT.  Returns u_x.
__device__ __forceinline__ double collide_node(double (&f)[9], double Fx, double Fy, bool forced,
                                               const LbParams &p) {
    // sums and differences of the four antiparallel pairs serve the moments and the relaxation
    // (the order of LatticeBoltzmann.cu:396-405 is kept where it decides a float rounding: the
    // interpolation's moments in ib_device.cuh / fused_step.cu; here the result feeds FP64 only)
    const double s13 = DA(f[1], f[3]), d13 = DS(f[1], f[3]);
    const double s24 = DA(f[2], f[4]), d24 = DS(f[2], f[4]);
    const double s57 = DA(f[5], f[7]), d57 = DS(f[5], f[7]);
    const double s68 = DA(f[6], f[8]), d68 = DS(f[6], f[8]);
    const double rho = DA(DA(f[0], DA(s13, s24)), DA(s57, s68));
    const double mx = DS(DA(d13, d57), d68);
    const double my = DA(DA(d24, d57), d68);
    const double ir = __drcp_rn(rho);
    const double ux = DM(DF(p.half, Fx, mx), ir);
    const double uy = DM(DF(p.half, Fy, my), ir);
    const double base = DF(-DF(uy, uy, DM(ux, ux)), kIcs2h, 1.0);
    const double wp = p.omega_plus, wm = p.omega_minus;
    const double uF = DF(uy, Fy, DM(ux, Fx));
    const double u1 = DM(p.k1I2, uF), u2 = DM(p.k2I2, uF);
    const double r1 = DM(rho, kW1), r2 = DM(rho, kW2);
    // rest population: no force term (LatticeBoltzmann.cu:86)
    f[0] = DF(-wp, DS(f[0], DM(DM(rho, kW0), base)), f[0]);
    relax_pair(f[1], f[3], s13, d13, r1, ux, base, wp, wm, forced, Fx, p.k1I4, p.k1I2, u1);
    relax_pair(f[2], f[4], s24, d24, r1, uy, base, wp, wm, forced, Fy, p.k1I4, p.k1I2, u1);
    relax_pair(f[5], f[7], s57, d57, r2, DA(ux, uy), base, wp, wm, forced, DA(Fx, Fy), p.k2I4, p.k2I2, u2);
    relax_pair(f[6], f[8], s68, d68, r2, DS(uy, ux), base, wp, wm, forced, DS(Fy, Fx), p.k2I4, p.k2I2, u2);
    return ux;
}

// running flux sum of one 64-row slot (ImmersedBoundary.cu:259-264: Q is cumulative), and its value
// after this iteration into the history ring (the reference copies Q back every step, main.cu:928;
// here a caller reads a whole interval at once: iblb_get_flux_history)
__device__ __forceinline__ void flux_add(const LbParams &p, int slot, double v) {
    const double q = p.qacc[slot] + v;
    p.qacc[slot] = q;
    if (p.qhist != nullptr) {
        const int it = (p.it_ptr ? __ldcg(p.it_ptr) : 0) + p.it;
        p.qhist[(long long)(it % kFluxHist) * p.nq + slot] = q;
    }
}

__device__ __forceinline__ void store_remote2(const Remote &r, int col, int y0, int i, double2 v) {
    IBLB_CHK(col + r.G >= 0 && col + r.G < r.W + 2 * r.G && y0 >= 0 && y0 + 1 < r.ld && i >= 0 && i < 9);
    *reinterpret_cast<double2 *>(r.base + (long long)i * r.plane + (long long)(col + r.G) * r.ld +
                                 y0) = v;
}


// The nine populations two vertically adjacent nodes (y0, y0+1) of column x pull, as loaded:
// one aligned 128-bit load per population from the column it streams from, plus the values from
// the row below lane 0 / above lane 31 of the warp (walls folded in).
struct Pulled {
    double2 v[9];
    double e2, e5, e6, e4, e7, e8;
    bool bot, top;  // lane 0 sits on the floor / lane 31 under the ceiling: e2 / e4 are v[4].x / v[2].y
};

__device__ __forceinline__ void load_pulled_at(const Lattice &L, const double *__restrict__ src,
                                               long long cm, long long c0, long long cp, int y0,
                                               int lane, Pulled &q);
__device__ __forceinline__ void load_pulled(const Lattice &L, const double *__restrict__ src, int x,
                                            int y0, int lane, Pulled &q) {
    load_pulled_at(L, src, L.off(x - 1, y0), L.off(x, y0), L.off(x + 1, y0), y0, lane, q);
}
// cm, c0, cp: offsets of (x-1, y0), (x, y0), (x+1, y0) within a population plane (a caller that has
// to wait for the previous iteration forms them before it waits)
__device__ __forceinline__ void load_pulled_at(const Lattice &L, const double *__restrict__ src,
                                               long long cm, long long c0, long long cp, int y0,
                                               int lane, Pulled &q) {
    const long long P = L.plane;
    const bool at_bottom = (y0 == 0);
    const bool at_top = (y0 + 2 == L.YDIM);
    q.v[0] = ldg2(src + 0 * P + c0);
    q.v[1] = ldg2(src + 1 * P + cm);  // c = (+1, 0): from x-1
    q.v[2] = ldg2(src + 2 * P + c0);  // c = ( 0,+1): from y-1
    q.v[3] = ldg2(src + 3 * P + cp);  // c = (-1, 0): from x+1
    q.v[4] = ldg2(src + 4 * P + c0);  // c = ( 0,-1): from y+1
    q.v[5] = ldg2(src + 5 * P + cm);  // c = (+1,+1)
    q.v[6] = ldg2(src + 6 * P + cp);  // c = (-1,+1)
    q.v[7] = ldg2(src + 7 * P + cp);  // c = (-1,-1)
    q.v[8] = ldg2(src + 8 * P + cm);  // c = (+1,-1)
    q.e2 = q.e5 = q.e6 = q.e4 = q.e7 = q.e8 = 0.;
    // f2(x,0) <- f4(x,0) and f4(x,Y-1) <- f2(x,Y-1) are values loaded above: they are picked in
    // shift_pulled, NOT here — touching v[] now would stall the issue of the loads that follow
    // (edge values, the IB ring) for a whole memory latency
    q.bot = at_bottom;
    q.top = at_top;
    if (lane == 0) {
        if (at_bottom) {
            q.e5 = __ldg(src + 7 * P + c0);
            q.e6 = __ldg(src + 8 * P + c0);
        } else {
            q.e2 = __ldg(src + 2 * P + c0 - 1);
            q.e5 = __ldg(src + 5 * P + cm - 1);
            q.e6 = __ldg(src + 6 * P + cp - 1);
        }
    }
    if (lane == 31) {
        if (at_top) {
            q.e7 = __ldg(src + 6 * P + c0 + 1);
            q.e8 = __ldg(src + 5 * P + c0 + 1);
        } else {
            q.e4 = __ldg(src + 4 * P + c0 + 2);
            q.e7 = __ldg(src + 7 * P + cp + 2);
            q.e8 = __ldg(src + 8 * P + cm + 2);
        }
    }
}

// shift the six populations with c_y != 0 by one row in registers: a = node y0, b = node y0+1
__device__ __forceinline__ void shift_pulled(const Pulled &q, int lane, double (&a)[9],
                                             double (&b)[9]) {
    a[0] = q.v[0].x; b[0] = q.v[0].y;
    a[1] = q.v[1].x; b[1] = q.v[1].y;
    a[3] = q.v[3].x; b[3] = q.v[3].y;
    double t;
    t = shfl_up1(q.v[2].y); a[2] = lane == 0 ? (q.bot ? q.v[4].x : q.e2) : t; b[2] = q.v[2].x;
    t = shfl_up1(q.v[5].y); a[5] = lane == 0 ? q.e5 : t; b[5] = q.v[5].x;
    t = shfl_up1(q.v[6].y); a[6] = lane == 0 ? q.e6 : t; b[6] = q.v[6].x;
    t = shfl_dn1(q.v[4].x); a[4] = q.v[4].y; b[4] = lane == 31 ? (q.top ? q.v[2].y : q.e4) : t;
    t = shfl_dn1(q.v[7].x); a[7] = q.v[7].y; b[7] = lane == 31 ? q.e7 : t;
    t = shfl_dn1(q.v[8].x); a[8] = q.v[8].y; b[8] = lane == 31 ? q.e8 : t;
}

// stores of one thread's two nodes: own buffer + (boundary columns) the neighbours' ghost columns
__device__ __forceinline__ void store_pair(const LbParams &p, int x, int y0, const double (&a)[9],
                                           const double (&b)[9]) {
    const Lattice &L = p.L;
    const long long c0 = L.off(x, y0);
    double *__restrict__ dst = p.dst;
#pragma unroll
    for (int i = 0; i < 9; i++)
        *reinterpret_cast<double2 *>(dst + i * L.plane + c0) = make_double2(a[i], b[i]);
    // Ghost columns: with one ghost column per side (pure LB) only the three populations that
    // stream across the face are ever pulled from it — c_x = -1 (3, 6, 7) into the left
    // neighbour's right ghost, c_x = +1 (1, 5, 8) into the right neighbour's left ghost; with
    // cilia (G = 3) the neighbour's interpolation forms moments of ghost nodes and needs all nine.
    if (x < p.left.G) {
        const bool all = p.left.G > 1;
#pragma unroll
        for (int i = 0; i < 9; i++)
            if (all || i == 3 || i == 6 || i == 7)
                store_remote2(p.left, p.left.W + x, y0, i, make_double2(a[i], b[i]));
    }
    if (x >= L.W - p.right.G) {
        const bool all = p.right.G > 1;
#pragma unroll
        for (int i = 0; i < 9; i++)
            if (all || i == 1 || i == 5 || i == 8)
                store_remote2(p.right, x - L.W, y0, i, make_double2(a[i], b[i]));
    }
}

// ---- four vertically adjacent nodes per thread (lb_fused_v4): 256-bit loads and stores ----------
struct __align__(32) double4v {
    double x, y, z, w;
};
__device__ __forceinline__ double4v ldg4(const double *p) {  // LDG.E.256, read-only path
    double4v v;
    asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];"
                 : "=d"(v.x), "=d"(v.y), "=d"(v.z), "=d"(v.w)
                 : "l"(p));
    return v;
}
__device__ __forceinline__ void stg4(double *p, double a, double b, double c, double d) {
    asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(a), "d"(b), "d"(c), "d"(d)
                 : "memory");
}

// The nine populations the nodes (y0 .. y0+3) of column x pull, as loaded: one aligned 256-bit load
// per population from the column it streams from, plus the row below lane 0 / above lane 31
// (walls folded in exactly as in Pulled / load_pulled).
struct Pulled4 {
    double4v v[9];
    double e2, e5, e6, e4, e7, e8;
    bool bot, top;
};

__device__ __forceinline__ void load_pulled4(const Lattice &L, const double *__restrict__ src, int x,
                                             int y0, int lane, Pulled4 &q) {
    const long long P = L.plane;
    const long long cm = L.off(x - 1, y0), c0 = L.off(x, y0), cp = L.off(x + 1, y0);
    q.bot = (y0 == 0);
    q.top = (y0 + 4 == L.YDIM);
    q.v[0] = ldg4(src + 0 * P + c0);
    q.v[1] = ldg4(src + 1 * P + cm);
    q.v[2] = ldg4(src + 2 * P + c0);
    q.v[3] = ldg4(src + 3 * P + cp);
    q.v[4] = ldg4(src + 4 * P + c0);
    q.v[5] = ldg4(src + 5 * P + cm);
    q.v[6] = ldg4(src + 6 * P + cp);
    q.v[7] = ldg4(src + 7 * P + cp);
    q.v[8] = ldg4(src + 8 * P + cm);
    q.e2 = q.e5 = q.e6 = q.e4 = q.e7 = q.e8 = 0.;
    if (lane == 0) {
        if (q.bot) {
            q.e5 = __ldg(src + 7 * P + c0);
            q.e6 = __ldg(src + 8 * P + c0);
        } else {
            q.e2 = __ldg(src + 2 * P + c0 - 1);
            q.e5 = __ldg(src + 5 * P + cm - 1);
            q.e6 = __ldg(src + 6 * P + cp - 1);
        }
    }
    if (lane == 31) {
        if (q.top) {
            q.e7 = __ldg(src + 6 * P + c0 + 3);
            q.e8 = __ldg(src + 5 * P + c0 + 3);
        } else {
            q.e4 = __ldg(src + 4 * P + c0 + 4);
            q.e7 = __ldg(src + 7 * P + cp + 4);
            q.e8 = __ldg(src + 8 * P + cm + 4);
        }
    }
}

// n0..n3 = nodes y0..y0+3: the six populations with c_y != 0 shifted by one row in registers
__device__ __forceinline__ void shift_pulled4(const Pulled4 &q, int lane, double (&n0)[9],
                                              double (&n1)[9], double (&n2)[9], double (&n3)[9]) {
#pragma unroll
    for (int i = 0; i < 9; i++) {
        if (i == 0 || i == 1 || i == 3) {
            n0[i] = q.v[i].x; n1[i] = q.v[i].y; n2[i] = q.v[i].z; n3[i] = q.v[i].w;
        } else if (i == 2 || i == 5 || i == 6) {  // c_y = +1: from the row below
            const double t = shfl_up1(q.v[i].w);
            const double e = i == 2 ? (q.bot ? q.v[4].x : q.e2) : i == 5 ? q.e5 : q.e6;
            n0[i] = lane == 0 ? e : t;
            n1[i] = q.v[i].x; n2[i] = q.v[i].y; n3[i] = q.v[i].z;
        } else {  // c_y = -1: from the row above
            const double t = shfl_dn1(q.v[i].x);
            const double e = i == 4 ? (q.top ? q.v[2].w : q.e4) : i == 7 ? q.e7 : q.e8;
            n0[i] = q.v[i].y; n1[i] = q.v[i].z; n2[i] = q.v[i].w;
            n3[i] = lane == 31 ? e : t;
        }
    }
}

__device__ __forceinline__ void store_quad(const LbParams &p, int x, int y0, const double (&n0)[9],
                                           const double (&n1)[9], const double (&n2)[9],
                                           const double (&n3)[9]) {
    const Lattice &L = p.L;
    const long long c0 = L.off(x, y0);
#pragma unroll
    for (int i = 0; i < 9; i++) stg4(p.dst + i * L.plane + c0, n0[i], n1[i], n2[i], n3[i]);
    if (x < p.left.G) {  // see store_pair: three crossing populations when G == 1
        const Remote &r = p.left;
        const bool all = r.G > 1;
#pragma unroll
        for (int i = 0; i < 9; i++)
            if (all || i == 3 || i == 6 || i == 7)
                stg4(r.base + (long long)i * r.plane + (long long)(r.W + x + r.G) * r.ld + y0, n0[i],
                     n1[i], n2[i], n3[i]);
    }
    if (x >= L.W - p.right.G) {
        const Remote &r = p.right;
        const bool all = r.G > 1;
#pragma unroll
        for (int i = 0; i < 9; i++)
            if (all || i == 1 || i == 5 || i == 8)
                stg4(r.base + (long long)i * r.plane + (long long)(x - L.W + r.G) * r.ld + y0, n0[i],
                     n1[i], n2[i], n3[i]);
    }
}

}  // namespace
}  // namespace iblb
