// Fused pull-stream + moments + TRT collide + Guo forcing for D2Q9/FP64 on sm_100a.
//
// One launch replaces the reference's chain  streaming -> macro -> [spread's velocity update] ->
// equilibrium -> collision  (LatticeBoltzmann.cu:173-373, 375-411; ImmersedBoundary.cu:249-264;
// LatticeBoltzmann.cu:30-62, 64-136): only post-collision populations live in HBM between steps,
// 144 B per lattice update (9 doubles read + 9 written) against the reference's ~816 B.
//
// Mapping: one thread owns two vertically adjacent nodes (y0, y0+1) of one column, one warp owns
// 64 consecutive rows.  Every population is fetched with ONE aligned 128-bit load per thread
// from the column it streams from (x - c_x); the +-1 row shift of the six populations with
// c_y != 0 is done in registers with a warp shuffle, and only lane 0 / lane 31 issue one extra
// 8-byte load each for the value that comes from the neighbouring warp's rows.  The walls are
// folded into that edge load: at y = 0 the three upward populations are the node's own
// downward ones (half-way bounce-back, LatticeBoltzmann.cu:231,287-291,309-313,328-340), at
// y = YDIM-1 the three downward ones are the node's own upward ones mirrored in y only
// (specular, LatticeBoltzmann.cu:225,243-247,265-269,341-353).  x-periodicity / slab coupling
// costs nothing: ghost columns hold the neighbour's values and the threads of the boundary
// columns store their results a second time into the neighbour's ghost columns (a peer
// pointer over NVLink when the neighbour is another GPU).
#include "iblb_device.cuh"
#include "lb_device.cuh"

namespace iblb {

// Fast path: YDIM % 64 == 0.  grid = ceil(W*YDIM/2 / 256), block = 256.
__global__ void __launch_bounds__(256, 3) lb_fused_v2(const LbParams p) {
    const Lattice &L = p.L;
    const int HP = ((p.row_end ? p.row_end : L.YDIM) - p.row_begin) >> 1;  // row pairs per column (multiple of 32)
    const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (gid >= (long long)p.col_count * HP) return;  // whole warps only
    const int c = (int)(gid / HP);
    const int yp = (int)(gid - (long long)c * HP);
    int x = p.col_begin + c;
    if (x >= p.split_at) x += p.split_skip;
    const int y0 = p.row_begin + (yp << 1);
    const int lane = threadIdx.x & 31;

    Pulled q;
    load_pulled(L, p.src, x, y0, lane, q);
    // force: IB band (rows < band_rows) + uniform body force
    double Fx0 = p.gx, Fy0 = p.gy, Fx1 = p.gx, Fy1 = p.gy;
    bool forced = (p.gx != 0.0) | (p.gy != 0.0);
    if (p.band != nullptr && y0 < p.band_rows) {  // warp-uniform: band_rows % 64 == 0
        IBLB_CHK(x >= 0 && x < L.W && y0 >= 0 && y0 + 1 < p.band_ld);
        const double *b = p.band + (long long)x * p.band_ld + y0;
        const double2 bx = ldg2(b);
        const double2 by = ldg2(b + (long long)L.W * p.band_ld);
        Fx0 += bx.x; Fx1 += bx.y; Fy0 += by.x; Fy1 += by.y;
        forced = true;
    }

    double a[9], b[9];  // node y0, node y0+1
    shift_pulled(q, lane, a, b);
    const double ux0 = collide_node(a, Fx0, Fy0, forced, p);
    const double ux1 = collide_node(b, Fx1, Fy1, forced, p);
    store_pair(p, x, y0, a, b);

    // flux through the plane x = XDIM-5 (ImmersedBoundary.cu:259-264): fixed-order warp tree,
    // one running sum per 64-row slot, summed on the host in slot order => deterministic
    if (x == p.xq && p.qacc != nullptr) {
        double s = ux0 + ux1;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
        if (lane == 0) flux_add(p, y0 >> 6, s / (double)L.YDIM);
    }
}

// Four nodes per thread, 256-bit loads and stores (LDG.E.256 / STG.E.256): rows [row_begin, row_end)
// in multiples of 128 (a warp owns 128 consecutive rows of one column), no IB force band.  Same
// device functions as lb_fused_v2, same bits; twice the bytes in flight per warp.  Used for the
// pure-LB channel (C2) and for the rows above the force band of tall grids (C4).
__global__ void __launch_bounds__(256, 2) lb_fused_v4(const LbParams p) {
    const Lattice &L = p.L;
    const int QP = ((p.row_end ? p.row_end : L.YDIM) - p.row_begin) >> 2;  // row quads per column
    const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (gid >= (long long)p.col_count * QP) return;  // whole warps only
    const int c = (int)(gid / QP);
    const int yq = (int)(gid - (long long)c * QP);
    int x = p.col_begin + c;
    if (x >= p.split_at) x += p.split_skip;
    const int y0 = p.row_begin + (yq << 2);
    const int lane = threadIdx.x & 31;

    Pulled4 q;
    load_pulled4(L, p.src, x, y0, lane, q);
    const bool forced = (p.gx != 0.0) | (p.gy != 0.0);
    double n0[9], n1[9], n2[9], n3[9];
    shift_pulled4(q, lane, n0, n1, n2, n3);
    const double u0 = collide_node(n0, p.gx, p.gy, forced, p);
    const double u1 = collide_node(n1, p.gx, p.gy, forced, p);
    const double u2 = collide_node(n2, p.gx, p.gy, forced, p);
    const double u3 = collide_node(n3, p.gx, p.gy, forced, p);
    store_quad(p, x, y0, n0, n1, n2, n3);

    // flux through x = XDIM-5: the summation tree of lb_fused_v2, bit for bit.  There a lane holds
    // the pair sum of rows (2l, 2l+1) and the 32 pair sums of a 64-row slot are folded with offsets
    // 16, 8, 4, 2, 1; here a lane holds two consecutive pair sums and a slot is 16 lanes: the same
    // fold is offsets 8, 4, 2, 1 on each of the two, then their sum.
    if (x == p.xq && p.qacc != nullptr) {
        double s0 = u0 + u1, s1 = u2 + u3;
#pragma unroll
        for (int o = 8; o > 0; o >>= 1) {
            s0 += __shfl_down_sync(0xffffffffu, s0, o);
            s1 += __shfl_down_sync(0xffffffffu, s1, o);
        }
        if ((lane & 15) == 0) flux_add(p, y0 >> 6, (s0 + s1) / (double)L.YDIM);
    }
}

// Generic path: any YDIM >= 2, one node per thread, scalar loads.  Same arithmetic.
__global__ void __launch_bounds__(256) lb_fused_v1(const LbParams p, double *qpart) {
    const Lattice &L = p.L;
    const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const bool active = gid < (long long)p.col_count * L.YDIM;
    int x = 0, y = 0;
    double ux = 0.;
    if (active) {
        const int c = (int)(gid / L.YDIM);
        y = (int)(gid - (long long)c * L.YDIM);
        x = p.col_begin + c;
        if (x >= p.split_at) x += p.split_skip;
        const double *__restrict__ src = p.src;
        const long long P = L.plane;
        double f[9];
#pragma unroll
        for (int i = 0; i < 9; i++) {
            int pi = i, xs = x - kCx[i], ys = y - kCy[i];
            if (ys < 0) {  // bottom wall: own opposite population
                pi = (i == 2) ? 4 : (i == 5) ? 7 : 8;
                xs = x; ys = y;
            } else if (ys >= L.YDIM) {  // top wall: own y-mirrored population
                pi = (i == 4) ? 2 : (i == 7) ? 6 : 5;
                xs = x; ys = y;
            }
            f[i] = __ldg(src + pi * P + L.off(xs, ys));
        }
        double Fx = p.gx, Fy = p.gy;
        bool forced = (p.gx != 0.0) | (p.gy != 0.0);
        if (p.band != nullptr && y < p.band_rows) {
            const double *b = p.band + (long long)x * p.band_ld + y;
            Fx += __ldg(b);
            Fy += __ldg(b + (long long)L.W * p.band_ld);
            forced = true;
        }
        ux = collide_node(f, Fx, Fy, forced, p);
        const long long c0 = L.off(x, y);
#pragma unroll
        for (int i = 0; i < 9; i++) p.dst[i * P + c0] = f[i];
        if (x < p.left.G) {  // see store_pair: three crossing populations when G == 1
#pragma unroll
            for (int i = 0; i < 9; i++)
                if (p.left.G > 1 || i == 3 || i == 6 || i == 7)
                    p.left.base[(long long)i * p.left.plane +
                                (long long)(p.left.W + x + p.left.G) * p.left.ld + y] = f[i];
        }
        if (x >= L.W - p.right.G) {
#pragma unroll
            for (int i = 0; i < 9; i++)
                if (p.right.G > 1 || i == 1 || i == 5 || i == 8)
                    p.right.base[(long long)i * p.right.plane +
                                 (long long)(x - L.W + p.right.G) * p.right.ld + y] = f[i];
        }
    }
    // flux: per-node values of the flux column go to a scratch column, reduced by flux_reduce_v1
    if (active && x == p.xq && qpart != nullptr) qpart[y] = ux / (double)L.YDIM;
}

// generic path's deterministic flux reduction: one block, fixed strided order + fixed tree
__global__ void flux_reduce_v1(const double *qpart, int n, double *qacc, double *qhist,
                               const int *it_ptr, int it) {
    __shared__ double sh[256];
    double v = 0.;
    for (int y = threadIdx.x; y < n; y += 256) v += qpart[y];
    sh[threadIdx.x] = v;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const double q = qacc[0] + sh[0];
        qacc[0] = q;
        if (qhist != nullptr) qhist[((it_ptr ? *it_ptr : 0) + it) % kFluxHist] = q;  // nq == 1
    }
}

// ---------------------------------------------------------------------------------------------
// Initial state (main.cu:636-654, 722-754): f = f0(rho,u); stored post-collision:
// f* = f0 + F(u, g) because TRT leaves an equilibrium unchanged.  rho/u come from device arrays
// in the reference layout (j = y*W + x) or, when null, rho = 1, u_x = u0_amp*sin(2pi(y+.5)/Y).
// Rows [row0, row0 + nrows) per launch: rho_in / u_in hold that chunk only (plane stride W*nrows),
// so that the host never needs a device copy of the whole field.
namespace {
__device__ __forceinline__ void init_node(const LbParams &p, int x, int y, double rho, double ux,
                                          double uy) {
    const Lattice &L = p.L;
    const double w[9] = {kW0, kW1, kW1, kW1, kW1, kW2, kW2, kW2, kW2};
    const double cs2 = kCs * kCs;
    const long long c0 = L.off(x, y);
#pragma unroll
    for (int i = 0; i < 9; i++) {
        const double cx = kCx[i], cy = kCy[i];
        const double cu = ux * cx + uy * cy;
        // LatticeBoltzmann.cu:47-50 / :53-56 evaluated as written
        double f0 = rho * w[i] *
                    (1 + cu / cs2 + cu * cu / (2 * cs2 * cs2) - (ux * ux + uy * uy) / (2 * cs2));
        if (i > 0) {
            const double vx = (cx - ux) / cs2 + cu / (cs2 * cs2) * cx;
            const double vy = (cy - uy) / cs2 + cu / (cs2 * cs2) * cy;
            f0 += p.kforce * w[i] * (vx * p.gx + vy * p.gy);
        }
        p.dst[i * L.plane + c0] = f0;
    }
}
}  // namespace

// rho_in == nullptr: the initial state of main.cu:636-654 (one thread per node, y fastest)
__global__ void lb_init(const LbParams p, const double *rho_in, const double *u_in, double u0_amp,
                        int row0, int nrows) {
    const Lattice &L = p.L;
    const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (gid >= (long long)L.W * nrows) return;
    const int x = (int)(gid / nrows), y = row0 + (int)(gid - (long long)x * nrows);
    double rho = 1.0, ux, uy = 0.0;
    const double kTwoPi = 6.283185307179586476925286766559;
    if (rho_in != nullptr) {
        const long long j = (long long)(y - row0) * L.W + x, size = (long long)L.W * nrows;
        rho = rho_in[j];
        ux = u_in[j];
        uy = u_in[size + j];
    } else {
        ux = u0_amp * sin(kTwoPi * (y + 0.5) / L.YDIM);
    }
    init_node(p, x, y, rho, ux, uy);
}

// The same from host fields (iblb_set_fields): the inputs are x-fastest (reference layout), the
// populations y-fastest, so a block transposes a 32x32 tile through shared memory and both sides
// move in full sectors (the one-thread-per-node form read 8 bytes per 32-byte sector: 0.95 ms per
// 128 MB chunk against 2.7 ms of PCIe copy, a quarter of the upload leg of e2e).
// grid = (ceil(W/32), ceil(nrows/32)), block = 256.
__global__ void __launch_bounds__(256) lb_init_t(const LbParams p, const double *rho_in,
                                                 const double *u_in, int row0, int nrows) {
    __shared__ double srho[32][33], sux[32][33], suy[32][33];
    const Lattice &L = p.L;
    const int tx0 = blockIdx.x * 32, ty0 = blockIdx.y * 32;
    const int a = threadIdx.x & 31, b = threadIdx.x >> 5;
    const long long size = (long long)L.W * nrows;
#pragma unroll
    for (int k = 0; k < 4; k++) {  // rows of the chunk, x fastest
        const int yl = ty0 + b + 8 * k, x = tx0 + a;
        if (yl < nrows && x < L.W) {
            const long long j = (long long)yl * L.W + x;
            srho[b + 8 * k][a] = rho_in[j];
            sux[b + 8 * k][a] = u_in[j];
            suy[b + 8 * k][a] = u_in[size + j];
        }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 4; k++) {  // columns of the lattice, y fastest
        const int x = tx0 + b + 8 * k, yl = ty0 + a;
        if (yl < nrows && x < L.W)
            init_node(p, x, row0 + yl, srho[a][b + 8 * k], sux[a][b + 8 * k], suy[a][b + 8 * k]);
    }
}

// copy the owned boundary columns of `buf` into the neighbours' ghost columns (used after init;
// the fused kernel does this itself every step)
__global__ void halo_push(const Lattice L, const double *buf, Remote left, Remote right) {
    const int y = blockIdx.x * blockDim.x + threadIdx.x;
    const int g = blockIdx.y;  // ghost column index
    const int i = blockIdx.z;
    if (y >= L.YDIM) return;
    if (g < left.G) {
        left.base[(long long)i * left.plane + (long long)(left.W + g + left.G) * left.ld + y] =
            buf[i * L.plane + L.off(g, y)];
    }
    if (g < right.G) {
        right.base[(long long)i * right.plane + (long long)(g - right.G + right.G) * right.ld + y] =
            buf[i * L.plane + L.off(L.W - right.G + g, y)];
    }
}

// ---------------------------------------------------------------------------------------------
// Macroscopic fields / streamed populations of the last completed iteration, recomputed from the
// buffer that iteration pulled from (still intact: the step wrote the other buffer).  Output in
// the reference layouts.  `streamed` false = iteration 0 (no step yet): local moments.
// Rows [row0, row0 + nrows) per launch, written to chunk-local arrays (plane stride W*nrows).
namespace {
// rho, u, force (and, f_wanted, the streamed populations) of node (x, y) as the reference's arrays
// hold them after the last completed iteration
__device__ __forceinline__ void moments_node(const LbParams &p, int streamed, int x, int y,
                                             bool f_wanted, double (&f)[9], double &rho_o,
                                             double &ux_o, double &uy_o, double &Fx, double &Fy) {
    const Lattice &L = p.L;
    const long long P = L.plane;
#pragma unroll
    for (int i = 0; i < 9; i++) {
        int pi = i, xs = x, ys = y;
        if (streamed) {
            xs = x - kCx[i];
            ys = y - kCy[i];
            if (ys < 0) {
                pi = (i == 2) ? 4 : (i == 5) ? 7 : 8;
                xs = x; ys = y;
            } else if (ys >= L.YDIM) {
                pi = (i == 4) ? 2 : (i == 7) ? 6 : 5;
                xs = x; ys = y;
            }
        }
        f[i] = p.src[pi * P + L.off(xs, ys)];
    }
    Fx = 0.;
    Fy = 0.;
    if (p.band != nullptr && y < p.band_rows) {
        Fx = p.band[(long long)x * p.band_ld + y];
        Fy = p.band[(long long)(L.W + x) * p.band_ld + y];
    }
    double rho = f[0] + f[1] + f[2] + f[3] + f[4] + f[5] + f[6] + f[7] + f[8];
    double mx = f[1] - f[3] + f[5] - f[6] - f[7] + f[8];
    double my = f[2] - f[4] + f[5] + f[6] - f[7] - f[8];
    if (!streamed) {
        // Before the first step the buffer holds f* = f0(rho,u) + F(u,g) (see lb_init), while the
        // reference's arrays hold the inputs rho, u and f = f0 (main.cu:636-754).  Undo it in
        // closed form.  With C_S = 0.57735 the equilibrium's moments are not exactly (rho, rho u):
        //   sum c f0 = rho u/(3 cs^2),  sum f0 = rho (1 + A u^2),  A = 1/(6cs^4) - 1/(2cs^2),
        //   sum_{i>=1} c F = k g/(3 cs^2),  sum_{i>=1} F = k (u.g) B,  B = 1/(3cs^4) - 5/(9cs^2).
        const double cs2 = kCs * kCs;
        const double A = 1. / (6. * cs2 * cs2) - 1. / (2. * cs2);
        const double B = 1. / (3. * cs2 * cs2) - 5. / (9. * cs2);
        const double rmx = 3. * cs2 * mx - p.kforce * p.gx, rmy = 3. * cs2 * my - p.kforce * p.gy;
        double r = rho, ux = 0., uy = 0.;
        for (int k = 0; k < 4; k++) {
            ux = rmx / r;
            uy = rmy / r;
            r = (rho - p.kforce * (ux * p.gx + uy * p.gy) * B) / (1. + (ux * ux + uy * uy) * A);
        }
        rho_o = r;
        ux_o = rmx / r;
        uy_o = rmy / r;
        if (f_wanted) {
            const double w[9] = {kW0, kW1, kW1, kW1, kW1, kW2, kW2, kW2, kW2};
#pragma unroll
            for (int i = 1; i < 9; i++) {
                const double cx = kCx[i], cy = kCy[i];
                const double cu = ux * cx + uy * cy;
                const double vx = (cx - ux) / cs2 + cu / (cs2 * cs2) * cx;
                const double vy = (cy - uy) / cs2 + cu / (cs2 * cs2) * cy;
                f[i] -= p.kforce * w[i] * (vx * p.gx + vy * p.gy);
            }
        }
    } else {
        rho_o = rho;
        ux_o = (mx + p.half * (Fx + p.gx)) / rho;
        uy_o = (my + p.half * (Fy + p.gy)) / rho;
    }
}
}  // namespace

// one thread per node, x fastest: the form that also delivers the populations (tests, AoS [9*j+i])
__global__ void lb_moments(const LbParams p, int streamed, int row0, int nrows, double *rho_out,
                           double *u_out, double *f_out, double *force_out) {
    const Lattice &L = p.L;
    const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long size = (long long)L.W * nrows;
    if (gid >= size) return;
    const int yl = (int)(gid / L.W), x = (int)(gid - (long long)yl * L.W);
    double f[9], rho, ux, uy, Fx, Fy;
    moments_node(p, streamed, x, row0 + yl, f_out != nullptr, f, rho, ux, uy, Fx, Fy);
    if (rho_out) rho_out[gid] = rho;
    if (u_out) {
        u_out[gid] = ux;
        u_out[size + gid] = uy;
    }
    if (f_out) {
#pragma unroll
        for (int i = 0; i < 9; i++) f_out[9 * gid + i] = f[i];
    }
    if (force_out) {
        force_out[gid] = Fx + p.gx;
        force_out[size + gid] = Fy + p.gy;
    }
}

// rho, u, force only (iblb_get_fields / iblb_get_force): 32x32 tiles through shared memory — the
// nodes are evaluated y fastest (the populations' layout), the results written x fastest (the
// reference's), both in full sectors.  grid = (ceil(W/32), ceil(nrows/32)), block = 256.
__global__ void __launch_bounds__(256) lb_moments_t(const LbParams p, int streamed, int row0,
                                                    int nrows, double *rho_out, double *u_out,
                                                    double *force_out) {
    __shared__ double sv[5][32][33];
    const Lattice &L = p.L;
    const int tx0 = blockIdx.x * 32, ty0 = blockIdx.y * 32;
    const int a = threadIdx.x & 31, b = threadIdx.x >> 5;
    const long long size = (long long)L.W * nrows;
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int x = tx0 + b + 8 * k, yl = ty0 + a;
        if (yl < nrows && x < L.W) {
            double f[9], rho, ux, uy, Fx, Fy;
            moments_node(p, streamed, x, row0 + yl, false, f, rho, ux, uy, Fx, Fy);
            sv[0][a][b + 8 * k] = rho;
            sv[1][a][b + 8 * k] = ux;
            sv[2][a][b + 8 * k] = uy;
            sv[3][a][b + 8 * k] = Fx + p.gx;
            sv[4][a][b + 8 * k] = Fy + p.gy;
        }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 4; k++) {
        const int yl = ty0 + b + 8 * k, x = tx0 + a;
        if (yl < nrows && x < L.W) {
            const long long gid = (long long)yl * L.W + x;
            if (rho_out) rho_out[gid] = sv[0][b + 8 * k][a];
            if (u_out) {
                u_out[gid] = sv[1][b + 8 * k][a];
                u_out[size + gid] = sv[2][b + 8 * k][a];
            }
            if (force_out) {
                force_out[gid] = sv[3][b + 8 * k][a];
                force_out[size + gid] = sv[4][b + 8 * k][a];
            }
        }
    }
}

IBLB_CHECKED_COUNTER(checked_count_lb)

// ---- slab ordering across GPUs (one process per GPU): step counters in peer-mapped memory.
// flags[0], flags[1]: "data ready" — the left / right neighbour has stored the boundary columns of
//   its previous iteration into this slab's ghost columns (written by its slab_signal);
// flags[2], flags[3]: "release" — that neighbour has BEGUN this iteration, so nothing of it (no
//   getter, no checkpoint) still reads the ghost columns of the buffer this slab's boundary kernel
//   is about to overwrite over there.  Without it a rank that has stopped to read its fields races
//   with a neighbour that is one step ahead (the getters pull from the previous source buffer).
// A slab first publishes its own release, then spins on ITS OWN memory; the writers are kernels on
// other GPUs, so both sides are always resident (never used between ranks that share a GPU).
// The spin is bounded: when a neighbour never arrives (a rank died, or the ranks were asked for
// different step counts) the kernel raises *err = 3 and falls through instead of hanging the GPU;
// every later wait returns at once and the next getter / iblb_sync reports it.  A slab in that
// state no longer publishes step numbers: it writes kSlabPoison into its neighbours' release and
// data-ready words, a neighbour's wait that finds the poison raises its own *err = 3 and passes it
// on, so the failure travels round the ring one hop per step and EVERY rank reports it -- nobody
// goes on computing with ghost columns that a failed slab filled from stale data.
namespace {
__device__ __forceinline__ void slab_fail(int *rel_left, int *rel_right, int *err) {
    atomicExch(err, 3);
    *reinterpret_cast<volatile int *>(rel_left) = kSlabPoison;
    *reinterpret_cast<volatile int *>(rel_right) = kSlabPoison;
    __threadfence_system();
}
}  // namespace

__global__ void slab_wait(const volatile int *flags, int *rel_left, int *rel_right, int need,
                          long long timeout_ns, int *err) {
    if (threadIdx.x != 0) return;
    if (*reinterpret_cast<volatile int *>(err) == 3) {
        slab_fail(rel_left, rel_right, err);
        return;
    }
    *reinterpret_cast<volatile int *>(rel_left) = need;
    *reinterpret_cast<volatile int *>(rel_right) = need;
    __threadfence_system();
    unsigned long long t0, t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (int i = 0; i < kSlabFlags; i++) {
        for (;;) {
            const int v = flags[i];
            if (v == kSlabPoison) {  // that neighbour has failed
                slab_fail(rel_left, rel_right, err);
                return;
            }
            if (v >= need) break;
            __nanosleep(100);
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            if ((long long)(t - t0) > timeout_ns) {
                slab_fail(rel_left, rel_right, err);
                return;
            }
        }
    }
    __threadfence_system();
}

// `err`: this slab's error word; once it holds 3 the neighbours get the poison instead of a step number
__global__ void slab_signal(int *left_flag, int *right_flag, int value, const int *err) {
    if (threadIdx.x == 0) {
        if (*reinterpret_cast<const volatile int *>(err) == 3) value = kSlabPoison;
        __threadfence_system();  // the boundary kernel's peer stores are ordered before the flag
        *reinterpret_cast<volatile int *>(left_flag) = value;
        *reinterpret_cast<volatile int *>(right_flag) = value;
        __threadfence_system();
    }
}

// iblb_set_fields on a ring that had failed: every slab starts again from a common state, so a
// poisoned word of this slab's own flags becomes the step number the new start expects
__global__ void slab_unpoison(int *flags, int value) {
    if (threadIdx.x < kSlabFlags && flags[threadIdx.x] == kSlabPoison) flags[threadIdx.x] = value;
}

}  // namespace iblb
