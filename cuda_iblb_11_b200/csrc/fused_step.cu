// One launch per iteration: immersed boundary + lattice Boltzmann fused (sm_100a).
//
// Replaces, for a solver that owns the whole channel, the chain
//   ib_kinematics -> ib_interpolate -> ib_spread -> lb_fused_v2
// (i.e. the reference's define_filament, boundary_check, interpolate, spread, streaming, macro,
// equilibrium, collision: main.cu:77-252, ImmersedBoundary.cu:94-267, LatticeBoltzmann.cu:30-411)
// by ONE kernel without any dependency between CTAs.  Everything a CTA needs is a function of the
// iteration number and of the source populations f* (read-only during the step):
//   * a CTA owns a tile of 8 columns x 64 rows (warp = column, lane = row pair);
//   * tiles of the forcing band (rows < 128) first rebuild, in shared memory, the part of the IB
//     state that can touch them: positions/velocities of the cilia whose reach overlaps the tile
//     (+ their left neighbours, for epsilon), the ascending list of markers whose 3x3 support
//     touches the tile, epsilon and F_s of those markers (81 pulled populations each) — then every
//     node sums its spreading terms in ascending marker order, exactly the order of
//     ImmersedBoundary.cu:178-203;
//   * all tiles then do the pull-stream + moments + TRT collide + Guo forcing of lb_fused_v2.
// F_s of a marker near a tile edge is recomputed by each tile it touches (~2x redundancy on 81
// loads per marker) — the price of having no grid-wide barrier; at 288x192 the step drops from four
// dependent kernels to one.  The arithmetic is the same device functions as the separate kernels
// (ib_device.cuh, lb_device.cuh): results are bit-identical (tests/test_gpu_parity.py).
#include "ib_device.cuh"
#include "iblb_device.cuh"
#include "lb_device.cuh"

namespace iblb {

namespace {

struct TileIb {
    double trig[kFusedMaxCilia][24];                    // per slot: [it, it-1][n][cos,sin]
    float2 pos[kFusedMaxCilia][kMarkersPerCilium];
    float2 vel[kFusedMaxCilia][kMarkersPerCilium];
    float2 l_s[kFusedMaxList], l_us[kFusedMaxList], l_F[kFusedMaxList];
    short l_slot[kFusedMaxList];
    unsigned char l_j[kFusedMaxList], l_eps[kFusedMaxList];
    int wcount[8];
};

__device__ __forceinline__ int pos_mod(int a, int b) {
    const int r = a % b;
    return r < 0 ? r + b : r;
}

// Force on this thread's two nodes (global column xg, rows y0 and y0+1) from the markers, for the
// tile (tile_x, tile_y).  All 256 threads of the CTA must call it.
__device__ __forceinline__ void tile_ib_force(const FusedParams &p, TileIb &sh, int it, int tile_x,
                                              int tile_y, int xg, int y0, double &fx0, double &fy0,
                                              double &fx1, double &fy1) {
    const Lattice &L = p.lb.L;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    // the tile's reach box, as ib_spread's
    const float xlo = (float)(tile_x * 8 + L.x0) - 1.5f, xhi = (float)(tile_x * 8 + L.x0 + 7) + 1.5f;
    const float ylo = (float)(tile_y * 64) - 1.5f, yhi = (float)(tile_y * 64 + 63) + 1.5f;

    // ---- which cilia: translated base of (unwrapped) cilium mu is c_space*(mu + 1/2); its markers
    // stay within +-reach of it; wrapping by +-XDIM is the same as mu -> mu +- c_num ----
    const int cs = p.c_space, cn = p.c_num;
    const int mu_lo = (int)ceilf((xlo - p.reach) / cs - 0.5f - 1e-3f);
    const int mu_hi = (int)floorf((xhi + p.reach) / cs - 0.5f + 1e-3f);
    const int n_span = mu_hi - mu_lo + 1;
    const bool all = n_span + p.r_max - 1 >= cn;
    const int nE = all ? cn : n_span + p.r_max - 1;     // slots: relevant cilia + left neighbours
    const int mu_s = all ? 0 : mu_lo - (p.r_max - 1);   // slot e holds cilium (mu_s + e) mod c_num
    const int rel0 = all ? 0 : p.r_max - 1;             // first relevant slot
    const int nrel = nE - rel0;
    if (nE > kFusedMaxCilia) {
        if (t == 0) atomicExch(p.err, 1);
        return;
    }
    // ---- trig tables and marker states of every slot ----
    for (int idx = t; idx < nE * 24; idx += 256) {
        const int e = idx / 24, r = idx - e * 24;
        sh.trig[e][r] = beat_trig(r % 12, it - r / 12, pos_mod(mu_s + e, cn), p.p_step, p.T);
    }
    __syncthreads();
    for (int idx = t; idx < nE * kMarkersPerCilium; idx += 256) {
        const int e = idx / kMarkersPerCilium, j = idx - e * kMarkersPerCilium;
        float2 sm, v;
        marker_state(p.coef + j * 28, reinterpret_cast<const double2 *>(sh.trig[e]),
                     reinterpret_cast<const double2 *>(sh.trig[e] + 12), pos_mod(mu_s + e, cn), it,
                     cs, cn, L.XDIM, sm, v);
        sh.pos[e][j] = sm;
        sh.vel[e][j] = v;
    }
    __syncthreads();
    // ---- ascending (cilium, marker) list of the markers whose support can touch the tile ----
    // relevant slots in ascending cilium index: the slot sequence wraps at most once
    const int m_first = pos_mod(mu_s + rel0, cn);
    const int rot = (cn - m_first < nrel) ? cn - m_first : 0;
    int n = 0;
    for (int c0 = 0; c0 < nrel * kMarkersPerCilium; c0 += 256) {
        const int c = c0 + t;
        bool take = false;
        int e = 0, j = 0;
        float2 sm = make_float2(0.f, 0.f);
        if (c < nrel * kMarkersPerCilium) {
            const int q = c / kMarkersPerCilium;
            j = c - q * kMarkersPerCilium;
            e = rel0 + (q + rot) % nrel;
            sm = sh.pos[e][j];
            take = sm.x >= xlo && sm.x <= xhi && sm.y >= ylo && sm.y <= yhi;
        }
        const unsigned tk = __ballot_sync(0xffffffffu, take);
        if (lane == 0) sh.wcount[warp] = __popc(tk);
        __syncthreads();
        int before = 0, total = 0;
#pragma unroll
        for (int w = 0; w < 8; w++) {
            const int cw = sh.wcount[w];
            if (w < warp) before += cw;
            total += cw;
        }
        if (take) {
            const int q = n + before + __popc(tk & ((1u << lane) - 1u));
            if (q < kFusedMaxList) {
                sh.l_s[q] = sm;
                sh.l_us[q] = sh.vel[e][j];
                sh.l_slot[q] = (short)e;
                sh.l_j[q] = (unsigned char)j;
            }
        }
        n += total;
        __syncthreads();
    }
    if (n > kFusedMaxList) {
        if (t == 0) atomicExch(p.err, 2);
        return;
    }
    // ---- epsilon (main.cu:217-248): warp per listed marker ----
    for (int q = warp; q < n; q += 8) {
        const float2 sm = sh.l_s[q];
        const int e = sh.l_slot[q];
        bool close = false;
        for (int r = 1; r < p.r_max; r++) {
            const int eo = all ? pos_mod(e - r, cn) : e - r;
            for (int l = lane; l < kMarkersPerCilium; l += 32) {
                const float2 o = sh.pos[eo][l];
                close |= (fabsf(o.x - sm.x) < 1) && (fabsf(o.y - sm.y) < 1);
            }
        }
        const bool any = __any_sync(0xffffffffu, close);
        if (lane == 0) sh.l_eps[q] = any ? 0 : 1;
    }
    __syncthreads();
    // ---- F_s (ImmersedBoundary.cu:94-133): nine lanes per marker, three markers per warp ----
    const int g = lane / 9, i9 = lane - 9 * g;
    for (int base = 0; base < n; base += 24) {
        const int q = base + warp * 3 + g;
        const bool on = (lane < 27) && (q < n) && sh.l_eps[q];
        double tx = 0., ty = 0.;
        if (on) interp_term(L, p.lb.src, sh.l_s[q], sh.l_us[q], i9, tx, ty);
        float Fx = 0.f, Fy = 0.f;
#pragma unroll
        for (int i = 0; i < 9; i++) {  // float re-rounded after each term, fixed order (Q6)
            const int srcl = (lane < 27) ? g * 9 + i : lane;
            const double ax = __shfl_sync(0xffffffffu, tx, srcl);
            const double ay = __shfl_sync(0xffffffffu, ty, srcl);
            Fx += ax;
            Fy += ay;
        }
        if (on && i9 == 0) sh.l_F[q] = make_float2(Fx, Fy);
    }
    __syncthreads();
    // ---- spreading: ascending marker order (ImmersedBoundary.cu:196-197) ----
    for (int q = 0; q < n; q++) {
        if (!sh.l_eps[q]) continue;
        const float2 mk = sh.l_s[q];
        const float dx = fabsf(xg - mk.x);
        if (dx > 1.5f) continue;
        const float2 F = sh.l_F[q];
        const float ddx = delta_1d(dx);
        const float dy0 = fabsf(y0 - mk.y), dy1 = fabsf(y0 + 1 - mk.y);
        if (dy0 <= 1.5f) {
            const float del = ddx * delta_1d(dy0);
            fx0 += (double)(F.x * del);
            fy0 += (double)(F.y * del);
        }
        if (dy1 <= 1.5f) {
            const float del = ddx * delta_1d(dy1);
            fx1 += (double)(F.x * del);
            fy1 += (double)(F.y * del);
        }
    }
}

}  // namespace

// grid = (YDIM/64, ceil(W/8)), block = 256.  Requires YDIM % 64 == 0 and a solver that owns the
// whole channel (ghost columns = its own opposite edge).
__global__ void __launch_bounds__(256, 2) iblb_fused_step(const FusedParams p) {
    __shared__ TileIb sh;
    const LbParams &lp = p.lb;
    const Lattice &L = lp.L;
    const int tile_y = blockIdx.x, tile_x = blockIdx.y;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int x = tile_x * 8 + warp;
    const int y0 = tile_y * 64 + 2 * lane;
    const bool active = x < L.W;  // warp-uniform

    Pulled q;
    if (active) load_pulled(L, lp.src, x, y0, lane, q);  // in flight while the IB part runs

    double Fx0 = lp.gx, Fy0 = lp.gy, Fx1 = lp.gx, Fy1 = lp.gy;
    bool forced = (lp.gx != 0.0) | (lp.gy != 0.0);
    if (tile_y * 64 < p.band_rows) {  // block-uniform
        const int it = (p.it_ptr ? *p.it_ptr : p.it) + p.it_off;
        double fx0 = 0., fy0 = 0., fx1 = 0., fy1 = 0.;
        tile_ib_force(p, sh, it, tile_x, tile_y, x + L.x0, y0, fx0, fy0, fx1, fy1);
        Fx0 += fx0; Fy0 += fy0; Fx1 += fx1; Fy1 += fy1;
        forced = true;
    }
    if (!active) return;

    double a[9], b[9];
    shift_pulled(q, lane, a, b);
    const double ux0 = collide_node(a, Fx0, Fy0, forced, lp);
    const double ux1 = collide_node(b, Fx1, Fy1, forced, lp);
    store_pair(lp, x, y0, a, b);

    // flux through x = XDIM-5 (ImmersedBoundary.cu:259-264), same slots and order as lb_fused_v2
    if (x == lp.xq && lp.qacc != nullptr) {
        double s = ux0 + ux1;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
        if (lane == 0) lp.qacc[tile_y] += s / (double)L.YDIM;
    }
}

// last node of a captured batch of fused steps: the iteration counter the batch reads
__global__ void it_advance(int *it_ptr, int by) {
    if (threadIdx.x == 0 && blockIdx.x == 0) *it_ptr += by;
}

}  // namespace iblb
