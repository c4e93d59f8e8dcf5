// LB step with the force spreading fused in: the second (and last) launch of an iteration.
//
//   launch 1  ib_markers      markers: kinematics, epsilon, F_s           (ib_kernels.cu)
//   launch 2  lb_band_step    every node: [spread F_s ->] pull-stream, moments, TRT collide + Guo
//
// Replaces the reference's spread, streaming, macro, equilibrium, collision launches
// (ImmersedBoundary.cu:138-267, LatticeBoltzmann.cu:30-411).  A CTA owns a tile of 8 columns x 64
// rows (warp = column, lane = row pair).  Tiles of the forcing band (rows < 128) first build, by
// ballot-ordered compaction (no atomics), the ascending list of markers whose 3x3 support can touch
// the tile — cilia preselected by their bounding boxes — and every node sums its spreading terms in
// ascending marker order, the order of ImmersedBoundary.cu:178-203.  The force never goes through
// memory: it is consumed by the collision of the same thread.  The LB loads are issued before the
// list is built, so their latency hides behind it.  Same device functions as lb_fused_v2 /
// ib_spread: bit-identical results (tests/test_gpu_parity.py).
#include "ib_device.cuh"
#include "iblb_device.cuh"
#include "lb_device.cuh"

namespace iblb {

namespace {

struct TileList {
    // the markers that can touch the tile: (F_x, F_y, bx, by) and the 1-D delta factors of the
    // columns bx..bx+3 / rows by..by+3 (tables written by ib_markers)
    float4 fb[kBandMaxList], wx[kBandMaxList], wy[kBandMaxList];
    short rel[kBandMaxRel];       // cilia whose bounding box overlaps the tile, ascending
    int wcount[8];
};

// ordered compaction helper: position of this thread's element among the taken ones of the CTA
__device__ __forceinline__ int cta_rank(bool take, int *wcount, int &total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned tk = __ballot_sync(0xffffffffu, take);
    if (lane == 0) wcount[warp] = __popc(tk);
    __syncthreads();
    int before = 0;
    total = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) {
        const int cw = wcount[w];
        if (w < warp) before += cw;
        total += cw;
    }
    const int r = before + __popc(tk & ((1u << lane) - 1u));
    __syncthreads();  // wcount may be reused
    return r;
}

// Force on this thread's two nodes (global column xg, rows y0 and y0+1) from the markers.
// All 256 threads of the CTA must call it.
__device__ __forceinline__ void tile_spread(const BandParams &p, TileList &sh, int tile_x,
                                            int tile_y, int xg, int y0, double &fx0, double &fy0,
                                            double &fx1, double &fy1) {
    const Lattice &L = p.lb.L;
    const int t = threadIdx.x;
    const float xlo = (float)(tile_x * 8 + L.x0) - 1.5f, xhi = (float)(tile_x * 8 + L.x0 + 7) + 1.5f;
    const float ylo = (float)(tile_y * 64) - 1.5f, yhi = (float)(tile_y * 64 + 63) + 1.5f;
    // ---- cilia whose bounding box of (wrapped) marker x overlaps the tile's reach ----
    int nrel = 0;
    for (int c0 = 0; c0 < p.c_num; c0 += 256) {
        const int m = c0 + t;
        bool rel = false;
        if (m < p.c_num) {
            const float2 bb = p.M.bbox[m];
            rel = (bb.x <= xhi) && (bb.y >= xlo);
        }
        int total;
        const int r = nrel + cta_rank(rel, sh.wcount, total);
        if (rel && r < kBandMaxRel) sh.rel[r] = (short)m;
        nrel += total;
    }
    if (nrel > kBandMaxRel) {
        if (t == 0) atomicExch(p.err, 1);
        return;
    }
    __syncthreads();
    // ---- ascending (cilium, marker) list ----
    int n = 0;
    for (int c0 = 0; c0 < nrel * kMarkersPerCilium; c0 += 256) {
        const int c = c0 + t;
        bool take = false;
        int k = 0;
        if (c < nrel * kMarkersPerCilium) {
            const int q = c / kMarkersPerCilium;
            k = sh.rel[q] * kMarkersPerCilium + (c - q * kMarkersPerCilium);
            const float2 sm = p.M.s[k];
            take = (p.M.eps[k] != 0) && sm.x >= xlo && sm.x <= xhi && sm.y >= ylo && sm.y <= yhi;
        }
        int total;
        const int r = n + cta_rank(take, sh.wcount, total);
        if (take && r < kBandMaxList) {
            const float2 F = p.M.F_s[k];
            const int2 b = p.M.base[k];
            sh.fb[r] = make_float4(F.x, F.y, __int_as_float(b.x), __int_as_float(b.y));
            sh.wx[r] = p.M.wx[k];
            sh.wy[r] = p.M.wy[k];
        }
        n += total;
    }
    if (n > kBandMaxList) {
        if (t == 0) atomicExch(p.err, 2);
        return;
    }
    __syncthreads();
    // ---- force = sum over markers (ascending) of (double)(F_s*delta [float product]) ----
    // 32 entries at a time: the warp filters the markers whose stencil contains its column
    // (ballot), every lane then keeps, as a bit mask, those that also reach one of its two rows
    // and walks only its own few hits, lowest entry first: ascending order is preserved and the
    // loop length is the ~8 markers that touch a node instead of the whole list.
    const int lane = t & 31;
    for (int q0 = 0; q0 < n; q0 += 32) {
        bool hit = false;
        int by_l = 0;
        if (q0 + lane < n) {
            const float4 fb = sh.fb[q0 + lane];
            hit = (unsigned)(xg - __float_as_int(fb.z)) <= 3u;
            by_l = __float_as_int(fb.w);
        }
        unsigned col = __ballot_sync(0xffffffffu, hit);
        unsigned mine = 0;
        while (col) {
            const int e = __ffs(col) - 1;
            col &= col - 1;
            const int by_e = __shfl_sync(0xffffffffu, by_l, e);
            if ((unsigned)(y0 + 1 - by_e) <= 4u) mine |= 1u << e;  // rows by..by+3 vs y0, y0+1
        }
        while (mine) {
            const int q = q0 + __ffs(mine) - 1;
            mine &= mine - 1;
            const float4 fb = sh.fb[q], wx = sh.wx[q], wy = sh.wy[q];
            const int ix = xg - __float_as_int(fb.z);
            const int iy = y0 - __float_as_int(fb.w);  // row y0; row y0+1 is iy+1
            const float ddx = ix == 0 ? wx.x : ix == 1 ? wx.y : ix == 2 ? wx.z : wx.w;
            if (iy >= 0 && iy <= 3) {
                const float del = ddx * (iy == 0 ? wy.x : iy == 1 ? wy.y : iy == 2 ? wy.z : wy.w);
                fx0 += (double)(fb.x * del);
                fy0 += (double)(fb.y * del);
            }
            if (iy >= -1 && iy <= 2) {
                const float del = ddx * (iy == -1 ? wy.x : iy == 0 ? wy.y : iy == 1 ? wy.z : wy.w);
                fx1 += (double)(fb.x * del);
                fy1 += (double)(fb.y * del);
            }
        }
    }
}

}  // namespace

// grid = (row tiles, ceil(W/8)), block = 256.  Requires YDIM % 64 == 0 and a solver that owns the
// whole channel (ghost columns = its own opposite edge).  Row tiles >= band_rows/64 are plain LB;
// on tall grids only the band tiles are launched here and lb_fused_v2 takes the rows above.
__global__ void __launch_bounds__(256, 3) lb_band_step(const BandParams p) {
    __shared__ TileList sh;
    const LbParams &lp = p.lb;
    const Lattice &L = lp.L;
    const int tile_y = blockIdx.x, tile_x = blockIdx.y;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int x = tile_x * 8 + warp;
    const int y0 = tile_y * 64 + 2 * lane;
    const bool active = x < L.W;  // warp-uniform

    Pulled q;
    if (active) load_pulled(L, lp.src, x, y0, lane, q);  // in flight while the list is built

    double Fx0 = lp.gx, Fy0 = lp.gy, Fx1 = lp.gx, Fy1 = lp.gy;
    bool forced = (lp.gx != 0.0) | (lp.gy != 0.0);
    if (tile_y * 64 < p.band_rows) {  // block-uniform
        double fx0 = 0., fy0 = 0., fx1 = 0., fy1 = 0.;
        tile_spread(p, sh, tile_x, tile_y, x + L.x0, y0, fx0, fy0, fx1, fy1);
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

// last node of a captured batch of iterations: the iteration counter the batch reads
__global__ void it_advance(int *it_ptr, int by) {
    if (threadIdx.x == 0 && blockIdx.x == 0) *it_ptr += by;
}

}  // namespace iblb
