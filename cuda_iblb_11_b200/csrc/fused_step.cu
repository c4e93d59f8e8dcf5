// One launch per iteration: immersed boundary and lattice Boltzmann in the same kernel.
//
//   every kKinBatch iterations   ib_kinematics_batch   markers: kinematics, epsilon, stencils  (ib_kernels.cu)
//                                ib_tile_lists         per band tile and iteration: the ascending list of
//                                                      markers whose 3x3 support can touch the tile
//   every iteration              lb_ib_tile_step       F_s of the tile's markers -> spread -> pull-stream,
//                                                      moments, TRT collide + Guo
//
// Replaces the reference's interpolate, spread, streaming, macro, equilibrium, collision launches
// (ImmersedBoundary.cu:94-267, LatticeBoltzmann.cu:30-411).  A tile is 8 (or 6) columns x 64 rows
// (warp = column, lane = row pair).  Everything about the markers that does not depend on the fluid
// — positions, velocities, epsilon, delta factors AND which markers reach which tile — is prescribed
// (main.cu:77-252 is a function of `it` alone) and is computed off the fluid's critical path for a
// batch of iterations.  What is left per iteration is done by the tile that needs it: a band
// tile forms rho, u* of the support nodes it holds from the populations it pulls anyway and pulls
// those of the two rings around it, evaluates F_s of every marker on its list (nine support nodes
// each, ImmersedBoundary.cu:94-133; a marker that reaches two tiles is evaluated by both: no
// inter-CTA dependency, no second launch, no atomics), then every node sums its spreading terms in
// ascending marker order, the order of ImmersedBoundary.cu:178-203.  Neither F_s nor the force goes
// through global memory: the force is consumed by the collision of the same thread.  Same device
// functions as lb_fused_v2 / ib_interp / ib_spread: bit-identical results (tests/test_gpu_parity.py).
// Two other forms of the iteration were built, measured and removed in round 2 (marker CTAs beside
// light LB CTAs in one launch; marker kernels beside LB kernels on two streams): DESIGN.md section 3.
#include "ib_device.cuh"
#include "iblb_device.cuh"
#include "lb_device.cuh"

namespace iblb {

namespace {

// programmatic dependent launch (PTX griddepcontrol); no-ops in a launch without the attribute
__device__ __forceinline__ void pdl_launch_dependents() {
    asm volatile("griddepcontrol.launch_dependents;");
}
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

struct TileMarkers {
    // the markers that can touch the tile: (F_x, F_y, bx, by), the 1-D delta factors of the columns
    // bx..bx+3 / rows by..by+3 (tables written by ib_kinematics_batch), position and velocity
    float4 fb[kTileCap], wx[kTileCap], wy[kTileCap], su[kTileCap];
    // rho and u* of every node a support of these markers can contain: table[(x - (tx0-2))*kTabY +
    // (y - (ty0-2))], x and y as the reference's interpolate counts them (unwrapped); rho = 0 marks
    // a node that contributes nothing
    double rho[kTabX * kTabY], ux[kTabX * kTabY], uy[kTabX * kTabY];
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

// moment-table cell of ring node idx in [0, 4*kTabY + 4*tw) of a tile tw columns wide
__device__ __forceinline__ int ring_cell(int idx, int tw) {
    int lx, ly;
    if (idx < 2 * kTabY) {  // two columns left of the tile
        lx = idx / kTabY;
        ly = idx - lx * kTabY;
    } else if (idx < 4 * kTabY) {  // two columns right of it
        const int r = idx - 2 * kTabY;
        lx = r / kTabY;
        ly = r - lx * kTabY;
        lx += tw + 2;
    } else {  // two rows below, two above
        const int r = idx - 4 * kTabY;
        const int row = r / tw;
        lx = 2 + (r - row * tw);
        ly = row < 2 ? row : 64 + row;
    }
    return lx * kTabY + ly;
}

// rho, u* of table cell c (a ring node, or an own node of a column past the channel), pulled from
// global memory as the reference's interpolate would address it
__device__ __forceinline__ void tile_pulled_cell(const Lattice &L, const double *__restrict__ src,
                                                 TileMarkers &sh, int tx0, int ty0, int c) {
    IBLB_CHK(c >= 0 && c < kTabCells);
    const int lx = c / kTabY, ly = c - lx * kTabY;
    double rho, ux, uy;
    if (!support_node_moments(L, src, tx0 - 2 + lx, ty0 - 2 + ly, rho, ux, uy)) rho = ux = uy = 0.;
    sh.rho[c] = rho;
    sh.ux[c] = ux;
    sh.uy[c] = uy;
}

// rho and u* = sum c f / rho of pulled populations, same order as support_node_moments
__device__ __forceinline__ void node_moments(const double (&f)[9], double &rho, double &ux,
                                             double &uy) {
    const double r = DA(DA(DA(DA(DA(DA(DA(DA(f[0], f[1]), f[2]), f[3]), f[4]), f[5]), f[6]), f[7]), f[8]);
    rho = r;
    ux = __ddiv_rn(DA(DS(DS(DA(DS(f[1], f[3]), f[5]), f[6]), f[7]), f[8]), r);
    uy = __ddiv_rn(DS(DS(DA(DA(DS(f[2], f[4]), f[5]), f[6]), f[7]), f[8]), r);
}

// table cell of the centre of a marker's 3x3 support (ImmersedBoundary.cu:108-109: nearbyint of the
// position); false when part of the support lies outside the table — only for a marker exactly 1.5
// off the tile, whose every weight on the tile is 0 whatever its F_s
__device__ __forceinline__ bool marker_centre_cell(float sx, float sy, int tx0, int ty0, int tw,
                                                   int &c) {
    const int x0 = (int)nearbyint((double)sx), y0 = (int)nearbyint((double)sy);
    const int lx0 = x0 - (tx0 - 2), ly0 = y0 - (ty0 - 2);
    c = lx0 * kTabY + ly0;
    return lx0 >= 1 && lx0 <= tw + 2 && ly0 >= 1 && ly0 <= kTabY - 2;
}

// which of the 32 staged markers q0..q0+31 force the node pair (xg; y0, y0+1): stencil columns
// bx..bx+3 contain xg and rows by..by+3 contain y0 or y0+1
__device__ __forceinline__ unsigned hit_mask32(const int2 *base, int q0, int n, int xg, int y0) {
    unsigned m = 0;
    const int cnt = min(32, n - q0);
    for (int e = 0; e < cnt; e++) {
        const int2 b = base[q0 + e];
        if ((unsigned)(xg - b.x) <= 3u && (unsigned)(y0 + 1 - b.y) <= 4u) m |= 1u << e;
    }
    return m;
}

// F_s of staged marker e (ImmersedBoundary.cu:94-133) from the moment table: nine terms in the
// fixed order i = 0..8, the float re-rounded after each (Q6).  The delta factors are those of the
// spreading stencil: support column x0+c is column bx + (x0+c-bx) of it, same expression, same bits.
__device__ __forceinline__ void tile_marker_force(TileMarkers &sh, int tx0, int ty0, int tw,
                                                  int e) {
    const float4 su = sh.su[e], fb = sh.fb[e], wx = sh.wx[e], wy = sh.wy[e];
    const int x0 = (int)nearbyint((double)su.x), y0 = (int)nearbyint((double)su.y);
    const int bx = __float_as_int(fb.z), by = __float_as_int(fb.w);
    const int ox = x0 - bx, oy = y0 - by;  // 1 or 2
    const float dx[3] = {ox == 1 ? wx.x : wx.y, ox == 1 ? wx.y : wx.z, ox == 1 ? wx.z : wx.w};
    const float dy[3] = {oy == 1 ? wy.x : wy.y, oy == 1 ? wy.y : wy.z, oy == 1 ? wy.z : wy.w};
    float Fx = 0.f, Fy = 0.f;
    int c0;
    if (marker_centre_cell(su.x, su.y, tx0, ty0, tw, c0)) {
        // all products first (independent), then the two chains of float-rounded additions
        double tx[9], ty[9];
#pragma unroll
        for (int i = 0; i < 9; i++) {
            const int cx = i == 0 ? 0 : i == 1 ? 1 : i == 2 ? 0 : i == 3 ? -1 : i == 4 ? 0 : (i == 5 || i == 8) ? 1 : -1;
            const int cy = i == 0 ? 0 : i == 1 ? 0 : i == 2 ? 1 : i == 3 ? 0 : i == 4 ? -1 : (i == 5 || i == 6) ? 1 : -1;
            const int c = c0 + cx * kTabY + cy;
            IBLB_CHK(c >= 0 && c < kTabCells);
            const double rho = sh.rho[c];
            const float del = dx[cx + 1] * dy[cy + 1];  // float product (ImmersedBoundary.cu:78)
            // no branch (the nine terms overlap); a node off the array (rho = 0) contributes +0
            const double px = interp_product(del, rho, su.z, sh.ux[c]);
            const double py = interp_product(del, rho, su.w, sh.uy[c]);
            tx[i] = rho != 0. ? px : 0.;
            ty[i] = rho != 0. ? py : 0.;
        }
#pragma unroll
        for (int i = 0; i < 9; i++) {
            Fx = (float)__dadd_rn((double)Fx, tx[i]);
            Fy = (float)__dadd_rn((double)Fy, ty[i]);
        }
    }
    sh.fb[e].x = Fx;
    sh.fb[e].y = Fy;
}

// Force on this thread's two nodes (global column xg, rows y0 and y0+1) from the n markers staged
// in sh: sum over markers (ascending) of (double)(F_s*delta [float product]),
// ImmersedBoundary.cu:196-197.  `hit` holds the precomputed masks of the first kHitWords*32 markers
// (which of them reach one of the two nodes); the thread walks only its own few hits, lowest entry
// first: ascending order is preserved and the loop length is the ~8 markers that touch a node.
__device__ __forceinline__ void tile_spread(const TileMarkers &sh, int n, int xg, int y0,
                                            const unsigned (&hit)[kHitWords], double &fx0,
                                            double &fy0, double &fx1, double &fy1) {
    for (int w = 0; w * 32 < n; w++) {
        unsigned mine;
        if (w < kHitWords) {
            mine = w == 0 ? hit[0] : w == 1 ? hit[1] : w == 2 ? hit[2] : hit[3];
        } else {  // long lists: found here
            mine = 0;
            const int cnt = min(32, n - w * 32);
            for (int e = 0; e < cnt; e++) {
                const float4 fb = sh.fb[w * 32 + e];
                if ((unsigned)(xg - __float_as_int(fb.z)) <= 3u &&
                    (unsigned)(y0 + 1 - __float_as_int(fb.w)) <= 4u)
                    mine |= 1u << e;
            }
        }
        // branch-free body: a row the marker does not reach gets delta 0, and adding the resulting
        // +-0 leaves the sum as it is (it starts at +0 and can never become -0) — the reference
        // itself adds every marker to every node, zeros included (ImmersedBoundary.cu:194-197)
        const float *__restrict__ wxs = reinterpret_cast<const float *>(sh.wx);
        const float *__restrict__ wys = reinterpret_cast<const float *>(sh.wy);
        // software-pipelined: the next hit's shared-memory loads are in flight while the current
        // one is added, so only the four DADD chains are serial
        int q = 0, ix = 0, iy = 0;
        float4 fb = make_float4(0.f, 0.f, 0.f, 0.f);
        if (mine) {
            q = w * 32 + __ffs(mine) - 1;
            mine &= mine - 1;
            fb = sh.fb[q];
            ix = xg - __float_as_int(fb.z);  // 0..3 for a hit
            iy = y0 - __float_as_int(fb.w);  // row y0; row y0+1 is iy+1; -1..3 for a hit
        } else {
            continue;
        }
        for (;;) {
            const float ddx = wxs[4 * q + ix];
            const float wa = wys[4 * q + max(iy, 0)];
            const float wb = wys[4 * q + min(iy + 1, 3)];
            const float Fx = fb.x, Fy = fb.y;
            const bool lo = iy >= 0, hi = iy <= 2;
            const bool more = mine != 0;
            if (more) {
                q = w * 32 + __ffs(mine) - 1;
                mine &= mine - 1;
                fb = sh.fb[q];
                ix = xg - __float_as_int(fb.z);
                iy = y0 - __float_as_int(fb.w);
            }
            const float del0 = lo ? ddx * wa : 0.f;
            const float del1 = hi ? ddx * wb : 0.f;
            fx0 += (double)(Fx * del0);
            fy0 += (double)(Fy * del0);
            fx1 += (double)(Fx * del1);
            fy1 += (double)(Fy * del1);
            if (!more) break;
        }
    }
}

// Per-thread spreading list, built in the kernel's prologue (before the previous iteration has
// finished): for the first kSpreadList markers that reach this thread's two nodes, the marker's
// staged index and the two float delta products (ImmersedBoundary.cu:78) — none of it depends on
// the fluid.  What is left on the critical path per hit is one F_s load, four float products and
// the four double adds.  Layout [entry][thread]: conflict-free.
constexpr int kSpreadList = 12;
struct SpreadLists {
    float d0[kSpreadList][256], d1[kSpreadList][256];
    unsigned short q[kSpreadList][256];
};

// Moves the lowest hits of this thread out of its masks into the list; returns how many.  The
// masks keep what did not fit (more than kSpreadList hits): tile_spread adds those afterwards,
// which preserves the ascending order.
__device__ __forceinline__ int build_spread_list(const TileMarkers &sh, SpreadLists &sl, int n,
                                                 int xg, int y0, unsigned (&hit)[kHitWords]) {
    const int t = threadIdx.x;
    const float *__restrict__ wxs = reinterpret_cast<const float *>(sh.wx);
    const float *__restrict__ wys = reinterpret_cast<const float *>(sh.wy);
    int cnt = 0;
#pragma unroll
    for (int w = 0; w < kHitWords; w++) {
        if (w * 32 >= n) break;
        unsigned mine = hit[w];
        while (mine && cnt < kSpreadList) {
            const int q = w * 32 + __ffs(mine) - 1;
            mine &= mine - 1;
            const float4 fb = sh.fb[q];
            const int ix = xg - __float_as_int(fb.z);  // 0..3 for a hit
            const int iy = y0 - __float_as_int(fb.w);  // row y0; row y0+1 is iy+1; -1..3 for a hit
            const float ddx = wxs[4 * q + ix];
            const float wa = wys[4 * q + max(iy, 0)];
            const float wb = wys[4 * q + min(iy + 1, 3)];
            IBLB_CHK(q >= 0 && q < kTileCap && ix >= 0 && ix <= 3 && iy >= -1 && iy <= 3 && t < 256);
            sl.d0[cnt][t] = iy >= 0 ? ddx * wa : 0.f;
            sl.d1[cnt][t] = iy <= 2 ? ddx * wb : 0.f;
            sl.q[cnt][t] = (unsigned short)q;
            cnt++;
        }
        hit[w] = mine;
    }
    return cnt;
}

// the listed part of the spreading sum (see tile_spread for the arithmetic)
__device__ __forceinline__ void list_spread(const TileMarkers &sh, const SpreadLists &sl, int cnt,
                                            double &fx0, double &fy0, double &fx1, double &fy1) {
    const int t = threadIdx.x;
    const float2 *__restrict__ fs = reinterpret_cast<const float2 *>(sh.fb);  // (F_x, F_y) at [2q]
#pragma unroll 4
    for (int i = 0; i < cnt; i++) {
        const float2 F = fs[2 * (int)sl.q[i][t]];
        const float del0 = sl.d0[i][t], del1 = sl.d1[i][t];
        fx0 += (double)(F.x * del0);
        fy0 += (double)(F.y * del0);
        fx1 += (double)(F.x * del1);
        fy1 += (double)(F.y * del1);
    }
}

}  // namespace

// The fluid-independent plan of every band tile, for every iteration slot of a kinematics batch:
// grid = (band row tiles, column tiles, slots).  Cilia are preselected by the bounding box of their
// (wrapped) marker positions, then the markers with epsilon != 0 within 1.5 of the tile are kept by
// ballot-ordered compaction (no atomics): the list is ascending in the marker index, which is the
// summation order of the reference's spread (ImmersedBoundary.cu:178-203).  From the list follow
// the per-thread hit masks of the spreading, and which nodes in and around the tile are support
// nodes of the interpolation (ImmersedBoundary.cu:108-124) — the only ones whose moments the step
// kernel has to form.
namespace {

template <int MPC>
__device__ __forceinline__ void tile_lists_body(const BandParams &p) {
    __shared__ short rel[kBandMaxRel];  // cilia whose bounding box overlaps the tile, ascending
    __shared__ int wcount[8];
    __shared__ int slist[kTileCap];
    __shared__ int2 sbase[kTileCap];
    __shared__ unsigned char need[kTabCells];
    const Lattice &L = p.lb.L;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int tile_y = blockIdx.x, tile_x = blockIdx.y, slot = blockIdx.z;
    const long long mo = (long long)slot * p.Ns;
    const float2 *__restrict__ bbox = p.M.bbox + (long long)slot * p.c_num;
    const float2 *__restrict__ ms = p.M.s + mo;
    const int *__restrict__ meps = p.M.eps + mo;
    const int2 *__restrict__ mbase = p.M.base + mo;
    TileProgram &pg = p.prog[((long long)slot * gridDim.x + tile_y) * p.tiles_x + tile_x];
    const int tw = p.tw;  // columns per tile = warps of lb_ib_tile_step
    const int tx0 = tile_x * tw + L.x0, ty0 = tile_y * 64;
    const float xlo = (float)tx0 - 1.5f, xhi = (float)(tx0 + tw - 1) + 1.5f;
    const float ylo = (float)ty0 - 1.5f, yhi = (float)(ty0 + 63) + 1.5f;
    int nrel = 0;
    for (int c0 = 0; c0 < p.c_num; c0 += 256) {
        const int m = c0 + t;
        bool take = false;
        if (m < p.c_num) {
            const float2 bb = bbox[m];
            take = (bb.x <= xhi) && (bb.y >= xlo);
        }
        int total;
        const int r = nrel + cta_rank(take, wcount, total);
        if (take && r < kBandMaxRel) rel[r] = (short)m;
        nrel += total;
    }
    if (nrel > kBandMaxRel) {
        if (t == 0) {
            atomicExch(p.err, 1);
            pg.n = 0;
            pg.nring = 0;
        }
        return;
    }
    __syncthreads();
    int n = 0;
    const int mpc = MPC ? MPC : p.mpc;
    for (int c0 = 0; c0 < nrel * mpc; c0 += 256) {
        const int c = c0 + t;
        bool take = false;
        int k = 0;
        if (c < nrel * mpc) {
            const int q = c / mpc;
            k = rel[q] * mpc + (c - q * mpc);
            const float2 sm = ms[k];
            take = (meps[k] != 0) && sm.x >= xlo && sm.x <= xhi && sm.y >= ylo && sm.y <= yhi;
        }
        int total;
        const int r = n + cta_rank(take, wcount, total);
        IBLB_CHK(!take || (k >= 0 && k < p.Ns));
        if (take && r < kTileCap) slist[r] = k;
        n += total;
    }
    if (n > kTileCap) {
        if (t == 0) atomicExch(p.err, 2);
        n = 0;
    }
    for (int c = t; c < kTabCells; c += 256) need[c] = 0;
    __syncthreads();
    for (int e = t; e < n; e += 256) {
        const int k = slist[e];
        pg.list[e] = k;
        sbase[e] = mbase[k];
        const float2 sm = ms[k];
        int c0;
        if (marker_centre_cell(sm.x, sm.y, tx0, ty0, p.tw, c0)) {
#pragma unroll
            for (int dx = -1; dx <= 1; dx++)
#pragma unroll
                for (int dy = -1; dy <= 1; dy++) need[c0 + dx * kTabY + dy] = 1;
        }
    }
    __syncthreads();
    // this thread's node pair in lb_ib_tile_step: column tx0 + warp, rows ty0 + 2*lane, +1
#pragma unroll
    for (int w = 0; w < kHitWords; w++)
        pg.hit[w][t] = w * 32 < n ? hit_mask32(sbase, w * 32, n, tx0 + warp, ty0 + 2 * lane) : 0u;
    // the support nodes: per-thread flags for the tile's own nodes, a list for the ring nodes
    int nring = 0;
    {
        const int c = (warp + 2) * kTabY + 2 * lane + 2;
        pg.own[t] = (unsigned char)((need[c] ? 1 : 0) | (need[c + 1] ? 2 : 0));
    }
    const int nring_all = 4 * kTabY + 4 * tw;
    for (int i0 = 0; i0 < nring_all; i0 += 256) {
        const int idx = i0 + t;
        const bool take = idx < nring_all && need[ring_cell(idx, tw)] != 0;
        int total;
        const int r = nring + cta_rank(take, wcount, total);
        if (take) pg.ring[r] = (short)idx;
        nring += total;
    }
    if (t == 0) {
        pg.n = n;
        pg.nring = nring;
    }
}

}  // namespace

__global__ void __launch_bounds__(256) ib_tile_lists(const BandParams p) {
    if (p.mpc == kMarkersPerCilium) tile_lists_body<kMarkersPerCilium>(p);
    else tile_lists_body<0>(p);
}

// grid = row tiles x column tiles (one dimension), block = 32*tw.  Requires YDIM % 64 == 0 and a solver that owns the
// whole channel (ghost columns = its own opposite edge).  Row tiles >= band_rows/64 are plain LB;
// on tall grids only the band tiles are launched here and lb_fused_v2 takes the rows above.
// p.M and p.prog point at the iteration's slot.
// Development build (-DIBLB_TILE_TIMING, tools/tile_timing.sh): thread 0 of every CTA stamps clock64()
// at the phase boundaries of lb_ib_tile_step into p.dbg[cta][16] (slot 13: latest exit over all
// threads, 14 / 15: markers / ring nodes of the tile); iblb_destroy prints the averages.
#ifdef IBLB_TILE_TIMING
#define TSTAMP(i)                                                                     \
    do {                                                                              \
        if (threadIdx.x == 0 && p.dbg) p.dbg[(long long)blockIdx.x * 16 + (i)] = clock64(); \
    } while (0)
#define TSTAMP_EXIT()                                                                                  \
    do {                                                                                               \
        if (p.dbg) atomicMax((unsigned long long *)&p.dbg[(long long)blockIdx.x * 16 + 13],           \
                             (unsigned long long)clock64());                                           \
    } while (0)
#else
#define TSTAMP(i)
#define TSTAMP_EXIT()
#endif

namespace {
__device__ __forceinline__ void tile_step_body(const BandParams &p, unsigned char *smem_raw) {
    TileMarkers &sh = *reinterpret_cast<TileMarkers *>(smem_raw);
    SpreadLists &sl = *reinterpret_cast<SpreadLists *>(smem_raw + sizeof(TileMarkers));
    const LbParams &lp = p.lb;
    const Lattice &L = lp.L;
    // Block index -> tile: the band's tiles first, the row tiles above the band last.  With
    // programmatic dependent launch the next iteration's CTAs take the slots this iteration's CTAs
    // free, in block-index order: the early slots (freed by tiles without markers) then go to band
    // tiles, whose prologue (marker staging, spreading lists) runs under this iteration's tail; the
    // last slots (freed by the slow marker tiles) go to tiles that have no prologue at all.
    int tile_y, tile_x;
    {
        const int bt = min(p.rows_lb, p.band_rows >> 6), b = blockIdx.x, nb = bt * p.tiles_x;
        if (!p.band_first) {  // grids that leave SMs empty: row tile fastest (measured faster there)
            tile_x = b / p.rows_lb;
            tile_y = b - tile_x * p.rows_lb;
        } else if (b < nb) {
            tile_x = b / bt;
            tile_y = b - tile_x * bt;
        } else {
            const int r = b - nb, rt = p.rows_lb - bt;
            tile_x = r / rt;
            tile_y = bt + (r - tile_x * rt);
        }
    }
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int tw = p.tw, NT = tw * 32;  // columns per tile = warps of this CTA; threads
    const int tx0 = tile_x * tw, ty0 = tile_y * 64;
    const int x = tx0 + warp;
    const int y0 = ty0 + 2 * lane;
    const bool active = x < L.W;  // warp-uniform
    const bool band_tile = ty0 < p.band_rows;  // block-uniform

    // Programmatic dependent launch: the next iteration's CTAs may start as soon as every CTA of
    // this one is running.  Everything up to pdl_wait() reads only the prescribed marker data (the
    // tile's plan and its markers, written by kernels that completed before the first kernel of this
    // PDL chain started), so it overlaps the tail of the previous iteration; pdl_wait() returns once
    // the previous iteration has completed and its populations are visible.
    pdl_launch_dependents();
    TSTAMP(0);
#ifdef IBLB_TILE_TIMING
    if (threadIdx.x == 0 && p.dbg) p.dbg[(long long)blockIdx.x * 16 + 13] = 0;
#endif
    int n = 0, nring = 0;
    const TileProgram *__restrict__ pg = nullptr;
    if (band_tile) {
        pg = p.prog + ((long long)tile_y * p.tiles_x + tile_x);
        n = __ldg(&pg->n);
        nring = __ldg(&pg->nring);
    }
    int r0 = 0, r1 = 0, own = 0;
    unsigned hit[kHitWords] = {0u, 0u, 0u, 0u};
    if (n > 0) {  // block-uniform
        int k0 = 0, k1 = 0;
        if (t < n) k0 = __ldg(pg->list + t);
        if (t + NT < n) k1 = __ldg(pg->list + t + NT);
        if (t < nring) r0 = __ldg(pg->ring + t);
        if (t + NT < nring) r1 = __ldg(pg->ring + t + NT);
        own = __ldg(pg->own + t);
#pragma unroll
        for (int w = 0; w < kHitWords; w++)
            if (w * 32 < n) hit[w] = __ldg(&pg->hit[w][t]);
#pragma unroll
        for (int r = 0; r < 2; r++) {
            const int e = t + NT * r;
            if (e < n) {
                const int k = r ? k1 : k0;
                const float2 sm = p.M.s[k], us = p.M.u_s[k];
                const int2 b = p.M.base[k];
                sh.su[e] = make_float4(sm.x, sm.y, us.x, us.y);
                sh.fb[e] = make_float4(0.f, 0.f, __int_as_float(b.x), __int_as_float(b.y));
                sh.wx[e] = p.M.wx[k];
                sh.wy[e] = p.M.wy[k];
            }
        }
    }
    int nlist = 0;
    if (n > 0) {
        __syncthreads();  // the staged markers
        nlist = build_spread_list(sh, sl, n, x + L.x0, y0, hit);
    }
    // (index arithmetic of this thread's ring node: still fluid-independent, still before the wait)
    int rc0 = 0;
    long long ro[9];
    bool rvalid = false;
    if (t < nring) {
        rc0 = ring_cell(r0, tw);
        const int lx = rc0 / kTabY, ly = rc0 - lx * kTabY;
        rvalid = support_node_offsets(L, tx0 + L.x0 - 2 + lx, ty0 - 2 + ly, ro);
    }
    // ... and of this thread's own nine vector loads
    const long long pcm = active ? L.off(x - 1, y0) : 0, pc0 = active ? L.off(x, y0) : 0,
                    pcp = active ? L.off(x + 1, y0) : 0;
    TSTAMP(1);
    pdl_wait();
    TSTAMP(2);
#ifdef IBLB_TILE_TIMING
    if (threadIdx.x == 0 && p.dbg) {
        p.dbg[(long long)blockIdx.x * 16 + 14] = n;
        p.dbg[(long long)blockIdx.x * 16 + 15] = nring;
    }
#endif
    // The support nodes in the two rings around the tile: pulled into registers and turned into
    // moments after this thread's own nodes.  Their (few, scattered) loads are issued BEFORE the
    // tile's own 37 KB: behind those they queued for the whole load phase and the marker chain
    // started ~2000 cycles late; in front of them they delay the bulk by a few hundred.
    double rf[9] = {1., 0., 0., 0., 0., 0., 0., 0., 0.};
    if (rvalid) {
#pragma unroll
        for (int i = 0; i < 9; i++) rf[i] = lp.src[ro[i]];
    }
    Pulled q;
    if (active) load_pulled_at(L, lp.src, pcm, pc0, pcp, y0, lane, q);
    if (n > 0) {
        if (t + NT < nring) tile_pulled_cell(L, lp.src, sh, tx0 + L.x0, ty0, ring_cell(r1, tw));
        if (!active) {  // columns past the channel (W % 8 != 0): as the reference would address them
            const int c = (warp + 2) * kTabY + 2 * lane + 2;
            if (own & 1) tile_pulled_cell(L, lp.src, sh, tx0 + L.x0, ty0, c);
            if (own & 2) tile_pulled_cell(L, lp.src, sh, tx0 + L.x0, ty0, c + 1);
        }
    }

    TSTAMP(3);
    // the ring nodes' moments now, while the tile's own loads are still in flight (the ring pulls
    // were issued first and are back first)
    if (n > 0 && t < nring) {
        double rho, ux, uy;
        node_moments(rf, rho, ux, uy);  // rf = (1, 0, ...) placeholder when !rvalid
        sh.rho[rc0] = rvalid ? rho : 0.;
        sh.ux[rc0] = rvalid ? ux : 0.;
        sh.uy[rc0] = rvalid ? uy : 0.;
    }

    TSTAMP(4);
    double a[9], b[9];
    if (active) shift_pulled(q, lane, a, b);
    TSTAMP(5);

    double Fx0 = lp.gx, Fy0 = lp.gy, Fx1 = lp.gx, Fy1 = lp.gy;
    bool forced = (lp.gx != 0.0) | (lp.gy != 0.0);
    if (band_tile) forced = true;  // as the reference: spread's velocity update covers every node
    if (n > 0) {
        if (active) {  // this thread's support nodes: moments of the populations it pulled anyway
            const int c = (warp + 2) * kTabY + 2 * lane + 2;
            // a warp with any support node forms both nodes' moments in every lane, without
            // branches (the two divisions per node overlap); cells nobody needs are never read
            if (__any_sync(0xffffffffu, own != 0)) {
                double r0_, x0_, y0_, r1_, x1_, y1_;
                node_moments(a, r0_, x0_, y0_);
                node_moments(b, r1_, x1_, y1_);
                sh.rho[c] = r0_; sh.ux[c] = x0_; sh.uy[c] = y0_;
                sh.rho[c + 1] = r1_; sh.ux[c + 1] = x1_; sh.uy[c + 1] = y1_;
            }
        }
        TSTAMP(6);
        __syncthreads();
        TSTAMP(7);
        if (t < n) tile_marker_force(sh, tx0 + L.x0, ty0, tw, t);
        if (t + NT < n) tile_marker_force(sh, tx0 + L.x0, ty0, tw, t + NT);
        TSTAMP(8);
        __syncthreads();
        TSTAMP(9);
        double fx0 = 0., fy0 = 0., fx1 = 0., fy1 = 0.;
        list_spread(sh, sl, nlist, fx0, fy0, fx1, fy1);
        tile_spread(sh, n, x + L.x0, y0, hit, fx0, fy0, fx1, fy1);  // what did not fit the list
        Fx0 += fx0; Fy0 += fy0; Fx1 += fx1; Fy1 += fy1;
        TSTAMP(10);
    }
    if (!active) return;
    const double ux0 = collide_node(a, Fx0, Fy0, forced, lp);
    const double ux1 = collide_node(b, Fx1, Fy1, forced, lp);
    TSTAMP(11);
    store_pair(lp, x, y0, a, b);
    TSTAMP(12);
    TSTAMP_EXIT();

    // flux through x = XDIM-5 (ImmersedBoundary.cu:259-264), same slots and order as lb_fused_v2
    if (x == lp.xq && lp.qacc != nullptr) {
        double s = ux0 + ux1;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
        if (lane == 0) flux_add(lp, tile_y, s / (double)L.YDIM);
    }
}
}  // namespace

__global__ void __launch_bounds__(256, 2) lb_ib_tile_step(const BandParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];  // kTileSmemBytes, dynamic
    tile_step_body(p, smem_raw);
}

// dynamic shared memory of lb_ib_tile_step (above the 48 KB static limit: opt-in per device)
size_t tile_step_smem_bytes() { return sizeof(TileMarkers) + sizeof(SpreadLists); }
cudaError_t tile_step_prepare() {
    return cudaFuncSetAttribute(lb_ib_tile_step, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                (int)tile_step_smem_bytes());
}

IBLB_CHECKED_COUNTER(checked_count_fused)

// last node of a captured batch of iterations: the iteration counter the batch reads
__global__ void it_advance(int *it_ptr, int by) {
    if (threadIdx.x == 0 && blockIdx.x == 0) *it_ptr += by;
}

}  // namespace iblb
