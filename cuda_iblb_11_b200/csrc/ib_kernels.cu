// Immersed-boundary side of the step: cilia kinematics, velocity interpolation, force spreading.
//
// Replaces define_filament + boundary_check (main.cu:77-252) and interpolate + spread
// (ImmersedBoundary.cu:94-267).  What changes against the reference:
//  * kinematics are evaluated only for the 96 markers of a cilium (the reference evaluates 9600
//    fine points per cilium and keeps 96), one block per cilium, trig shared through smem, and
//    are stateless: the velocity is pos(it) - pos(it-1) recomputed, not a stored `lasts` array;
//  * interpolation runs one warp per marker straight from the post-collision populations (the
//    nine support nodes pull their nine populations; rho/u never exist as full-grid arrays);
//  * spreading is a deterministic gather: a CTA owns an 8x32-node tile of the force band, builds
//    (by ballot-ordered compaction, no atomics) the ascending list of markers that touch it
//    and every node sums its contributions in ascending marker order — the reference's order
//    (ImmersedBoundary.cu:178-203) at O(9 Ns) instead of O(size*Ns) work.
// The float/double rounding points of the reference are kept on purpose (SURVEY Q3,Q6,Q7,Q14).
#include "iblb_device.cuh"
#include "ib_device.cuh"

namespace iblb {

// Per-marker shape coefficients a_n(arcl), b_n(arcl) (main.cu:109-135), once at create.
// k_of_j[j]: the fine point that wins marker j (main.cu:158-172).  coef[j][0..13] = a_n[2n+c],
// coef[j][14..27] = b_n[2n+c].  Float accumulators, double terms, as the reference.
__global__ void ib_coef(const int *k_of_j, float *coef, int mpc, int fine_points, int mucus) {
    const int j = threadIdx.x;
    if (j >= mpc) return;
    const float arcl = 1. * k_of_j[j] / fine_points;  // main.cu:100
    const double *A = mucus ? kAmucus : kA, *B = mucus ? kBmucus : kB;
    for (int n = 0; n < 7; n++)
        for (int c = 0; c < 2; c++) {
            float a = 0.f, b = 0.f;
            for (int pw = 0; pw < 3; pw++) {
                a += A[n + 14 * pw + 7 * c] * pow(arcl, pw + 1);
                b += B[n + 14 * pw + 7 * c] * pow(arcl, pw + 1);
            }
            coef[j * 28 + 2 * n + c] = a;
            coef[j * 28 + 14 + 2 * n + c] = b;
        }
}

// Everything about the markers that does not depend on the fluid, for a BATCH of iterations in one
// launch (the cilia are prescribed: main.cu:77-252 is a function of `it` alone): grid =
// (c_num, iterations), one block per cilium and iteration.
//   1. kinematics (main.cu:77-173,193-209) of the cilium and of the r_max-1 cilia to its left;
//   2. epsilon (main.cu:217-248): own markers against the left neighbours' markers;
//   3. the spreading stencil of every marker (columns/rows within 1.5 and their delta factors) and
//      the cilium's bounding box.
// Slot b of the marker arrays (stride Ns) receives iteration it0 + b.  Replaces the reference's
// define_filament and boundary_check launches, off the fluid's critical path: one launch per
// kKinBatch iterations.
// MPC: markers per cilium as a compile-time constant (the reference's 96: divisions by it fold
// away — they cost 13 us per batch as run-time divisions), or 0 for the run-time value p.mpc.
struct KinematicsShared {
    __align__(16) double trig[kMaxLeftCilia][24];  // slot 0 = own, e = cilium m-e
    float2 pos[kMaxLeftCilia * kMaxMarkersPerCilium];  // [e][j] with row pitch mpc: cilium m-e, marker j
    float2 vel[kMaxMarkersPerCilium];
    int closef[kMaxMarkersPerCilium];
    float wgt[kMaxMarkersPerCilium][8];
    float red_min[4], red_max[4];
};

template <int MPC>
__device__ __forceinline__ void kinematics_body(const IbParams &p, KinematicsShared &sh) {
    auto &trig = sh.trig;
    auto &pos = sh.pos;
    auto &vel = sh.vel;
    auto &closef = sh.closef;
    auto &wgt = sh.wgt;
    auto &red_min = sh.red_min;
    auto &red_max = sh.red_max;
    const int mpc = MPC ? MPC : p.mpc;  // markers per cilium, <= kMaxMarkersPerCilium
    const int m = blockIdx.x, slot = blockIdx.y;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int it = (p.it_ptr ? *p.it_ptr : p.it) + p.it_off + slot;
    const int r_max = p.r_max < 1 ? 1 : p.r_max;  // cilia whose positions are needed
    if (t < mpc) closef[t] = 0;
    for (int idx = t; idx < r_max * 24; idx += 1024) {
        const int e = idx / 24, r = idx - e * 24;
        const int mm = (m - e < 0) ? (m - e + p.c_num) : (m - e);
        if (e == 0 || r < 12) trig[e][r] = beat_trig(r % 12, it - r / 12, mm, p.p_step, p.T);
    }
    __syncthreads();
    for (int idx = t; idx < r_max * mpc; idx += 1024) {
        const int e = idx / mpc, j = idx - e * mpc;
        const int mm = (m - e < 0) ? (m - e + p.c_num) : (m - e);
        float2 sm, v;
        // neighbours only need the position: their trig[e][12..23] is not filled, it = 0 skips it
        marker_state(p.coef + j * 28, reinterpret_cast<const double2 *>(trig[e]),
                     reinterpret_cast<const double2 *>(trig[e] + 12), mm, e == 0 ? it : 0,
                     p.c_space, p.c_num, p.L.XDIM, p.scale, sm, v);
        if (p.replay != nullptr && j == p.replay_j && it >= 0 && it < p.replay_steps) {
            // the reference's racing marker exactly as a recorded run produced it (test hook)
            const float4 r = p.replay[(long long)it * p.c_num + mm];
            sm = make_float2(r.x, r.y);
            v = make_float2(r.z, r.w);
        }
        pos[idx] = sm;  // idx = e * mpc + j
        if (e == 0) vel[j] = v;
    }
    __syncthreads();
    // ---- epsilon: any marker of the r_max-1 cilia to the left within 1 unit on both axes ----
    {
        const int j = t % mpc, part = t / mpc, nparts = 1024 / mpc;  // 10 scanners a marker at 96
        if (part < nparts) {
            const float2 sm = pos[j];
            bool close = false;
            for (int idx = part; idx < (r_max - 1) * mpc; idx += nparts) {
                const float2 o = pos[mpc + idx];
                close |= (fabsf(o.x - sm.x) < 1) && (fabsf(o.y - sm.y) < 1);
            }
            if (close) closef[j] = 1;
        }
    }
    // ---- spreading stencil: columns bx..bx+3, rows by..by+3 and their 1-D delta factors ----
    if (t < mpc * 8) {
        const int j = t >> 3, i = t & 7;
        const float2 sm = pos[j];
        const int bx = (int)ceilf(sm.x - 1.5f), by = (int)ceilf(sm.y - 1.5f);
        wgt[j][i] = (i < 4) ? delta_1d(fabsf((bx + i) - sm.x)) : delta_1d(fabsf((by + i - 4) - sm.y));
    }
    __syncthreads();
    if (t < mpc) {
        const long long k = (long long)slot * p.Ns + m * mpc + t;
        IBLB_CHK(slot >= 0 && slot < kKinBatch && m >= 0 && m < p.c_num);
        const float2 sm = pos[t];
        p.M.s[k] = sm;
        p.M.u_s[k] = vel[t];
        p.M.eps[k] = closef[t] ? 0 : 1;
        p.M.wx[k] = make_float4(wgt[t][0], wgt[t][1], wgt[t][2], wgt[t][3]);
        p.M.wy[k] = make_float4(wgt[t][4], wgt[t][5], wgt[t][6], wgt[t][7]);
        p.M.base[k] = make_int2((int)ceilf(sm.x - 1.5f), (int)ceilf(sm.y - 1.5f));
    }
    // bbox of the wrapped x positions (for the spreading tiles)
    if (t < 128) {
        const float sx = (t < mpc) ? pos[t].x : 0.f;
        float lo = (t < mpc) ? sx : 3.0e38f, hi = (t < mpc) ? sx : -3.0e38f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
            hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
        }
        if (lane == 0) {
            red_min[warp] = lo;
            red_max[warp] = hi;
        }
    }
    __syncthreads();
    if (t == 0)
        p.M.bbox[slot * p.c_num + m] =
            make_float2(fminf(fminf(red_min[0], red_min[1]), fminf(red_min[2], red_min[3])),
                        fmaxf(fmaxf(red_max[0], red_max[1]), fmaxf(red_max[2], red_max[3])));
}

__global__ void __launch_bounds__(1024, 2) ib_kinematics_batch(const IbParams p) {
    __shared__ KinematicsShared sh;
    if (p.mpc == kMarkersPerCilium) kinematics_body<kMarkersPerCilium>(p, sh);
    else kinematics_body<0>(p, sh);
}

// F_s of every marker (ImmersedBoundary.cu:94-133), the only per-marker work that needs the fluid:
// nine lanes per marker (one per support node, each pulling the node's nine populations), three
// markers per warp, 24 per block so that the markers spread over all SMs; the float accumulation in
// the fixed order i = 0..8 (Q6) runs over warp shuffles.  `p.M` points at the iteration's slot.
__global__ void __launch_bounds__(256) ib_interp(const IbParams p) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane / 9, i9 = lane - 9 * g;
    const int k = (blockIdx.x * 8 + warp) * 3 + g;
    const bool on = lane < 27 && k < p.Ns;
    double tx = 0., ty = 0.;
    if (on) interp_term(p.L, p.src, p.M.s[k], p.M.u_s[k], i9, tx, ty);
    float Fx = 0.f, Fy = 0.f;
#pragma unroll
    for (int i = 0; i < 9; i++) {  // float re-rounded after each term, fixed order (Q6)
        const int srcl = lane < 27 ? g * 9 + i : lane;
        const double ax = __shfl_sync(0xffffffffu, tx, srcl);
        const double ay = __shfl_sync(0xffffffffu, ty, srcl);
        Fx += ax;
        Fy += ay;
    }
    if (on && i9 == 0) p.M.F_s[k] = make_float2(Fx, Fy);
}

// CTA per 8x32-node tile of the band: force = sum over markers (ascending) of
// (double)(F_s*delta [float product]) * eps   (ImmersedBoundary.cu:196-197).
__global__ void __launch_bounds__(256) ib_spread(const IbParams p) {
    __shared__ unsigned relmask[64];      // relevant cilia bitmap (c_num <= 2048)
    __shared__ float4 list[kMaxMarkersPerCilium];
    __shared__ int wcount[4];
    const int mpc = p.mpc;
    const Lattice &L = p.L;
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int tx0 = blockIdx.x * kSpreadTileX, ty0 = blockIdx.y * kSpreadTileY;
    const int xl = tx0 + warp, y = ty0 + lane;  // this thread's node (local column)
    const int xg = xl + L.x0;                   // global column: no periodic image (Q5)
    const float xlo = (float)(tx0 + L.x0) - 1.5f, xhi = (float)(tx0 + L.x0 + kSpreadTileX - 1) + 1.5f;
    const float ylo = (float)ty0 - 1.5f, yhi = (float)(ty0 + kSpreadTileY - 1) + 1.5f;

    const int nwords = (p.c_num + 31) >> 5;
    for (int w0 = 0; w0 < nwords; w0 += 8) {  // 8 warps x 32 cilia per pass
        const int mcil = (w0 + warp) * 32 + lane;
        bool rel = false;
        if (mcil < p.c_num) {
            const float2 bb = p.M.bbox[mcil];
            rel = (bb.x <= xhi) && (bb.y >= xlo);
        }
        const unsigned bits = __ballot_sync(0xffffffffu, rel);
        if (lane == 0 && w0 + warp < nwords) relmask[w0 + warp] = bits;
    }
    __syncthreads();

    double fx = 0., fy = 0.;
    for (int w = 0; w < nwords; w++) {
        unsigned bits = relmask[w];
        while (bits) {
            const int mcil = (w << 5) + __ffs(bits) - 1;
            bits &= bits - 1;
            // ordered compaction of this cilium's markers that can touch the tile
            bool take = false;
            float4 e = make_float4(0.f, 0.f, 0.f, 0.f);
            if (t < mpc) {
                const int k = mcil * mpc + t;
                const float2 sm = p.M.s[k];
                take = (p.M.eps[k] != 0) && sm.x >= xlo && sm.x <= xhi && sm.y >= ylo && sm.y <= yhi;
                if (take) {
                    const float2 F = p.M.F_s[k];
                    e = make_float4(sm.x, sm.y, F.x, F.y);
                }
            }
            const unsigned tk = __ballot_sync(0xffffffffu, take);
            if (lane == 0 && warp < 4) wcount[warp] = __popc(tk);
            __syncthreads();
            const int n = wcount[0] + wcount[1] + wcount[2] + wcount[3];
            if (take) {
                int pos = __popc(tk & ((1u << lane) - 1u));
                if (warp >= 1) pos += wcount[0];
                if (warp >= 2) pos += wcount[1];
                if (warp >= 3) pos += wcount[2];
                IBLB_CHK(pos >= 0 && pos < kMaxMarkersPerCilium);
                list[pos] = e;
            }
            __syncthreads();
            for (int q = 0; q < n; q++) {
                const float4 mk = list[q];
                const float dx = fabsf(xg - mk.x), dy = fabsf(y - mk.y);
                if (dx <= 1.5f && dy <= 1.5f) {
                    const float del = delta_1d(dx) * delta_1d(dy);
                    fx += (double)(mk.z * del);
                    fy += (double)(mk.w * del);
                }
            }
            __syncthreads();  // list reused by the next cilium
        }
    }
    if (xl < L.W && y < p.band_rows) {
        IBLB_CHK(y < p.band_ld);
        p.band[(long long)xl * p.band_ld + y] = fx;
        p.band[(long long)(L.W + xl) * p.band_ld + y] = fy;
    }
}

IBLB_CHECKED_COUNTER(checked_count_ib)

}  // namespace iblb
