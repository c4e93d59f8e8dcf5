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
__global__ void ib_coef(const int *k_of_j, float *coef) {
    const int j = threadIdx.x;
    if (j >= kMarkersPerCilium) return;
    const float arcl = 1. * k_of_j[j] / kFinePoints;  // main.cu:100
    for (int n = 0; n < 7; n++)
        for (int c = 0; c < 2; c++) {
            float a = 0.f, b = 0.f;
            for (int pw = 0; pw < 3; pw++) {
                a += kA[n + 14 * pw + 7 * c] * pow(arcl, pw + 1);
                b += kB[n + 14 * pw + 7 * c] * pow(arcl, pw + 1);
            }
            coef[j * 28 + 2 * n + c] = a;
            coef[j * 28 + 14 + 2 * n + c] = b;
        }
}

// One block (128 threads) per cilium: positions at `it` and `it-1`, velocity, wrap, bbox.
__global__ void __launch_bounds__(128) ib_kinematics(const IbParams p) {
    __shared__ __align__(16) double trig[24];  // [t][n][cos,sin]: t=0 current, t=1 previous
    __shared__ float red_min[4], red_max[4];
    const int m = blockIdx.x;
    const int t = threadIdx.x;
    const int T = p.T;
    const int it = p.it_ptr ? *p.it_ptr : p.it;
    if (t < 24) trig[t] = beat_trig(t % 12, it - t / 12, m, p.p_step, T);  // main.cu:143
    __syncthreads();
    float sx = 0.f;
    if (t < kMarkersPerCilium) {
        float2 sm, v;
        marker_state(p.coef + t * 28, reinterpret_cast<const double2 *>(trig),
                     reinterpret_cast<const double2 *>(trig + 12), m, it, p.c_space, p.c_num,
                     p.L.XDIM, sm, v);
        sx = sm.x;
        const int k = m * kMarkersPerCilium + t;
        p.M.s[k] = sm;
        p.M.u_s[k] = v;
    }
    // bbox of the wrapped x positions (for the spreading tiles)
    float lo = (t < kMarkersPerCilium) ? sx : 3.0e38f, hi = (t < kMarkersPerCilium) ? sx : -3.0e38f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if ((t & 31) == 0) {
        red_min[t >> 5] = lo;
        red_max[t >> 5] = hi;
    }
    __syncthreads();
    if (t == 0)
        p.M.bbox[m] = make_float2(fminf(fminf(red_min[0], red_min[1]), red_min[2]),
                                  fmaxf(fmaxf(red_max[0], red_max[1]), red_max[2]));
}

// Warp per marker: epsilon (main.cu:217-248), then F_s (ImmersedBoundary.cu:94-133) with the
// macroscopic values of the nine support nodes built from pulled populations
// (LatticeBoltzmann.cu:173-373 pull form + :396-405).
__global__ void __launch_bounds__(256) ib_interpolate(const IbParams p) {
    const int lane = threadIdx.x & 31;
    const int k = (int)(((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    if (k >= p.Ns) return;
    const Lattice &L = p.L;
    const float2 sm = p.M.s[k];
    const float2 us = p.M.u_s[k];

    // ---- epsilon: any marker of the r_max-1 cilia to the left within 1 unit on both axes ----
    const int m = k / kMarkersPerCilium;
    bool close = false;
    for (int r = 1; r < p.r_max; r++) {
        const int mm = (m - r < 0) ? (m - r + p.c_num) : (m - r);
        for (int l = lane; l < kMarkersPerCilium; l += 32) {
            const float2 o = p.M.s[l + mm * kMarkersPerCilium];
            close |= (fabsf(o.x - sm.x) < 1) && (fabsf(o.y - sm.y) < 1);
        }
    }
    const int eps = __any_sync(0xffffffffu, close) ? 0 : 1;

    // ---- lanes 0..8: support node i of the marker ----
    double tx = 0., ty = 0.;
    if (lane < 9) interp_term(L, p.src, sm, us, lane, tx, ty);
    // F_s is a float re-rounded after each of the nine terms, fixed order (Q6)
    float Fx = 0.f, Fy = 0.f;
#pragma unroll
    for (int i = 0; i < 9; i++) {
        const double ax = __shfl_sync(0xffffffffu, tx, i);
        const double ay = __shfl_sync(0xffffffffu, ty, i);
        Fx += ax;
        Fy += ay;
    }
    if (lane == 0) {
        p.M.F_s[k] = make_float2(Fx, Fy);
        p.M.eps[k] = eps;
        // graph mode: the kinematics of this step have read the counter (previous kernel)
        if (k == 0 && p.it_ptr) *p.it_ptr = *p.it_ptr + 1;
    }
}

// CTA per 8x32-node tile of the band: force = sum over markers (ascending) of
// (double)(F_s*delta [float product]) * eps   (ImmersedBoundary.cu:196-197).
__global__ void __launch_bounds__(256) ib_spread(const IbParams p) {
    __shared__ unsigned relmask[64];      // relevant cilia bitmap (c_num <= 2048)
    __shared__ float4 list[kMarkersPerCilium];
    __shared__ int wcount[4];
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
            if (t < kMarkersPerCilium) {
                const int k = mcil * kMarkersPerCilium + t;
                const float2 sm = p.M.s[k];
                take = (p.M.eps[k] != 0) && sm.x >= xlo && sm.x <= xhi && sm.y >= ylo && sm.y <= yhi;
                if (take) {
                    const float2 F = p.M.F_s[k];
                    e = make_float4(sm.x, sm.y, F.x, F.y);
                }
            }
            const unsigned tk = __ballot_sync(0xffffffffu, take);
            if (lane == 0 && warp < 3) wcount[warp] = __popc(tk);
            __syncthreads();
            const int n = wcount[0] + wcount[1] + wcount[2];
            if (take) {
                int pos = __popc(tk & ((1u << lane) - 1u));
                if (warp >= 1) pos += wcount[0];
                if (warp >= 2) pos += wcount[1];
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
        p.band[(long long)xl * p.band_ld + y] = fx;
        p.band[(long long)(L.W + xl) * p.band_ld + y] = fy;
    }
}

}  // namespace iblb
