// libiblb: solver lifecycle and the time-step loop (replaces main.cu:389-587, 634-754, 817-934).
#include <cmath>
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "iblb_internal.h"

namespace iblb {

int cuda_fail(cudaError_t e, const char *what) {
    set_error(std::string(what) + ": " + cudaGetErrorString(e));
    cudaGetLastError();  // clear the sticky-free error state
    return IBLB_ECUDA;
}

namespace {

// which fine points can write marker j (main.cu:158-172 evaluated for every k in float, as the
// kernel does); the winner of the write race is the lowest k (measured on B200) unless PICK_HIGHEST_K.
void build_marker_table(std::vector<int> &k_of_j, bool lowest, int mpc, int fine_points,
                        double scale) {
    k_of_j.assign(mpc, -1);
    for (int k = 0; k < fine_points; k++) {
        const float arcl = 1. * k / fine_points;  // main.cu:100
        const float s2 = scale * arcl;            // main.cu:139 (the product of two floats is exact
                                                  // in double: same bits as the float multiply)
        for (int j = 0; j < mpc; j++) {
            const float b_length = j;
            if (fabsf(s2 - b_length) < 0.01) {
                if (k_of_j[j] < 0 || !lowest) k_of_j[j] = k;
            }
        }
    }
}

}  // namespace

LbParams make_lb_params(const iblb_solver *s, int src_idx) {
    LbParams q;
    memset(&q, 0, sizeof q);
    q.L = s->L;
    q.src = s->buf[src_idx];
    q.dst = s->buf[src_idx ^ 1];
    q.left = s->nb[0][src_idx ^ 1];
    q.right = s->nb[1][src_idx ^ 1];
    q.band = s->band;
    q.band_rows = s->band_rows;
    q.band_ld = s->band_ld;
    q.omega_plus = 1. / s->p.TAU;
    q.omega_minus = 1. / s->p.TAU2;
    q.kforce = 1. - 1. / (2. * s->p.TAU);
    {
        const double cs2 = kCs * kCs, k1 = q.kforce * (1.0 / 9.0), k2 = q.kforce * (1.0 / 36.0);
        q.k1I2 = k1 * (1.0 / cs2);
        q.k1I4 = k1 * (1.0 / (cs2 * cs2));
        q.k2I2 = k2 * (1.0 / cs2);
        q.k2I4 = k2 * (1.0 / (cs2 * cs2));
    }
    q.gx = s->p.gx;
    q.gy = s->p.gy;
    const bool half = s->p.c_num > 0 || (s->p.flags & IBLB_FLAG_HALF_FORCE_BODY);
    q.half = half ? 0.5 : 0.0;
    q.qacc = s->flux_on ? s->qacc : nullptr;
    q.qhist = s->flux_on ? s->qhist : nullptr;
    q.nq = s->nq;
    if (s->cap_off >= 0) {  // captured batches read the iteration number from the device
        q.it_ptr = s->it_dev;
        q.it = s->cap_off;
    } else {
        q.it_ptr = nullptr;
        q.it = (int)s->it;
    }
    q.xq = s->flux_on ? s->xq_local : -1000;
    q.col_begin = 0;
    q.col_count = s->L.W;
    q.split_at = 1 << 30;
    q.split_skip = 0;
    q.row_begin = 0;
    return q;
}

// marker arrays of iteration slot `slot` (F_s has a single slot)
MarkerArrays marker_slot(const iblb_solver *s, int slot) {
    MarkerArrays m = s->M;
    const size_t o = (size_t)slot * s->Ns;
    m.s += o; m.u_s += o; m.eps += o; m.wx += o; m.wy += o; m.base += o;
    m.bbox += (size_t)slot * s->p.c_num;
    return m;
}

IbParams make_ib_params(const iblb_solver *s, int src_idx, int slot) {
    IbParams q;
    memset(&q, 0, sizeof q);
    q.L = s->L;
    q.src = s->buf[src_idx];
    q.M = marker_slot(s, slot);
    q.coef = s->coef;
    q.c_num = (int)s->p.c_num;
    q.c_space = (int)s->p.c_space;
    q.T = (int)s->p.T;
    q.p_step = s->p.p_step;
    q.it = (int)s->it;
    q.it_off = 0;
    q.it_ptr = nullptr;
    q.Ns = s->Ns;
    q.mpc = s->mpc;
    q.scale = s->shape_scale;
    q.r_max = (int)(2 * s->mpc / (double)s->p.c_space);  // main.cu:185
    q.band = s->band;
    q.band_rows = s->band_rows;
    q.band_ld = s->band_ld;
    q.replay = s->replay_dev;
    q.replay_steps = s->replay_steps;
    q.replay_j = s->replay_j;
    return q;
}

// The staging buffer is used as two halves: while a kernel works on one (lb_init_t / lb_moments*),
// the copy stream moves the other over PCIe.
int stage_buffer(iblb_solver *s) {
    if (s->stage_dev) return IBLB_OK;
    IBLB_CUDA(cudaMalloc((void **)&s->stage_dev, sizeof(double) * kStageDoubles));
    IBLB_CUDA(cudaStreamCreateWithFlags(&s->stream_copy, cudaStreamNonBlocking));
    for (int h = 0; h < 2; h++) {
        IBLB_CUDA(cudaEventCreateWithFlags(&s->ev_copy[h], cudaEventDisableTiming));
        IBLB_CUDA(cudaEventCreateWithFlags(&s->ev_kern[h], cudaEventDisableTiming));
    }
    return IBLB_OK;
}

}  // namespace iblb

using namespace iblb;

// rho, u: HOST arrays in the reference layout (main.cu:469-483), or null for the initial state of
// main.cu:636-654.  Uploaded in chunks of rows through the staging buffer: no device copy of the
// whole field (24 B/node would not fit beside the populations of the largest grids).
static int init_state(iblb_solver *s, const double *rho, const double *u) {
    LbParams q = make_lb_params(s, 0);
    q.dst = s->buf[0];
    const int W = s->L.W, Y = s->L.YDIM;
    const size_t size = (size_t)W * Y;
    if (rho == nullptr) {
        lb_init<<<(unsigned)((size + 255) / 256), 256, 0, s->stream>>>(q, nullptr, nullptr,
                                                                       s->p.u0_amp, 0, Y);
        IBLB_CUDA(cudaGetLastError());
    } else {
        int rc = stage_buffer(s);
        if (rc) return rc;
        const size_t half = kStageDoubles / 2;
        const int rows = (int)std::max<size_t>(1, std::min<size_t>(Y, half / (3 * (size_t)W)));
        cudaStream_t cp = s->stream_copy;
        // the copies start behind whatever the main stream still does with the staging buffer
        IBLB_CUDA(cudaEventRecord(s->ev_kern[0], s->stream));
        IBLB_CUDA(cudaEventRecord(s->ev_kern[1], s->stream));
        int h = 0;
        for (int r0 = 0; r0 < Y; r0 += rows, h ^= 1) {
            const int nr = std::min(rows, Y - r0);
            const size_t cs = (size_t)W * nr, o = (size_t)W * r0;
            double *d = s->stage_dev + h * half;
            IBLB_CUDA(cudaStreamWaitEvent(cp, s->ev_kern[h], 0));  // the kernel that last read this half
            IBLB_CUDA(cudaMemcpyAsync(d, rho + o, sizeof(double) * cs, cudaMemcpyHostToDevice, cp));
            IBLB_CUDA(cudaMemcpyAsync(d + cs, u + o, sizeof(double) * cs, cudaMemcpyHostToDevice, cp));
            IBLB_CUDA(cudaMemcpyAsync(d + 2 * cs, u + size + o, sizeof(double) * cs, cudaMemcpyHostToDevice, cp));
            IBLB_CUDA(cudaEventRecord(s->ev_copy[h], cp));
            IBLB_CUDA(cudaStreamWaitEvent(s->stream, s->ev_copy[h], 0));
            lb_init_t<<<dim3((W + 31) / 32, (nr + 31) / 32), 256, 0, s->stream>>>(q, d, d + cs, r0, nr);
            IBLB_CUDA(cudaGetLastError());
            IBLB_CUDA(cudaEventRecord(s->ev_kern[h], s->stream));
        }
    }
    // ghost columns of buffer 0 (own opposite edge, or the connected neighbours)
    dim3 g((s->L.YDIM + 255) / 256, s->L.G, 9);
    halo_push<<<g, 256, 0, s->stream>>>(s->L, s->buf[0], s->nb[0][0], s->nb[1][0]);
    IBLB_CUDA(cudaGetLastError());
    if (s->band) IBLB_CUDA(cudaMemsetAsync(s->band, 0, sizeof(double) * 2 * s->L.W * s->band_ld, s->stream));
    IBLB_CUDA(cudaMemsetAsync(s->qacc, 0, sizeof(double) * s->nq, s->stream));
    if (s->it_dev) IBLB_CUDA(cudaMemsetAsync(s->it_dev, 0, sizeof(int), s->stream));
    s->cur = 0;
    s->ib_stale = false;
    s->kin_base = -1;
    s->lists_base = -1;
    s->flag_base += (int)s->it;  // all slabs reset together: later flag values stay above old ones
    s->it = 0;
    // a new state clears an earlier device-side failure (tile-list overflow, slab timeout)
    if (s->err_dev) IBLB_CUDA(cudaMemsetAsync(s->err_dev, 0, sizeof(int), s->stream));
    if (s->flags) {
        slab_unpoison<<<1, 32, 0, s->stream>>>(s->flags, s->flag_base);
        IBLB_CUDA(cudaGetLastError());
    }
    s->hist_from = 0;
    IBLB_CUDA(cudaStreamSynchronize(s->stream));
    return IBLB_OK;
}

extern "C" {

int iblb_create(const iblb_params *pp, iblb_solver **out) {
    if (!pp || !out) {
        set_error("iblb_create: null argument");
        return IBLB_EINVAL;
    }
    *out = nullptr;
    const iblb_params &p = *pp;
    if (p.XDIM < 4 || p.YDIM < 4 || (p.YDIM & 1)) {
        set_error("iblb_create: XDIM, YDIM must be >= 4 and YDIM even");
        return IBLB_EINVAL;
    }
    // cilium geometry: the reference's hard-wired values unless the caller gives others
    const int mpc = p.LENGTH ? (int)p.LENGTH : kMarkersPerCilium;
    const int fine_points = p.fine_points ? (int)p.fine_points : kFinePoints;
    const double shape_scale = p.shape_scale > 0. ? p.shape_scale : kShapeScale;
    const int flux_back = p.flux_back ? (int)p.flux_back : 5;
    if (p.c_num > 0 && (p.XDIM != p.c_num * p.c_space || p.c_num > 2048 || p.c_space == 0 ||
                        2 * mpc / (int)p.c_space > kMaxLeftCilia)) {
        set_error("iblb_create: with cilia XDIM must equal c_num*c_space, c_num <= 2048, "
                  "2*LENGTH/c_space <= 16");
        return IBLB_EINVAL;
    }
    if (p.c_num > 0 && (mpc < 1 || mpc > kMaxMarkersPerCilium || fine_points < mpc)) {
        set_error("iblb_create: LENGTH (markers per cilium) must be 1..128 and fine_points >= LENGTH");
        return IBLB_EINVAL;
    }
    if (p.flux_back > p.XDIM) {  // (the default 5 on a narrower channel: no flux plane, as before)
        set_error("iblb_create: flux_back must be in 1..XDIM");
        return IBLB_EINVAL;
    }
    if (!(p.TAU > 0.5) || !(p.TAU2 > 0.5)) {
        set_error("iblb_create: relaxation times must exceed 1/2 (call iblb_params_derive)");
        return IBLB_EINVAL;
    }
    const bool slab = p.slab_w != 0 && p.slab_w != p.XDIM;
    if (slab && (p.slab_x0 + p.slab_w > p.XDIM)) {
        set_error("iblb_create: slab exceeds the channel");
        return IBLB_EINVAL;
    }
    if (p.device >= 0) IBLB_CUDA(cudaSetDevice(p.device));
    iblb_solver *s = new iblb_solver();  // value-initialised: all members zero
    s->p = p;
    {
        cudaError_t e = cudaGetDevice(&s->device);
        if (e != cudaSuccess) {
            delete s;
            return cuda_fail(e, "cudaGetDevice");
        }
    }
    Lattice &L = s->L;
    L.XDIM = (int)p.XDIM;
    L.YDIM = (int)p.YDIM;
    L.W = slab ? (int)p.slab_w : (int)p.XDIM;
    L.x0 = slab ? (int)p.slab_x0 : 0;
    L.G = (slab && p.c_num > 0) ? 3 : 1;
    L.ld = ((int)p.YDIM + 15) / 16 * 16;
    L.plane = (long long)(L.W + 2 * L.G) * L.ld;
    if (L.W < 2 * L.G + 2) {
        delete s;
        set_error("iblb_create: slab too narrow");
        return IBLB_EINVAL;
    }
    s->vec2 = (p.YDIM % 64 == 0);
    s->vec4 = true;
    const bool dev_knobs = (p.flags & IBLB_FLAG_DEV_KNOBS) != 0;  // development A/B variables honoured
    if (const char *e = dev_knobs ? getenv("IBLB_V4") : nullptr) s->vec4 = atoi(e) != 0;
    s->mpc = mpc;
    s->fine_points = fine_points;
    s->shape_scale = shape_scale;
    s->Ns = mpc * (int)p.c_num;

    int rc = IBLB_OK;
#define TRY(call)                                   \
    do {                                            \
        cudaError_t e__ = (call);                   \
        if (e__ != cudaSuccess && rc == IBLB_OK) rc = cuda_fail(e__, #call); \
    } while (0)
    const size_t pop_bytes = sizeof(double) * 9 * (size_t)L.plane;
    TRY(cudaMalloc((void **)&s->buf[0], pop_bytes));
    if (rc == IBLB_OK) TRY(cudaMalloc((void **)&s->buf[1], pop_bytes));
    if (rc == IBLB_OK) TRY(cudaMemset(s->buf[0], 0, pop_bytes));
    if (rc == IBLB_OK) TRY(cudaMemset(s->buf[1], 0, pop_bytes));
    s->nq = s->vec2 ? (int)p.YDIM / 64 : 1;
    TRY(cudaMalloc((void **)&s->qacc, sizeof(double) * s->nq));
    TRY(cudaMalloc((void **)&s->qhist, sizeof(double) * s->nq * (size_t)kFluxHist));
    s->cap_off = -1;
    TRY(cudaMallocHost((void **)&s->q_host, sizeof(double) * (s->nq + 1)));
    if (!s->vec2) TRY(cudaMalloc((void **)&s->qpart, sizeof(double) * p.YDIM));
    TRY(cudaMalloc((void **)&s->flags, kSlabFlags * sizeof(int)));
    if (rc == IBLB_OK) TRY(cudaMemset(s->flags, 0, kSlabFlags * sizeof(int)));
    TRY(cudaMalloc((void **)&s->err_dev, sizeof(int)));
    if (rc == IBLB_OK) TRY(cudaMemset(s->err_dev, 0, sizeof(int)));
    TRY(cudaMalloc((void **)&s->it_dev, sizeof(int)));
    {
        const char *e = getenv("IBLB_SLAB_TIMEOUT_S");
        // long enough for what a rank legitimately does between two steps while its neighbours
        // wait (a multi-GB checkpoint with fsync, gathering and writing a BigData snapshot)
        const double sec = e ? atof(e) : 900.;
        s->slab_timeout_ns = (long long)((sec > 0. ? sec : 900.) * 1e9);
    }
    TRY(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
    TRY(cudaEventCreate(&s->ev0));
    TRY(cudaEventCreate(&s->ev1));
    if (p.c_num > 0 && rc == IBLB_OK) {
        // marker arrays: kKinBatch iteration slots (stride Ns); F_s only for the current iteration
        const size_t nm = (size_t)s->Ns * kKinBatch;
        TRY(cudaMalloc((void **)&s->M.s, sizeof(float2) * nm));
        TRY(cudaMalloc((void **)&s->M.u_s, sizeof(float2) * nm));
        TRY(cudaMalloc((void **)&s->M.F_s, sizeof(float2) * s->Ns));
        TRY(cudaMalloc((void **)&s->M.eps, sizeof(int) * nm));
        TRY(cudaMalloc((void **)&s->M.bbox, sizeof(float2) * p.c_num * kKinBatch));
        TRY(cudaMalloc((void **)&s->M.wx, sizeof(float4) * nm));
        TRY(cudaMalloc((void **)&s->M.wy, sizeof(float4) * nm));
        TRY(cudaMalloc((void **)&s->M.base, sizeof(int2) * nm));
        TRY(cudaMalloc((void **)&s->coef, sizeof(float) * mpc * 28));
        TRY(cudaMalloc((void **)&s->k_of_j_dev, sizeof(int) * mpc));
        if (rc == IBLB_OK) {
            TRY(cudaMemset(s->M.s, 0, sizeof(float2) * nm));
            TRY(cudaMemset(s->M.u_s, 0, sizeof(float2) * nm));
            TRY(cudaMemset(s->M.F_s, 0, sizeof(float2) * s->Ns));
            TRY(cudaMemset(s->M.eps, 0, sizeof(int) * nm));
            build_marker_table(s->k_of_j, (p.flags & IBLB_FLAG_PICK_HIGHEST_K) == 0, mpc,
                               fine_points, shape_scale);
            for (int j = 0; j < mpc && rc == IBLB_OK; j++)
                if (s->k_of_j[j] < 0) {
                    // main.cu:158-172 would leave this marker at the origin for ever
                    set_error("iblb_create: no fine point falls on marker " + std::to_string(j) +
                              " (|shape_scale*k/fine_points - j| < 0.01 has no solution)");
                    rc = IBLB_EINVAL;
                }
        }
        if (rc == IBLB_OK) {
            TRY(cudaMemcpy(s->k_of_j_dev, s->k_of_j.data(), sizeof(int) * mpc,
                           cudaMemcpyHostToDevice));
            ib_coef<<<1, 128, 0, s->stream>>>(s->k_of_j_dev, s->coef, mpc, fine_points,
                                              (p.flags & IBLB_FLAG_MUCUS) ? 1 : 0);
            TRY(cudaGetLastError());
            // envelope of every marker over the whole beat cycle, relative to its cilium's base
            // (main.cu:137-145): sampled at kEnvSamples phases, widened by the distance a marker can
            // move between two samples (bound on d/dphase from the coefficients) — rigorous
            std::vector<float> cf((size_t)mpc * 28);
            TRY(cudaMemcpyAsync(cf.data(), s->coef, sizeof(float) * cf.size(),
                                cudaMemcpyDeviceToHost, s->stream));
            TRY(cudaStreamSynchronize(s->stream));
            constexpr int kEnvSamples = 2048;
            const double two_pi = 6.283185307179586;
            std::vector<double> cs(kEnvSamples * 12);
            for (int k = 0; k < kEnvSamples; k++)
                for (int n = 1; n < 7; n++) {
                    cs[k * 12 + 2 * (n - 1)] = cos(n * two_pi * k / kEnvSamples);
                    cs[k * 12 + 2 * (n - 1) + 1] = sin(n * two_pi * k / kEnvSamples);
                }
            s->env.assign(4 * (size_t)mpc, 0.f);
            double rx = 0., ry = 0.;
            for (int j = 0; j < mpc && rc == IBLB_OK; j++) {
                const float *c = cf.data() + j * 28;
                double lo[2] = {1e30, 1e30}, hi[2] = {-1e30, -1e30}, lip[2] = {0., 0.};
                for (int a = 0; a < 2; a++) {
                    for (int n = 1; n < 7; n++)
                        lip[a] += shape_scale * n * two_pi * hypot((double)c[2 * n + a], (double)c[14 + 2 * n + a]);
                    for (int k = 0; k < kEnvSamples; k++) {
                        double v = shape_scale * c[a] * 0.5;
                        for (int n = 1; n < 7; n++)
                            v += shape_scale * (c[2 * n + a] * cs[k * 12 + 2 * (n - 1)] +
                                         c[14 + 2 * n + a] * cs[k * 12 + 2 * (n - 1) + 1]);
                        lo[a] = v < lo[a] ? v : lo[a];
                        hi[a] = v > hi[a] ? v : hi[a];
                    }
                    const double margin = 0.5 * lip[a] / kEnvSamples + 0.01;  // + float rounding
                    lo[a] -= margin;
                    hi[a] += margin;
                }
                s->env[4 * j + 0] = (float)lo[0];
                s->env[4 * j + 1] = (float)hi[0];
                s->env[4 * j + 2] = (float)(lo[1] + 1.);  // + the translation of main.cu:198
                s->env[4 * j + 3] = (float)(hi[1] + 1.);
                rx = std::max(rx, std::max(fabs(lo[0]), fabs(hi[0])));
                ry = std::max(ry, hi[1] + 1.);
            }
            s->reach_x = (float)rx;
            s->reach_y = (float)ry;
            // the force band: the rows a marker's delta support can ever reach, in whole row tiles
            // (the reference's cilia stay below y = 96: two tiles)
            s->band_rows = 64 * (int)((ry + 1.5 + 1. + 63.) / 64.);
            s->band_ld = s->band_rows;
            if (s->band_rows > (int)p.YDIM && rc == IBLB_OK) {
                set_error("iblb_create: the cilia reach y = " + std::to_string(ry) +
                          ": YDIM must be at least " + std::to_string(s->band_rows));
                rc = IBLB_EINVAL;
            }
            if (rc == IBLB_OK)
                TRY(cudaMalloc((void **)&s->band, sizeof(double) * 2 * (size_t)L.W * s->band_ld));
        }
    }
#undef TRY
    if (rc != IBLB_OK) {
        std::string keep = iblb_last_error();
        iblb_destroy(s);
        set_error(keep);
        return rc;
    }
    // flux plane x = XDIM-5 (ImmersedBoundary.cu:259); accumulated only where `spread` runs in
    // the reference (cilia present) or the half-force extension asks for it
    const int xq = (int)p.XDIM - flux_back - L.x0;
    s->xq_local = xq;
    s->flux_on = (p.c_num > 0 || (p.flags & IBLB_FLAG_HALF_FORCE_BODY)) && xq >= 0 && xq < L.W;
    // spreading fused into the LB kernel: whole channel on this GPU, vector path; in every mode the
    // markers must stay inside the force band (rigorous bound from the coefficient table)
    if (p.c_num > 0) {
        if (s->reach_y + 1.5f >= (float)s->band_rows) {
            iblb_destroy(s);
            set_error("iblb_create: cilia reach above the force band");
            return IBLB_EINVAL;
        }
        s->fuse_ok = !slab && s->vec2 && !(p.flags & IBLB_FLAG_NO_FUSE);
        if (s->fuse_ok) {
            s->band_tiles = s->band_rows / 64;
            {
                // Tile width: the step of a small grid ends with its busiest SM, so take the width
                // (8 or 6 columns = warps per CTA) that puts the fewest nodes on one SM; ties and
                // grids that fill the machine many times over keep 8.
                cudaDeviceProp prop;
                int sms = 148;
                if (cudaGetDeviceProperties(&prop, s->device) == cudaSuccess) sms = prop.multiProcessorCount;
                const int rows = (L.YDIM / 64 > 8) ? s->band_tiles : L.YDIM / 64;
                long long best = -1;
                for (int tw : {8, 6}) {
                    const long long ctas = (long long)((L.W + tw - 1) / tw) * rows;
                    if (tw != 8 && (long long)((L.W + 7) / 8) * rows >= 4 * sms) break;
                    const long long load = ((ctas + sms - 1) / sms) * tw;
                    if (best < 0 || load < best) {
                        best = load;
                        s->tile_w = tw;
                    }
                }
                if (const char *e = dev_knobs ? getenv("IBLB_TILE_W") : nullptr) {
                    const int v = atoi(e);
                    if (v == 6 || v == 8) s->tile_w = v;
                }
            }
            s->tiles_x = (L.W + s->tile_w - 1) / s->tile_w;
            {
                // Block order of lb_ib_tile_step (fused_step.cu): band tiles first once the grid
                // has more CTAs than SMs (C3: 7.4 -> 6.8 us, 1152x192: 13.0 -> 9.9 us), row tile
                // fastest below, where SMs stay empty (C1: no difference to 5 %, box dependent).
                cudaDeviceProp prop;
                int sms = 148;
                if (cudaGetDeviceProperties(&prop, s->device) == cudaSuccess) sms = prop.multiProcessorCount;
                const int rows = (L.YDIM / 64 > 8) ? s->band_tiles : L.YDIM / 64;
                s->band_first = (long long)s->tiles_x * rows > sms ? 1 : 0;
                if (const char *e = dev_knobs ? getenv("IBLB_BAND_FIRST") : nullptr)
                    s->band_first = atoi(e) != 0;
            }
            // Capacity of the per-tile plans, decided here and not at run time: an upper bound on
            // the markers / cilia that can ever be within 1.5 of a tile, from the per-marker
            // envelopes (positions are wrapped into [0, XDIM], main.cu:193-196).  A carpet too
            // dense for kTileCap / kBandMaxRel runs the unfused chain (same results bit for bit).
            {
                const int tw = s->tile_w, ntx = s->tiles_x, nty = s->band_tiles;
                std::vector<int> cnt((size_t)ntx * nty, 0), ccnt((size_t)ntx, 0), seen((size_t)ntx, -1);
                for (int m = 0; m < (int)p.c_num; m++) {
                    const double base = (double)p.c_space * (m + 0.5);
                    for (int j = 0; j < mpc; j++) {
                        const double xl = base + s->env[4 * j] - 1.5, xh = base + s->env[4 * j + 1] + 1.5;
                        const double yl = s->env[4 * j + 2] - 1.5, yh = s->env[4 * j + 3] + 1.5;
                        const int ty_lo = std::max(0, (int)floor(yl / 64.)), ty_hi = std::min(nty - 1, (int)floor(yh / 64.));
                        for (int im = -1; im <= 1; im++) {  // periodic images of the envelope
                            const double a = xl + im * (double)L.XDIM, b = xh + im * (double)L.XDIM;
                            const int tx_lo = std::max(0, (int)floor(a / tw)), tx_hi = std::min(ntx - 1, (int)floor(b / tw));
                            for (int tx = tx_lo; tx <= tx_hi; tx++) {
                                for (int ty = ty_lo; ty <= ty_hi; ty++) cnt[(size_t)ty * ntx + tx]++;
                                if (seen[tx] != m) { seen[tx] = m; ccnt[tx]++; }
                            }
                        }
                    }
                }
                int worst = 0, worst_c = 0;
                for (int v : cnt) worst = std::max(worst, v);
                for (int v : ccnt) worst_c = std::max(worst_c, v);
                s->tile_bound = worst;
                if (worst > kTileCap || worst_c > kBandMaxRel) s->fuse_ok = false;
            }
        }
        if (s->fuse_ok) {
            const size_t nt = (size_t)kKinBatch * s->band_tiles * s->tiles_x;
            cudaError_t e = tile_step_prepare();
            if (e == cudaSuccess) e = cudaMalloc((void **)&s->prog, sizeof(TileProgram) * nt);
            if (e == cudaSuccess) e = cudaMemset(s->prog, 0, sizeof(TileProgram) * nt);
#ifdef IBLB_TILE_TIMING
            const size_t nd = (size_t)(L.YDIM / 64) * s->tiles_x * 16;
            if (e == cudaSuccess) e = cudaMalloc((void **)&s->dbg_dev, sizeof(long long) * nd);
            if (e == cudaSuccess) e = cudaMemset(s->dbg_dev, 0, sizeof(long long) * nd);
#endif
            if (e != cudaSuccess) {
                iblb_destroy(s);
                return cuda_fail(e, "cudaMalloc(tile lists)");
            }
        }
    }
    // a lone solver is periodic with itself
    for (int b = 0; b < 2; b++) {
        Remote self{s->buf[b], L.plane, L.ld, L.G, L.W};
        s->nb[0][b] = self;
        s->nb[1][b] = self;
    }
    s->self_left = s->self_right = true;
    s->nb_flag[0] = &s->flags[1];
    s->nb_flag[1] = &s->flags[0];
    s->nb_rel[0] = &s->flags[3];
    s->nb_rel[1] = &s->flags[2];
    rc = init_state(s, nullptr, nullptr);
    if (rc != IBLB_OK) {
        std::string keep = iblb_last_error();
        iblb_destroy(s);
        set_error(keep);
        return rc;
    }
    *out = s;
    return IBLB_OK;
}

int iblb_destroy(iblb_solver *s) {
    if (!s) return IBLB_OK;
    cudaSetDevice(s->device);
    if (s->stream) cudaStreamSynchronize(s->stream);
#ifdef IBLB_TILE_TIMING
    if (s->dbg_dev) {  // phase stamps of the LAST lb_ib_tile_step launch, per CTA (tools/tile_timing.sh)
        const int ty = s->L.YDIM / 64 > 8 ? s->band_tiles : s->L.YDIM / 64;
        const size_t nc = (size_t)ty * s->tiles_x;
        std::vector<long long> h(nc * 16);
        cudaMemcpy(h.data(), s->dbg_dev, sizeof(long long) * nc * 16, cudaMemcpyDeviceToHost);
        static const char *name[13] = {"entry", "prologue done", "wait returned", "loads issued",
                                       "ring moments", "bulk arrived+shift", "own moments",
                                       "barrier 1", "F_s", "barrier 2", "spread", "collide", "stores issued"};
        for (int cls = 0; cls < 3; cls++) {  // tiles without markers / with markers / the 10 % with most
            std::vector<size_t> sel;
            long long nmax = 0;
            for (size_t c = 0; c < nc; c++) nmax = std::max(nmax, h[c * 16 + 14]);
            for (size_t c = 0; c < nc; c++) {
                const long long n = h[c * 16 + 14];
                if ((cls == 0 && n == 0) || (cls == 1 && n > 0) || (cls == 2 && n > 0 && n * 10 >= nmax * 7))
                    sel.push_back(c);
            }
            if (sel.empty()) continue;
            fprintf(stderr, "tile timing, %zu CTAs %s, cycles after the wait returned (thread 0), avg:\n",
                    sel.size(), cls == 0 ? "without markers" : cls == 1 ? "with markers" : "with >= 70 % of the most markers");
            for (int i = 0; i < 13; i++) {
                double a = 0;
                int cnt = 0;
                for (size_t c : sel)
                    if (h[c * 16 + i] != 0) { a += (double)(h[c * 16 + i] - h[c * 16 + 2]); cnt++; }
                if (cnt) fprintf(stderr, "  %-20s %9.0f  (%d CTAs)\n", name[i], a / cnt, cnt);
            }
            double ex = 0, nn = 0, nr = 0;
            for (size_t c : sel) { ex += (double)(h[c * 16 + 13] - h[c * 16 + 2]); nn += h[c * 16 + 14]; nr += h[c * 16 + 15]; }
            fprintf(stderr, "  %-20s %9.0f   markers avg %.1f, ring nodes avg %.1f\n", "last thread exit", ex / sel.size(), nn / sel.size(), nr / sel.size());
        }
        cudaFree(s->dbg_dev);
    }
#endif
    if (s->graph_exec) cudaGraphExecDestroy(s->graph_exec);
    cudaFree(s->err_dev);
    for (int f = 0; f < 2; f++) {
        for (int b = 0; b < 2; b++)
            if (s->nb_mapped[f][b]) cudaIpcCloseMemHandle(s->nb_mapped[f][b]);
        if (s->nb_flag_mapped[f]) cudaIpcCloseMemHandle(s->nb_flag_mapped[f]);
    }
    cudaFree(s->buf[0]); cudaFree(s->buf[1]); cudaFree(s->qacc); cudaFree(s->qpart); cudaFree(s->qhist);
    if (s->q_host) cudaFreeHost(s->q_host);
    cudaFree(s->stage_dev);
    if (s->stream_copy) {
        cudaStreamSynchronize(s->stream_copy);
        cudaStreamDestroy(s->stream_copy);
        for (int h = 0; h < 2; h++) {
            cudaEventDestroy(s->ev_copy[h]);
            cudaEventDestroy(s->ev_kern[h]);
        }
    }
    cudaFree(s->flags); cudaFree(s->band); cudaFree(s->M.s); cudaFree(s->M.u_s);
    cudaFree(s->M.F_s); cudaFree(s->M.eps); cudaFree(s->M.bbox); cudaFree(s->coef);
    cudaFree(s->M.wx); cudaFree(s->M.wy); cudaFree(s->M.base);
    cudaFree(s->k_of_j_dev); cudaFree(s->it_dev);
    cudaFree(s->prog);
    cudaFree(s->replay_dev);
    if (s->ev0) cudaEventDestroy(s->ev0);
    if (s->ev1) cudaEventDestroy(s->ev1);
    if (s->ev_boundary) cudaEventDestroy(s->ev_boundary);
    if (s->ev_b) cudaEventDestroy(s->ev_b);
    if (s->ev_i) cudaEventDestroy(s->ev_i);
    if (s->ev_ib) cudaEventDestroy(s->ev_ib);
    if (s->stream_b) cudaStreamDestroy(s->stream_b);
    if (s->stream) cudaStreamDestroy(s->stream);
    delete s;
    return IBLB_OK;
}

int iblb_sync(iblb_solver *s) {
    if (!s) return IBLB_EINVAL;
    IBLB_CUDA(cudaSetDevice(s->device));
    IBLB_CUDA(cudaStreamSynchronize(s->stream));
    if (s->connected && s->err_dev) {  // a slab whose neighbour never arrived says so here
        int flag = 0;
        IBLB_CUDA(cudaMemcpy(&flag, s->err_dev, sizeof(int), cudaMemcpyDeviceToHost));
        if (flag == 3) {
            set_error("slab_wait timed out: a neighbouring slab never reached this iteration");
            return IBLB_ECUDA;
        }
    }
    return IBLB_OK;
}

unsigned iblb_iteration(const iblb_solver *s) { return s ? s->it : 0; }

int iblb_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

// Development: index violations counted on the device by an IBLB_CHECKED build (all solvers of the
// current device, since start); -1 when the library was built without the checks.
long long iblb_debug_violations(void) {
#ifdef IBLB_CHECKED
    cudaDeviceSynchronize();
    return (long long)(checked_count_lb() + checked_count_ib() + checked_count_fused());
#else
    return -1;
#endif
}

int iblb_marker_fine_index(const iblb_solver *s, unsigned j) {
    if (!s || j >= (unsigned)s->mpc) return -1;
    if (s->k_of_j.empty()) {
        std::vector<int> t;
        build_marker_table(t, (s->p.flags & IBLB_FLAG_PICK_HIGHEST_K) == 0, s->mpc, s->fine_points,
                           s->shape_scale);
        return t[j];
    }
    return s->k_of_j[j];
}

// Test hook, the product-side twin of the oracle's oracle_set_marker32_replay: marker `marker_j`
// of every cilium takes recorded values instead of computed ones.  The reference's define_filament
// has one genuinely timing-dependent write race (marker 32, main.cu:158-172); with the
// values a recorded run of the reference produced, the product can be compared with that run
// directly at the stated tolerance (tests/test_gpu_parity.py).
int iblb_set_marker_replay(iblb_solver *s, unsigned marker_j, const float *vals, unsigned n_steps) {
    if (!s || s->p.c_num == 0 || marker_j >= (unsigned)s->mpc) {
        set_error("iblb_set_marker_replay: bad argument");
        return IBLB_EINVAL;
    }
    IBLB_CUDA(cudaSetDevice(s->device));
    IBLB_CUDA(cudaStreamSynchronize(s->stream));
    cudaFree(s->replay_dev);
    s->replay_dev = nullptr;
    s->replay_steps = 0;
    if (vals && n_steps > 0) {
        const size_t nb = sizeof(float4) * (size_t)n_steps * s->p.c_num;
        IBLB_CUDA(cudaMalloc((void **)&s->replay_dev, nb));
        IBLB_CUDA(cudaMemcpy(s->replay_dev, vals, nb, cudaMemcpyHostToDevice));
        s->replay_steps = (int)n_steps;
        s->replay_j = (int)marker_j;
    }
    // marker slots, tile plans and captured launches hold the old kinematics
    s->kin_base = -1;
    s->lists_base = -1;
    if (s->graph_exec) {
        cudaGraphExecDestroy(s->graph_exec);
        s->graph_exec = nullptr;
    }
    return IBLB_OK;
}

int iblb_set_fields(iblb_solver *s, const double *rho, const double *u) {
    if (!s || !rho || !u) {
        set_error("iblb_set_fields: null argument");
        return IBLB_EINVAL;
    }
    IBLB_CUDA(cudaSetDevice(s->device));
    return init_state(s, rho, u);
}

}  // extern "C"
