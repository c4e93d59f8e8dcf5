// libiblb: the time-step loop (replaces main.cu:817-934) and the state getters
// (replace the D2H copies of main.cu:928, 944-982).
#include <algorithm>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "iblb_internal.h"

namespace iblb {
LbParams make_lb_params(const iblb_solver *s, int src_idx);
IbParams make_ib_params(const iblb_solver *s, int src_idx, int slot);
MarkerArrays marker_slot(const iblb_solver *s, int slot);
}  // namespace iblb
using namespace iblb;

namespace {

constexpr unsigned kGraphBatch = kKinBatch;  // iterations per captured batch (even: A/B swap back)

BandParams make_band_params(const iblb_solver *s, int src_idx, int slot) {
    BandParams q;
    memset(&q, 0, sizeof q);
    q.lb = make_lb_params(s, src_idx);
    q.M = marker_slot(s, slot);
    q.c_num = (int)s->p.c_num;
    q.Ns = s->Ns;
    q.mpc = s->mpc;
    q.band_rows = s->band_rows;
    q.tw = s->tile_w;
    q.tiles_x = s->tiles_x;
    q.prog = s->prog + (size_t)slot * s->band_tiles * s->tiles_x;
    q.err = s->err_dev;
#ifdef IBLB_TILE_TIMING
    q.dbg = s->dbg_dev;
#endif
    q.rows_lb = s->L.YDIM / 64;
    q.band_first = s->band_first;
    return q;
}

// spreading fused into the LB kernel: whole channel on this GPU, vector path
bool fused_path(const iblb_solver *s) {
    const bool coupled = s->connected && !(s->self_left && s->self_right);
    return s->p.c_num > 0 && s->fuse_ok && !coupled;
}

struct PhaseEvents {
    cudaEvent_t e[5];
};

// Launch the fused LB kernel over the column range encoded in q.
int launch_lb(iblb_solver *s, const LbParams &q, cudaStream_t st, unsigned long long &launches) {
    if (q.col_count <= 0) return IBLB_OK;
    const int rows = (q.row_end ? q.row_end : s->L.YDIM) - q.row_begin;
    if (s->vec2 && s->vec4 && q.band == nullptr && rows % 128 == 0 && q.col_count >= 64) {
        // four nodes per thread, 256-bit accesses: no force band on these rows
        const long long n = (long long)q.col_count * (rows / 4);
        lb_fused_v4<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(q);
        launches += 1;
    } else if (s->vec2) {
        const long long n = (long long)q.col_count * (rows / 2);
        lb_fused_v2<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(q);
        launches += 1;
    } else {
        const long long n = (long long)q.col_count * s->L.YDIM;
        lb_fused_v1<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(q, s->flux_on ? s->qpart : nullptr);
        launches += 1;
        const bool has_q = s->flux_on && s->xq_local >= 0;
        if (has_q) {
            // the flux column is covered by exactly one of the launches of a step
            int c = s->xq_local;
            bool covered;
            if (q.split_skip)
                covered = (c < q.split_at) || (c >= q.split_at + q.split_skip);
            else
                covered = c >= q.col_begin && c < q.col_begin + q.col_count;
            if (covered) {
                flux_reduce_v1<<<1, 256, 0, st>>>(s->qpart, s->L.YDIM, s->qacc, q.qhist, q.it_ptr, q.it);
                launches += 1;
            }
        }
    }
    IBLB_CUDA(cudaGetLastError());
    return IBLB_OK;
}

// One iteration.
//   whole channel on this GPU:  lb_ib_tile_step: ONE launch (2 on tall grids, where lb_fused_v2
//                               takes the rows above the force band), plus, once per kKinBatch
//                               iterations, ib_kinematics_batch + ib_tile_lists
//   x-slab coupled to others:   [slab wait] -> ib_kinematics_batch (1 in 32) -> ib_interp -> ib_spread -> lb_fused_v2 on the
//                               boundary columns -> signal -> lb_fused_v2 on the interior
// `pe`: time the phases separately (markers / spread / LB), which forces the unfused chain.
int enqueue_iteration(iblb_solver *s, cudaStream_t st, bool graph_mode, PhaseEvents *pe,
                      unsigned long long &launches, int graph_it_off = 0) {
    const int src = s->cur;
    s->cap_off = graph_mode ? graph_it_off : -1;  // make_lb_params: where the iteration number lives
    const bool coupled = s->connected && !(s->self_left && s->self_right);
    const bool local = coupled && s->peer[0] != nullptr;
    // Slabs coupled across processes run two streams: `sb` carries what the neighbours wait for
    // (flag wait -> marker kernels -> boundary columns -> signal), `st` the interior columns, which
    // read no ghost column and need not sit behind the wait.  Events order the two: the boundary
    // kernel needs the previous iteration's interior (column G-1 pulls from column G; the band),
    // the interior needs the previous iteration's boundary kernel and this iteration's spreading.
    const bool two = coupled && !local && !pe && !graph_mode && s->stream_b != nullptr;
    cudaStream_t sb = two ? s->stream_b : st;
    if (two) IBLB_CUDA(cudaStreamWaitEvent(st, s->ev_b, 0));  // boundary kernel of the PREVIOUS iteration
    if (local) {
        // same-process neighbours: their boundary kernels of the previous step (host lockstep)
        for (int f = 0; f < 2; f++)
            if (s->peer[f] != s && s->peer[f]->ev_boundary)
                IBLB_CUDA(cudaStreamWaitEvent(st, s->peer[f]->ev_boundary, 0));
    } else if (coupled) {
        slab_wait<<<1, 32, 0, sb>>>(s->flags, s->nb_rel[0], s->nb_rel[1],
                                    s->flag_base + (int)s->it, s->slab_timeout_ns, s->err_dev);
        launches += 1;
        if (two) IBLB_CUDA(cudaStreamWaitEvent(sb, s->ev_i, 0));  // interior of the previous iteration
    }
    if (pe) cudaEventRecord(pe->e[0], st);
    const bool fused = !pe && fused_path(s);
    const bool chain = s->pdl_chain;  // the previous kernel in the stream is lb_ib_tile_step
    s->pdl_chain = false;
    bool fresh_plan = false;          // this iteration launched what the fused prologue reads
    bool ib_pending = false;          // two streams: spreading of this iteration recorded in ev_ib
    int slot = 0;
    if (s->p.c_num > 0) {
        // marker kinematics, epsilon, spreading stencils: a batch of kKinBatch iterations at once
        bool need_kin;
        if (graph_mode) {
            slot = graph_it_off;
            need_kin = (graph_it_off == 0);
        } else {
            need_kin = s->kin_base < 0 || (long long)s->it < s->kin_base ||
                       (long long)s->it >= s->kin_base + kKinBatch;
            if (need_kin) s->kin_base = s->it;
            slot = (int)((long long)s->it - s->kin_base);
        }
        if (need_kin) {
            IbParams kp = make_ib_params(s, src, 0);
            if (graph_mode) {
                kp.it_ptr = s->it_dev;
                kp.it = 0;
            }
            ib_kinematics_batch<<<dim3(kp.c_num, kKinBatch), 1024, 0, sb>>>(kp);
            fresh_plan = true;
            launches += 1;
            IBLB_CUDA(cudaGetLastError());
        }
        if (fused) {
            // which markers reach which tile, for the same batch of iterations
            const bool need_lists = graph_mode ? need_kin : (s->lists_base != s->kin_base);
            if (need_lists) {
                BandParams lp = make_band_params(s, src, 0);
                ib_tile_lists<<<dim3(s->band_tiles, s->tiles_x, kKinBatch), 256, 0, st>>>(lp);
                fresh_plan = true;
                launches += 1;
                IBLB_CUDA(cudaGetLastError());
                if (!graph_mode) s->lists_base = s->kin_base;
            }
        }
        if (pe) cudaEventRecord(pe->e[1], st);
        IbParams ib = make_ib_params(s, src, slot);
        if (!fused) {
            ib_interp<<<(ib.Ns + 23) / 24, 256, 0, sb>>>(ib);
            launches += 1;
            IBLB_CUDA(cudaGetLastError());
        }
        if (pe) cudaEventRecord(pe->e[2], st);
        if (!fused) {
            dim3 g((s->L.W + kSpreadTileX - 1) / kSpreadTileX, s->band_rows / kSpreadTileY);
            ib_spread<<<g, 256, 0, sb>>>(ib);
            launches += 1;
            IBLB_CUDA(cudaGetLastError());
            s->ib_stale = false;
            if (two) {  // the interior kernel's band rows read the force band (waited for below)
                IBLB_CUDA(cudaEventRecord(s->ev_ib, sb));
                ib_pending = true;
            }
        }
    } else if (pe) {
        cudaEventRecord(pe->e[1], st);
        cudaEventRecord(pe->e[2], st);
    }
    if (pe) cudaEventRecord(pe->e[3], st);
    LbParams q = make_lb_params(s, src);
    if (fused) {
        BandParams bp = make_band_params(s, src, slot);
        const int tiles_y = s->L.YDIM / 64, band_tiles = s->band_rows / 64;
        // short channels: one launch for all row tiles; tall ones: band tiles here, the rest in
        // lb_fused_v2
        const bool split = tiles_y > 8;
        bp.rows_lb = split ? band_tiles : tiles_y;
        dim3 g((unsigned)(bp.rows_lb * s->tiles_x));  // one CTA per tile
        // Programmatic dependent launch behind the previous iteration's lb_ib_tile_step: the
        // marker CTAs' fluid-independent prologue overlaps that kernel's tail.  Not right behind the
        // kernels that write what the prologue reads (kinematics, tile lists), and not on tall
        // grids, where lb_fused_v2 of the previous iteration is the predecessor.
        cudaLaunchConfig_t cfg;
        memset(&cfg, 0, sizeof cfg);
        cfg.gridDim = g;
        cfg.blockDim = dim3(32 * s->tile_w);
        cfg.dynamicSmemBytes = tile_step_smem_bytes();
        cfg.stream = st;
        cudaLaunchAttribute attr;
        attr.id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr.val.programmaticStreamSerializationAllowed = 1;
        const bool pdl = !split && !fresh_plan && chain && !(s->p.flags & IBLB_FLAG_NO_PDL);
        cfg.attrs = &attr;
        cfg.numAttrs = pdl ? 1 : 0;
        IBLB_CUDA(cudaLaunchKernelEx(&cfg, lb_ib_tile_step, bp));
        s->pdl_chain = !split;
        launches += 1;
        if (split) {
            q.band = nullptr;
            q.row_begin = s->band_rows;
            int rc = launch_lb(s, q, st, launches);
            if (rc) return rc;
        }
        s->ib_stale = true;
    } else if (coupled) {
        const int G = s->L.G, W = s->L.W;
        LbParams qb = q;  // boundary columns [0,G) and [W-G,W)
        qb.col_begin = 0;
        qb.col_count = 2 * G;
        qb.split_at = G;
        qb.split_skip = W - 2 * G;
        int rc = launch_lb(s, qb, sb, launches);
        if (rc) return rc;
        if (local) {
            IBLB_CUDA(cudaEventRecord(s->ev_boundary, st));
        } else {
            slab_signal<<<1, 32, 0, sb>>>(s->nb_flag[0], s->nb_flag[1],
                                        s->flag_base + (int)s->it + 1, s->err_dev);
            launches += 1;
            if (two) IBLB_CUDA(cudaEventRecord(s->ev_b, sb));
        }
        LbParams qi = q;  // interior: never touches a ghost column
        qi.col_begin = G;
        qi.col_count = W - 2 * G;
        qi.left.G = 0;
        qi.right.G = 0;
        if (ib_pending && s->vec2 && s->band_rows < s->L.YDIM) {
            // the rows above the force band need nothing from the markers: they start at once
            // (four nodes per thread where the row count allows), beside interpolation and spreading
            // on the boundary stream; the band's rows follow the spreading
            LbParams qt = qi;
            qt.band = nullptr;
            qt.row_begin = s->band_rows;
            rc = launch_lb(s, qt, st, launches);
            if (rc) return rc;
            qi.row_end = s->band_rows;
        }
        if (ib_pending) IBLB_CUDA(cudaStreamWaitEvent(st, s->ev_ib, 0));
        rc = launch_lb(s, qi, st, launches);
        if (rc) return rc;
        if (two) IBLB_CUDA(cudaEventRecord(s->ev_i, st));
    } else {
        int rc = launch_lb(s, q, st, launches);
        if (rc) return rc;
    }
    if (pe) cudaEventRecord(pe->e[4], st);
    s->cur ^= 1;
    s->it += 1;
    return IBLB_OK;
}

// kGraphBatch iterations reading the iteration number from device memory, then it_advance
int build_graph(iblb_solver *s) {
    cudaGraph_t graph = nullptr;
    const int cur0 = s->cur;
    const unsigned it0 = s->it;
    const bool stale0 = s->ib_stale;
    unsigned long long dummy = 0;
    s->pdl_chain = false;
    IBLB_CUDA(cudaStreamBeginCapture(s->stream, cudaStreamCaptureModeThreadLocal));
    int rc = IBLB_OK;
    for (unsigned i = 0; i < kGraphBatch && rc == IBLB_OK; i++)
        rc = enqueue_iteration(s, s->stream, true, nullptr, dummy, (int)i);
    if (rc == IBLB_OK && s->it_dev) it_advance<<<1, 32, 0, s->stream>>>(s->it_dev, (int)kGraphBatch);
    s->cap_off = -1;
    cudaError_t e = cudaStreamEndCapture(s->stream, &graph);
    s->graph_launches = dummy + (s->it_dev ? 1 : 0);
    s->cur = cur0;
    s->it = it0;
    s->ib_stale = stale0;
    if (rc != IBLB_OK) {
        if (graph) cudaGraphDestroy(graph);
        return rc;
    }
    if (e != cudaSuccess) return cuda_fail(e, "cudaStreamEndCapture");
    e = cudaGraphInstantiate(&s->graph_exec, graph, 0);
    cudaGraphDestroy(graph);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGraphInstantiate");
    return IBLB_OK;
}

// lb_ib_tile_step keeps F_s in shared memory and consumes the spread force in registers: rebuild
// F_s and the force band of the last completed iteration with the stand-alone kernels when
// somebody asks for them
int materialise_ib(iblb_solver *s) {
    if (!s->ib_stale || s->p.c_num == 0 || s->it == 0) return IBLB_OK;
    const long long last = (long long)s->it - 1;
    if (s->kin_base < 0 || last < s->kin_base || last >= s->kin_base + kKinBatch) {
        // the marker slots do not hold the last iteration (right after iblb_load_checkpoint)
        IbParams kp = make_ib_params(s, s->cur ^ 1, 0);
        kp.it = (int)last;
        ib_kinematics_batch<<<dim3(kp.c_num, kKinBatch), 1024, 0, s->stream>>>(kp);
        IBLB_CUDA(cudaGetLastError());
        s->kin_base = last;
        s->lists_base = -1;
    }
    IbParams ib = make_ib_params(s, s->cur ^ 1, (int)(last - s->kin_base));
    ib_interp<<<(ib.Ns + 23) / 24, 256, 0, s->stream>>>(ib);
    IBLB_CUDA(cudaGetLastError());
    dim3 g((s->L.W + kSpreadTileX - 1) / kSpreadTileX, s->band_rows / kSpreadTileY);
    ib_spread<<<g, 256, 0, s->stream>>>(ib);
    IBLB_CUDA(cudaGetLastError());
    s->ib_stale = false;
    return IBLB_OK;
}

int check_device_flag(iblb_solver *s) {
    if (!s->err_dev) return IBLB_OK;
    int flag = 0;
    IBLB_CUDA(cudaMemcpyAsync(&flag, s->err_dev, sizeof(int), cudaMemcpyDeviceToHost, s->stream));
    IBLB_CUDA(cudaStreamSynchronize(s->stream));
    if (flag != 0) {
        set_error(flag == 1 ? "ib_tile_lists: more cilia overlap a tile than kBandMaxRel "
                              "(use IBLB_FLAG_NO_FUSE)"
                  : flag == 2 ? "ib_tile_lists: more markers touch a tile than kTileCap "
                                "(use IBLB_FLAG_NO_FUSE)"
                              : "slab_wait timed out: a neighbouring slab never reached this "
                                "iteration (rank lost, ranks stepped different counts, or host work "
                                "between steps outlasted IBLB_SLAB_TIMEOUT_S), or a neighbour "
                                "reported that failure; the state of this slab is no longer valid");
        return flag == 3 ? IBLB_ECUDA : IBLB_EINVAL;
    }
    return IBLB_OK;
}

}  // namespace

extern "C" {

int iblb_step(iblb_solver *s, unsigned n_steps) {
    if (!s) {
        set_error("iblb_step: null solver");
        return IBLB_EINVAL;
    }
    IBLB_CUDA(cudaSetDevice(s->device));
    cudaStream_t st = s->stream;
    unsigned long long launches = 0;
    s->pdl_chain = false;  // anything may have been enqueued since the last call
    IBLB_CUDA(cudaEventRecord(s->ev0, st));
    const bool coupled = s->connected && !(s->self_left && s->self_right);
    const bool two = coupled && s->peer[0] == nullptr && s->stream_b != nullptr;
    // the boundary stream starts behind everything enqueued on the main stream so far
    if (two) IBLB_CUDA(cudaEventRecord(s->ev_i, st));
    // small grids are launch-bound: replay captured batches of kGraphBatch iterations (captured
    // with buffer A as the source: cur == 0)
    const long long nodes = (long long)s->L.W * s->L.YDIM;
    const bool use_graph = !(s->p.flags & IBLB_FLAG_NO_GRAPH) && !coupled && nodes <= (1 << 22) &&
                           n_steps >= kGraphBatch;
    unsigned left = n_steps;
    if (use_graph) {
        if (s->cur != 0) {
            int rc = enqueue_iteration(s, st, false, nullptr, launches);
            if (rc) return rc;
            left--;
        }
        if (left >= kGraphBatch) {
            if (!s->graph_exec) {
                int rc = build_graph(s);
                if (rc) return rc;
            }
            if (s->it_dev)
                IBLB_CUDA(cudaMemcpyAsync(s->it_dev, &s->it, sizeof(int), cudaMemcpyHostToDevice, st));
            while (left >= kGraphBatch) {
                IBLB_CUDA(cudaGraphLaunch(s->graph_exec, st));
                s->it += kGraphBatch;
                left -= kGraphBatch;
                launches += s->graph_launches;
            }
            s->pdl_chain = false;  // a batch ends with it_advance
            if (s->p.c_num > 0) {
                s->ib_stale = fused_path(s);
                s->kin_base = (long long)s->it - kGraphBatch;
                s->lists_base = fused_path(s) ? s->kin_base : -1;
            }
        }
    }
    while (left > 0) {
        int rc = enqueue_iteration(s, st, false, nullptr, launches);
        if (rc) return rc;
        left--;
    }
    if (two) IBLB_CUDA(cudaStreamWaitEvent(st, s->ev_b, 0));  // join: getters use the main stream
    IBLB_CUDA(cudaEventRecord(s->ev1, st));
    s->timed = true;
    s->last_launches = launches;
    return IBLB_OK;
}

int iblb_last_step_seconds(iblb_solver *s, double *seconds) {
    if (!s || !seconds || !s->timed) {
        set_error("iblb_last_step_seconds: no timed step");
        return IBLB_EINVAL;
    }
    IBLB_CUDA(cudaEventSynchronize(s->ev1));
    float ms = 0.f;
    IBLB_CUDA(cudaEventElapsedTime(&ms, s->ev0, s->ev1));
    *seconds = 1e-3 * (double)ms;
    return IBLB_OK;
}

int iblb_last_step_launches(iblb_solver *s, unsigned long long *n) {
    if (!s || !n) return IBLB_EINVAL;
    *n = s->last_launches;
    return IBLB_OK;
}

int iblb_profile_phases(iblb_solver *s, unsigned reps, double *t_kin, double *t_interp,
                        double *t_spread, double *t_lb) {
    if (!s || reps == 0) return IBLB_EINVAL;
    IBLB_CUDA(cudaSetDevice(s->device));
    PhaseEvents pe;
    for (auto &e : pe.e) IBLB_CUDA(cudaEventCreate(&e));
    double acc[4] = {0, 0, 0, 0};
    unsigned long long launches = 0;
    int rc = IBLB_OK;
    for (unsigned r = 0; r < reps && rc == IBLB_OK; r++) {
        rc = enqueue_iteration(s, s->stream, false, &pe, launches);
        if (rc) break;
        cudaError_t e = cudaStreamSynchronize(s->stream);
        if (e != cudaSuccess) {
            rc = cuda_fail(e, "profile sync");
            break;
        }
        for (int i = 0; i < 4; i++) {
            float ms = 0.f;
            cudaEventElapsedTime(&ms, pe.e[i], pe.e[i + 1]);
            acc[i] += 1e-3 * ms;
        }
    }
    for (auto &e : pe.e) cudaEventDestroy(e);
    if (rc) return rc;
    if (t_kin) *t_kin = acc[0] / reps;
    if (t_interp) *t_interp = acc[1] / reps;
    if (t_spread) *t_spread = acc[2] / reps;
    if (t_lb) *t_lb = acc[3] / reps;
    return IBLB_OK;
}

// ---- getters: recompute the macroscopic state of the last completed iteration from the buffer
// that iteration pulled from (the step wrote the other one) ----
static int run_moments(iblb_solver *s, double *rho, double *u, double *f, double *force) {
    IBLB_CUDA(cudaSetDevice(s->device));
    {
        int rc = check_device_flag(s);
        if (rc == IBLB_OK) rc = materialise_ib(s);  // u and force need the force band
        if (rc) return rc;
    }
    // rows [r0, r0 + nr) at a time through the staging buffer: no device copy of a whole field
    {
        int rc = stage_buffer(s);
        if (rc) return rc;
    }
    const int W = s->L.W, Y = s->L.YDIM;
    const size_t size = (size_t)W * Y;
    const size_t per_node = (rho ? 1 : 0) + (u ? 2 : 0) + (f ? 9 : 0) + (force ? 2 : 0);
    // two halves of the staging buffer: the kernel fills one while the copy stream drains the other
    const size_t half = kStageDoubles / 2;
    const int rows = (int)std::max<size_t>(1, std::min<size_t>(Y, half / (per_node * W)));
    const int streamed = s->it > 0;
    LbParams lp = make_lb_params(s, streamed ? (s->cur ^ 1) : s->cur);
    cudaError_t e = cudaSuccess;
    const cudaMemcpyKind D2H = cudaMemcpyDeviceToHost;
    cudaStream_t cp = s->stream_copy;
    int h = 0;
    bool used[2] = {false, false};
    for (int r0 = 0; r0 < Y && e == cudaSuccess; r0 += rows, h ^= 1) {
        const int nr = std::min(rows, Y - r0);
        const size_t cs = (size_t)W * nr, o = (size_t)W * r0, cb = sizeof(double) * cs;
        double *d_rho = nullptr, *d_u = nullptr, *d_f = nullptr, *d_force = nullptr;
        double *q = s->stage_dev + h * half;
        if (rho) { d_rho = q; q += cs; }
        if (u) { d_u = q; q += 2 * cs; }
        if (f) { d_f = q; q += 9 * cs; }
        if (force) { d_force = q; q += 2 * cs; }
        if (used[h]) e = cudaStreamWaitEvent(s->stream, s->ev_copy[h], 0);  // this half has been drained
        if (e != cudaSuccess) break;
        if (f)
            lb_moments<<<(unsigned)((cs + 255) / 256), 256, 0, s->stream>>>(lp, streamed, r0, nr, d_rho,
                                                                            d_u, d_f, d_force);
        else
            lb_moments_t<<<dim3((W + 31) / 32, (nr + 31) / 32), 256, 0, s->stream>>>(
                lp, streamed, r0, nr, d_rho, d_u, d_force);
        e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaEventRecord(s->ev_kern[h], s->stream);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(cp, s->ev_kern[h], 0);
        if (e == cudaSuccess && rho) e = cudaMemcpyAsync(rho + o, d_rho, cb, D2H, cp);
        if (e == cudaSuccess && u) e = cudaMemcpyAsync(u + o, d_u, cb, D2H, cp);
        if (e == cudaSuccess && u) e = cudaMemcpyAsync(u + size + o, d_u + cs, cb, D2H, cp);
        if (e == cudaSuccess && f) e = cudaMemcpyAsync(f + 9 * o, d_f, 9 * cb, D2H, cp);
        if (e == cudaSuccess && force) e = cudaMemcpyAsync(force + o, d_force, cb, D2H, cp);
        if (e == cudaSuccess && force)
            e = cudaMemcpyAsync(force + size + o, d_force + cs, cb, D2H, cp);
        if (e == cudaSuccess) e = cudaEventRecord(s->ev_copy[h], cp);
        used[h] = true;
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(cp);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s->stream);
    if (e != cudaSuccess) return cuda_fail(e, "iblb_get_*");
    return IBLB_OK;
}

int iblb_get_fields(iblb_solver *s, double *rho, double *u) {
    if (!s || (!rho && !u)) {
        set_error("iblb_get_fields: null argument");
        return IBLB_EINVAL;
    }
    return run_moments(s, rho, u, nullptr, nullptr);
}

int iblb_get_force(iblb_solver *s, double *force) {
    if (!s || !force) return IBLB_EINVAL;
    return run_moments(s, nullptr, nullptr, nullptr, force);
}

int iblb_get_populations(iblb_solver *s, double *f) {
    if (!s || !f) return IBLB_EINVAL;
    return run_moments(s, nullptr, nullptr, f, nullptr);
}

int iblb_get_markers(iblb_solver *s, float *pos, float *vel, float *force, int *eps) {
    if (!s) return IBLB_EINVAL;
    if (s->Ns == 0) return IBLB_OK;
    IBLB_CUDA(cudaSetDevice(s->device));
    {
        int rc = check_device_flag(s);
        if (rc == IBLB_OK) rc = materialise_ib(s);
        if (rc) return rc;
    }
    IBLB_CUDA(cudaStreamSynchronize(s->stream));
    // marker state of the last completed iteration (slot of it-1; before the first step: zeros)
    const size_t nb = sizeof(float2) * s->Ns;
    const int slot = (s->it > 0 && s->kin_base >= 0) ? (int)((long long)s->it - 1 - s->kin_base) : 0;
    const MarkerArrays M = marker_slot(s, slot);
    if (pos) IBLB_CUDA(cudaMemcpy(pos, M.s, nb, cudaMemcpyDeviceToHost));
    if (vel) IBLB_CUDA(cudaMemcpy(vel, M.u_s, nb, cudaMemcpyDeviceToHost));
    if (force) IBLB_CUDA(cudaMemcpy(force, M.F_s, nb, cudaMemcpyDeviceToHost));
    if (eps) IBLB_CUDA(cudaMemcpy(eps, M.eps, sizeof(int) * s->Ns, cudaMemcpyDeviceToHost));
    return IBLB_OK;
}

int iblb_get_flux(iblb_solver *s, double *Q) {
    if (!s || !Q) return IBLB_EINVAL;
    IBLB_CUDA(cudaSetDevice(s->device));
    // the reference copies Q back every step (main.cu:928): one pinned staging buffer, one sync
    double *part = s->q_host;
    int *flag = reinterpret_cast<int *>(s->q_host + s->nq);
    *flag = 0;
    IBLB_CUDA(cudaMemcpyAsync(part, s->qacc, sizeof(double) * s->nq, cudaMemcpyDeviceToHost,
                              s->stream));
    if (s->err_dev)
        IBLB_CUDA(cudaMemcpyAsync(flag, s->err_dev, sizeof(int), cudaMemcpyDeviceToHost, s->stream));
    IBLB_CUDA(cudaStreamSynchronize(s->stream));
    if (*flag != 0) return check_device_flag(s);  // sets the message
    double q = 0.;
    for (int i = 0; i < s->nq; i++) q += part[i];  // fixed slot order
    *Q = q;
    return IBLB_OK;
}

// Q after each of the iterations first_it .. first_it + n - 1 (0-based), i.e. what the reference's
// per-step copy of d_Q (main.cu:928) would have delivered, without a synchronisation per step.
int iblb_get_flux_history(iblb_solver *s, unsigned first_it, unsigned n, double *Q) {
    if (!s || !Q || n == 0) return IBLB_EINVAL;
    if ((unsigned long long)first_it + n > s->it || s->it - first_it > (unsigned)kFluxHist ||
        first_it < s->hist_from) {
        set_error("iblb_get_flux_history: the solver keeps the last 4096 completed iterations "
                  "(since the last restart)");
        return IBLB_EINVAL;
    }
    IBLB_CUDA(cudaSetDevice(s->device));
    if (!s->flux_on) {  // no flux plane in this solver (no cilia, or another slab owns it)
        for (unsigned i = 0; i < n; i++) Q[i] = 0.;
        return iblb_sync(s);
    }
    std::vector<double> rows((size_t)n * s->nq);
    for (unsigned done = 0; done < n;) {  // the ring may wrap once
        const unsigned r0 = (first_it + done) % kFluxHist;
        const unsigned len = std::min<unsigned>(n - done, kFluxHist - r0);
        IBLB_CUDA(cudaMemcpyAsync(rows.data() + (size_t)done * s->nq, s->qhist + (size_t)r0 * s->nq,
                                  sizeof(double) * len * s->nq, cudaMemcpyDeviceToHost, s->stream));
        done += len;
    }
    IBLB_CUDA(cudaStreamSynchronize(s->stream));
    for (unsigned i = 0; i < n; i++) {
        double q = 0.;
        for (int k = 0; k < s->nq; k++) q += rows[(size_t)i * s->nq + k];  // fixed slot order
        Q[i] = q;
    }
    return IBLB_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// iblb_get_fields_static: the field getter for callers that must not allocate device memory.
//
// iblb_get_fields stages its result in a 128 MB device buffer that it cudaMallocs on the first
// call (stage_buffer, solver.cu).  That is no good for a host program which handed the whole GPU
// to its own arena at start-up (a late cudaMalloc then fails at the first output interval, maybe
// hours into a run), nor for one in which another host thread holds a stream capture open in the
// global capture mode (the runtime refuses cudaMalloc from every thread while it lasts, and the
// refusal breaks that capture).  Here the staging area is part of the library image instead: a
// file-scope device array of fixed size, reserved on every GPU that loads libiblb.so from the
// moment the module is loaded until the process exits.
//
// Host code must not hand `g_static_fields` itself to a kernel: in host code that name is a
// host-side stand-in of the symbol, not a device address.  static_fields_ptr() asks the runtime for
// the real address once per device (with that device current) and keeps it.
namespace {

// rho + u_x + u_y of the largest grid served: 3 * 2^18 doubles = 6 MiB per GPU
constexpr size_t kStaticFieldNodes = IBLB_STATIC_FIELDS_MAX_NODES;
static __device__ __align__(256) double g_static_fields[3 * kStaticFieldNodes];

constexpr int kStaticMaxDevices = 64;
double *g_static_ptr[kStaticMaxDevices];  // device address of g_static_fields on device d, or null
std::mutex g_static_mu;

// The caller has made `device` current.  The first call per device resolves the symbol (which may
// also load this module onto the device); later calls return the stored address.
int static_fields_ptr(int device, double **out) {
    std::lock_guard<std::mutex> lock(g_static_mu);
    double *ptr = (device >= 0 && device < kStaticMaxDevices) ? g_static_ptr[device] : nullptr;
    if (ptr == nullptr) {
        const cudaError_t e = cudaGetSymbolAddress((void **)&ptr, g_static_fields);
        if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_static_fields)");
        if (ptr == nullptr) {
            set_error("iblb_get_fields_static: the runtime returned no address for the staging array");
            return IBLB_ECUDA;
        }
        if (device >= 0 && device < kStaticMaxDevices) g_static_ptr[device] = ptr;
    }
    *out = ptr;
    return IBLB_OK;
}

}  // namespace

extern "C" int iblb_get_fields_static(iblb_solver *s, double *rho, double *u) {
    if (!s || (!rho && !u)) {
        set_error("iblb_get_fields_static: null argument");
        return IBLB_EINVAL;
    }
    const int W = s->L.W, Y = s->L.YDIM;
    const size_t size = (size_t)W * Y;
    if (size > kStaticFieldNodes) {
        set_error("iblb_get_fields_static: " + std::to_string(W) + " x " + std::to_string(Y) + " = " +
                  std::to_string(size) + " nodes exceed the fixed staging array of " +
                  std::to_string(kStaticFieldNodes) +
                  " nodes (IBLB_STATIC_FIELDS_MAX_NODES); use iblb_get_fields, which allocates");
        return IBLB_EINVAL;
    }
    IBLB_CUDA(cudaSetDevice(s->device));
    double *stage = nullptr;
    {
        int rc = static_fields_ptr(s->device, &stage);
        if (rc == IBLB_OK) rc = check_device_flag(s);
        if (rc == IBLB_OK) rc = materialise_ib(s);  // u needs the force band
        if (rc) return rc;
    }
    // One launch over all rows.  The kernel writes every element that is copied out below, so the
    // previous call's contents never show and nothing has to be cleared.
    double *d_rho = rho ? stage : nullptr;
    double *d_u = u ? stage + (rho ? size : 0) : nullptr;
    const int streamed = s->it > 0;
    const LbParams lp = make_lb_params(s, streamed ? (s->cur ^ 1) : s->cur);
    lb_moments_t<<<dim3((W + 31) / 32, (Y + 31) / 32), 256, 0, s->stream>>>(lp, streamed, 0, Y, d_rho,
                                                                          d_u, nullptr);
    cudaError_t e = cudaGetLastError();
    const size_t cb = sizeof(double) * size;
    if (e == cudaSuccess && rho) e = cudaMemcpyAsync(rho, d_rho, cb, cudaMemcpyDeviceToHost, s->stream);
    if (e == cudaSuccess && u) e = cudaMemcpyAsync(u, d_u, 2 * cb, cudaMemcpyDeviceToHost, s->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s->stream);
    if (e != cudaSuccess) return cuda_fail(e, "iblb_get_fields_static");
    return IBLB_OK;
}
