// Host-side solver object behind the C ABI of include/iblb.h.
#pragma once
#include <cuda_runtime.h>
#include <string>
#include <vector>

#include "iblb.h"
#include "iblb_device.cuh"

namespace iblb {

void set_error(const std::string &msg);
int cuda_fail(cudaError_t e, const char *what);

#define IBLB_CUDA(call)                                         \
    do {                                                        \
        cudaError_t e__ = (call);                               \
        if (e__ != cudaSuccess) return cuda_fail(e__, #call);   \
    } while (0)

// kernels (lb_kernels.cu / ib_kernels.cu)
__global__ void lb_fused_v2(const LbParams p);
__global__ void lb_fused_v4(const LbParams p);
__global__ void lb_fused_v1(const LbParams p, double *qpart);
__global__ void flux_reduce_v1(const double *qpart, int n, double *qacc, double *qhist,
                               const int *it_ptr, int it);
__global__ void lb_init(const LbParams p, const double *rho_in, const double *u_in, double u0_amp,
                        int row0, int nrows);
__global__ void lb_init_t(const LbParams p, const double *rho_in, const double *u_in, int row0,
                          int nrows);
__global__ void halo_push(const Lattice L, const double *buf, Remote left, Remote right);
__global__ void lb_moments_t(const LbParams p, int streamed, int row0, int nrows, double *rho_out,
                             double *u_out, double *force_out);
__global__ void lb_moments(const LbParams p, int streamed, int row0, int nrows, double *rho_out,
                           double *u_out, double *f_out, double *force_out);
__global__ void ib_coef(const int *k_of_j, float *coef, int mpc, int fine_points, int mucus);
__global__ void ib_kinematics_batch(const IbParams p);
__global__ void ib_interp(const IbParams p);
__global__ void ib_spread(const IbParams p);
__global__ void ib_tile_lists(const BandParams p);
__global__ void lb_ib_tile_step(const BandParams p);
__global__ void it_advance(int *it_ptr, int by);
unsigned long long checked_count_lb();     // IBLB_CHECKED builds: index violations per
unsigned long long checked_count_ib();     // translation unit (0 in product builds)
unsigned long long checked_count_fused();
constexpr size_t kStageDoubles = (size_t)16 << 20;  // 128 MB
int stage_buffer(iblb_solver *s);  // lazily allocates s->stage_dev
size_t tile_step_smem_bytes();
cudaError_t tile_step_prepare();  // once per device before the first launch
__global__ void slab_wait(const volatile int *flags, int *rel_left, int *rel_right, int need,
                          long long timeout_ns, int *err);
__global__ void slab_signal(int *left_flag, int *right_flag, int value, const int *err);
__global__ void slab_unpoison(int *flags, int value);

}  // namespace iblb

struct iblb_solver {
    iblb_params p;
    iblb::Lattice L;
    int device;
    bool vec2;                 // fast path usable (YDIM % 64 == 0)
    bool vec4;                 // lb_fused_v4 (256-bit accesses) where a launch's rows allow it
    double *buf[2];            // population buffers A/B
    int cur;                   // buffer the NEXT step pulls from
    unsigned it;               // completed iterations
    // IB
    int Ns;
    int mpc, fine_points;      // markers / fine points per cilium (main.cu:81-83: 96 / 9600)
    double shape_scale;        // main.cu:137-145: 111
    iblb::MarkerArrays M;
    float *coef;
    int *k_of_j_dev;
    std::vector<int> k_of_j;
    double *band;
    int band_rows, band_ld;
    int *it_dev;               // device copy of `it` read by the graph-captured kinematics
    long long kin_base;        // marker slots hold iterations [kin_base, kin_base + kKinBatch); -1: none
    // single-launch iteration: lb_ib_tile_step (interpolation and spreading inside the LB kernel)
    bool fuse_ok;              // geometry allows it (whole channel, YDIM % 64 == 0)
    bool ib_stale;             // F_s / force band not materialised for the last step
    bool pdl_chain;            // the last kernel enqueued was lb_ib_tile_step over the whole grid
    iblb::TileProgram *prog;   // per slot and band tile: fluid-independent plan (ib_tile_lists)
    int band_first;            // lb_ib_tile_step's block order (see fused_step.cu), chosen per grid
    int tile_w;                // columns per tile of lb_ib_tile_step: 8, or 6 where that fills the SMs better
    int tiles_x, band_tiles;   // column tiles, band row tiles
    long long lists_base;      // tile lists hold iterations [lists_base, lists_base + kKinBatch); -1: none
    float reach_x, reach_y;    // bounds on the marker excursion from the cilium base
    std::vector<float> env;    // per marker of a cilium: x lo, x hi, y lo, y hi over the beat cycle
    int tile_bound;            // upper bound on the markers that can ever touch one tile
    int *err_dev;
#ifdef IBLB_TILE_TIMING
    long long *dbg_dev;        // development: phase stamps of lb_ib_tile_step
#endif
    float4 *replay_dev;        // iblb_set_marker_replay (test hook) or null
    int replay_steps, replay_j;
    // flux
    double *qacc, *qpart;
    double *qhist;             // [kFluxHist][nq] running flux sums after each of the last iterations
    unsigned hist_from;        // first iteration whose row of qhist is valid (restart point)
    int cap_off;               // >= 0 while a graph batch is being captured: offset within the batch
    double *q_host;            // pinned: nq flux partial sums + the device error flag (one sync)
    double *stage_dev;         // kStageDoubles: row chunks of fields on their way to / from the host
    cudaStream_t stream_copy;  // PCIe copies of one half of it while a kernel works on the other
    cudaEvent_t ev_copy[2], ev_kern[2];
    int nq;
    bool flux_on;
    int xq_local;
    // execution
    cudaStream_t stream;
    cudaEvent_t ev0, ev1;
    bool timed;
    cudaGraphExec_t graph_exec;  // a batch of iterations + it_advance, built lazily
    unsigned long long graph_launches;  // kernel nodes in it
    unsigned long long last_launches;
    // slab coupling
    bool connected;
    iblb::Remote nb[2][2];     // [face 0=left,1=right][buffer parity]
    void *nb_mapped[2][2];
    int *flags;                // kSlabFlags step counters written by the neighbours (see slab_wait)
    int *nb_flag[2];           // left neighbour's flags[1], right neighbour's flags[0] (data ready)
    int *nb_rel[2];            // left neighbour's flags[3], right neighbour's flags[2] (release)
    cudaStream_t stream_b;     // coupled slabs: wait -> boundary columns -> signal run here,
    cudaEvent_t ev_b, ev_i, ev_ib;  // beside the interior kernel on `stream` (events order them)
    long long slab_timeout_ns; // bound on slab_wait's spin (IBLB_SLAB_TIMEOUT_S, default 900 s)
    void *nb_flag_mapped[2];
    bool self_left, self_right;
    int flag_base;             // step flags carry flag_base + it (monotonic across set_fields)
    iblb_solver *peer[2];      // same-process neighbours (iblb_connect_local)
    cudaEvent_t ev_boundary;   // recorded after this slab's boundary-column kernel
};
