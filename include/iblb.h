/* libiblb — C ABI of the B200-native IB-LBM time-step path.
 *
 * This is the drop-in boundary for the time-step loop of ptheywood/CUDA_IBLB_11
 * (CUDA_IBLB_11/main.cu:817-934 and the eight kernels it launches: LatticeBoltzmann.cuh:4-10,
 * ImmersedBoundary.cuh:4-8, main.cu:77, main.cu:176).  The reference has no plugin/FFI layer; its
 * replaceable surface is main()'s argv (main.cu:284-296), the derived run parameters
 * (main.cu:298-321,336), the device state main() copies back (main.cu:944-982) and the Data/
 * files it writes (main.cu:589-631, 936-1005, 1030-1060).  Each entry point below names the
 * reference lines it replaces.  Plain pointers and sizes only; every function returns 0 on
 * success or an IBLB_E* code and never aborts (the reference prints and carries on,
 * main.cu:724-728); iblb_last_error() gives the message.
 *
 * Host arrays use the REFERENCE's layouts (main.cu:469-501): node index j = y*XDIM + x;
 * rho[size]; u/force plane-major [c*size + j]; populations AoS [9*j + i]; markers interleaved
 * (x,y) floats [2*k + c].  The device layout (SoA, y-fastest, ghost columns) is internal.
 */
#ifndef IBLB_H
#define IBLB_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define IBLB_OK 0
#define IBLB_EINVAL 1   /* bad argument / unsupported geometry */
#define IBLB_ECUDA 2    /* a CUDA runtime call failed (message has cudaGetErrorString) */
#define IBLB_EARGS 3    /* too few driver arguments (main.cu:284-289) */
#define IBLB_ESIZE 4    /* XDIM < 2*LENGTH (main.cu:303-308) */
#define IBLB_EIO 5      /* an output file could not be written */

/* flags */
#define IBLB_FLAG_HALF_FORCE_BODY 1u /* Guo half-force velocity also for the uniform body force
                                        when no cilia are present (extension; the reference's LB
                                        chain alone applies none, LatticeBoltzmann.cu:404-405) */
#define IBLB_FLAG_PICK_HIGHEST_K 2u  /* resolve the define_filament marker write race
                                        (main.cu:158-172) for the HIGHER fine-point index, i.e.
                                        sequential semantics.  Default: the lower index, which is
                                        what the reference's CUDA build does on a B200 (measured:
                                        95 of 96 markers per cilium; marker 32's two candidates
                                        sit in different warps and its winner varies per step) */
#define IBLB_FLAG_NO_GRAPH 4u        /* launch kernels directly instead of CUDA-graph batches */
#define IBLB_FLAG_NO_FUSE 8u         /* keep the immersed-boundary pre-pass as separate kernels
                                        (kinematics, interpolate, spread) instead of the single
                                        fused IB+LB launch per iteration; same results bit for bit */
#define IBLB_FLAG_NO_PDL 16u         /* no programmatic dependent launch between consecutive fused
                                        iterations (full kernel-to-kernel serialisation) */
#define IBLB_FLAG_MUCUS 32u          /* the reference's other beat-shape coefficient set, "WITH
                                        MUCUS PRESENT" (main.cu:36-54, commented out there); default:
                                        the active "WITHOUT MUCUS" set (main.cu:56-74) */

#define IBLB_FLAG_DEV_KNOBS 64u       /* development only: let the environment variables IBLB_V4,
                                        IBLB_TILE_W and IBLB_BAND_FIRST override the kernel variant,
                                        tile width and block order that iblb_create picks (A/B
                                        measurements, tools/).  Without this flag they are ignored, so
                                        that no stray variable changes what a production run launches;
                                        results are bit-identical either way */

typedef struct iblb_params {
    /* --- the ten driver arguments, main.cu:291-296 --- */
    unsigned c_fraction, c_num, c_space;
    double Re;
    float T_num;
    unsigned T_pow;
    float I_pow;
    unsigned P_num;
    int ShARC, BigData;
    /* --- derived, main.cu:298-321,336 (filled by iblb_params_derive) --- */
    unsigned XDIM, YDIM, T, ITERATIONS, INTERVAL, LENGTH;
    int p_step;
    double SPEED, t_scale, x_scale, s_scale, TAU, TAU2, Ma;
    /* --- extensions (all 0 in a reference run) --- */
    double gx, gy;     /* uniform body force (config C2) */
    double u0_amp;     /* initial u_x(y) = u0_amp*sin(2*pi*(y+0.5)/YDIM) */
    unsigned flags;    /* IBLB_FLAG_* */
    int device;        /* CUDA device ordinal; -1 = current device */
    /* x-slab of a wider periodic channel (multi-GPU): this solver owns global columns
     * [slab_x0, slab_x0+slab_w); slab_w == 0 means the whole channel (single GPU). */
    unsigned slab_x0, slab_w;
    /* cilium geometry the reference hard-wires; 0 = the reference's value.  LENGTH above is the
     * number of markers per cilium (main.cu:83,180) as well as the length scale of TAU and of the
     * output units (main.cu:279,312,320): 1..128. */
    unsigned fine_points;  /* fine points per cilium that define_filament evaluates, main.cu:81
                              (9600); marker j is the fine point k with |shape_scale*k/fine_points
                              - j| < 0.01 (main.cu:158-172) — every marker must have one */
    double shape_scale;    /* filament length scale in lattice units, main.cu:137-145 (111) */
    unsigned flux_back;    /* the flux plane is x = XDIM - flux_back, ImmersedBoundary.cu:259 (5) */
} iblb_params;

typedef struct iblb_solver iblb_solver;

/* ---- driver surface ---------------------------------------------------------------------- */

/* Defaults of main.cu:267-282 (c_fraction 1, c_num 6, c_space 48, Re 1, T_num 1, T_pow 1 ...). */
void iblb_params_default(iblb_params *p);
/* argv -> the ten arguments, exactly as main.cu:284-296 parses them (one stream, same types).
 * Also accepts the scripts' legacy 8-value line `c_fraction c_num Re T ITERATIONS P_num ShARC
 * BigData` (app.sh:7, cilia6.sh:7-17) and maps it to the 10-value form with c_space = 48.
 * Returns IBLB_EARGS with the reference's message in iblb_last_error() when too few are given.
 * Calls iblb_params_derive on success. */
int iblb_params_from_args(int argc, const char *const *argv, iblb_params *p);
/* main.cu:298-321,336: XDIM, T, ITERATIONS, INTERVAL, scales, TAU, TAU2, Ma, p_step.  YDIM and
 * LENGTH keep their values when non-zero (reference: 192 and 96, main.cu:271,279).
 * Returns IBLB_ESIZE when XDIM < 2*LENGTH with cilia present (main.cu:303-308). */
int iblb_params_derive(iblb_params *p);

/* ---- lifecycle (replaces main.cu:389-587 allocation, 634-754 initial state) ------------- */
int iblb_create(const iblb_params *p, iblb_solver **out);
int iblb_destroy(iblb_solver *s);
/* Advance n_steps iterations of the loop main.cu:817-934.  Asynchronous on the solver's stream;
 * every getter synchronises. */
int iblb_step(iblb_solver *s, unsigned n_steps);
int iblb_sync(iblb_solver *s);
unsigned iblb_iteration(const iblb_solver *s);

/* ---- state access (replaces the D2H copies of main.cu:928, 944-982) ---------------------- */
/* rho[size], u[2*size] of the last completed iteration, reference layout; either may be NULL.
 * For a slab solver, size = slab_w*YDIM and x is slab-local. */
int iblb_get_fields(iblb_solver *s, double *rho, double *u);
int iblb_get_force(iblb_solver *s, double *force /*2*size*/);
/* post-streaming populations f[9*size] of the last completed iteration (what d_f holds after
 * `streaming`, LatticeBoltzmann.cu:173-373) */
int iblb_get_populations(iblb_solver *s, double *f);
int iblb_get_markers(iblb_solver *s, float *pos, float *vel, float *force /*2*Ns each*/,
                     int *eps /*Ns*/);
/* iblb_get_fields for callers that must not allocate device memory.  No counterpart in the
 * reference (its main() allocates everything up front, main.cu:389-587, and copies straight out of
 * those arrays); it exists because iblb_get_fields reserves its 128 MB device staging buffer with
 * cudaMalloc on first use.  Two kinds of host program cannot live with that: one that gave all of
 * the GPU's memory to its own arena at start-up, where a late cudaMalloc fails in the middle of a
 * run; and one in which some other host thread keeps a stream capture open in the global capture
 * mode, where the runtime refuses cudaMalloc from every thread and the refusal spoils that capture.
 * This variant stages through a fixed array inside the library image (6 MiB on every GPU that loads
 * libiblb.so, held until the process exits) and makes no cudaMalloc / cudaFree or other allocating
 * call.  Same arguments, same values bit for bit, same error codes as iblb_get_fields, except:
 *  - grids of more than IBLB_STATIC_FIELDS_MAX_NODES nodes (slab_w*YDIM for a slab) are refused
 *    with IBLB_EINVAL and a message that names both sizes; C1, C3 and C5 (288x192, 576x192) fit;
 *  - the first call on each device looks up the array's device address (and may make the runtime
 *    load the library's device code there): make that first call before any stream capture begins;
 *  - every call on a GPU, whichever solver it is for, goes through the same array.  A call returns
 *    only when its copies have landed, so calls made one after another are safe; calls running at
 *    the same time on one GPU from several host threads overwrite each other's data and must be
 *    serialised by the caller.  Nothing in the array needs clearing between calls: each call
 *    writes every value it reads back. */
#define IBLB_STATIC_FIELDS_MAX_NODES 262144 /* 2^18 nodes, e.g. 1365 x 192 */
int iblb_get_fields_static(iblb_solver *s, double *rho, double *u);
/* cumulative flux Q of main.cu:928 / ImmersedBoundary.cu:259-264 */
int iblb_get_flux(iblb_solver *s, double *Q);
/* Q as it stood after each of the iterations first_it .. first_it+n-1 (0-based): the values the
 * reference's per-step copy of d_Q delivers (main.cu:928), kept on the device for the last 4096
 * completed iterations, so that a caller who wants Q every step need not synchronise every step. */
int iblb_get_flux_history(iblb_solver *s, unsigned first_it, unsigned n, double *Q /*n*/);
/* Replace the state by the equilibrium of (rho,u) as main.cu:722-754 does for rho=1,u=0;
 * resets the iteration counter and Q.  Host arrays, reference layout. */
int iblb_set_fields(iblb_solver *s, const double *rho, const double *u);
/* which fine point k in [0,9600) feeds marker j in [0,96) (main.cu:158-172) */
int iblb_marker_fine_index(const iblb_solver *s, unsigned j);
/* Test hook: marker `marker_j` of cilium m takes (s_x, s_y, u_s_x, u_s_y) = vals[(it*c_num + m)*4
 * + 0..3] at iteration it < n_steps instead of the computed values (NULL / 0 switches it off).
 * The reference's define_filament has one timing-dependent write race (marker 32, main.cu:158-172);
 * replaying what a recorded run of the reference's CUDA build produced there lets the product be
 * compared with that run at full tolerance.  Call before stepping. */
int iblb_set_marker_replay(iblb_solver *s, unsigned marker_j, const float *vals,
                           unsigned n_steps);

/* ---- measurement helpers ------------------------------------------------------------------ */
/* Device time (CUDA events on the solver's stream) of the last iblb_step call, seconds. */
int iblb_last_step_seconds(iblb_solver *s, double *seconds);
/* Kernel launches issued by the last iblb_step call (graph nodes count one each). */
int iblb_last_step_launches(iblb_solver *s, unsigned long long *n);
/* Time the phases of one iteration separately (kinematics, interpolate, spread, fused LB),
 * averaged over `reps` iterations that DO advance the state; seconds per phase. */
int iblb_profile_phases(iblb_solver *s, unsigned reps, double *t_kin, double *t_interp,
                        double *t_spread, double *t_lb);

/* ---- x-slab coupling (multi-GPU; one solver per process/GPU) ----------------------------- */
#define IBLB_IPC_HANDLE_BYTES 64
typedef struct iblb_halo_handle {
    unsigned char mem[2][IBLB_IPC_HANDLE_BYTES]; /* cudaIpcMemHandle_t of population buffers A,B */
    unsigned char flag[IBLB_IPC_HANDLE_BYTES];   /* cudaIpcMemHandle_t of the step flags */
    unsigned long long plane;  /* doubles per population plane */
    unsigned ld, ghost, width; /* column pitch (doubles), ghost columns, owned columns */
    int device, pid;           /* exporter's CUDA ordinal (local to ITS process) and process id */
    unsigned char uuid[16];    /* exporter's GPU (cudaDeviceProp::uuid): ordinals mean nothing across
                                  processes whose CUDA_VISIBLE_DEVICES differ, this does */
} iblb_halo_handle;
/* Export this slab's buffers; import the left/right neighbours' (periodic ring).  After
 * iblb_connect the fused LB kernel stores its outgoing boundary columns straight into the
 * neighbours' ghost columns over NVLink and steps are ordered by device-side flags.
 * A slab waits on the device for its neighbours for at most IBLB_SLAB_TIMEOUT_S seconds (environment,
 * read by iblb_create; default 900): whatever a rank does on the host between two iblb_step calls
 * while the others step on (checkpoints, snapshot files) must finish within that.  A slab that gives
 * up marks its neighbours' flags, they mark theirs: within a few steps every rank's next getter or
 * iblb_sync returns IBLB_ECUDA ("slab_wait timed out").  The run's state is then invalid on all ranks;
 * iblb_set_fields on every rank (a collective restart) clears the condition when all ranks had been
 * asked for the same number of steps; otherwise destroy the solvers and build the ring again. */
int iblb_export_halo(iblb_solver *s, iblb_halo_handle *h);
int iblb_connect(iblb_solver *s, const iblb_halo_handle *left, const iblb_halo_handle *right);
/* Same-process variant (several solvers driven by one host thread, on one or several GPUs):
 * neighbours are given as solver handles, ordering uses CUDA events instead of flags.  The caller
 * must advance all slabs in lockstep: iblb_step(slab, 1) for every slab, then the next step. */
int iblb_connect_local(iblb_solver *s, iblb_solver *left, iblb_solver *right);

/* Checkpoint / restart (new; the reference restarts its 100 000-step runs from scratch).  The file
 * holds the complete time-stepping state of this solver (both population buffers, iteration,
 * flux sums); marker state is a function of the iteration and is not stored.  Load into a solver
 * created with the same parameters (flags may differ), before iblb_connect for x-slabs; stepping
 * on and every getter then continue bit for bit as if the run had never stopped. */
int iblb_save_checkpoint(iblb_solver *s, const char *path);
int iblb_load_checkpoint(iblb_solver *s, const char *path);

/* ---- driver-level run (replaces main() as a whole: SimLog, flux, fluid/cilia files) ------ */
/* Runs ITERATIONS steps writing the reference's Data/ tree under `root` (NULL: "Data/").
 * main.cu:589-631, 758-790, 936-1005, 1030-1060. */
int iblb_run(const iblb_params *p, const char *root);
/* The same on n_gpus GPUs of this process: the channel is cut into n_gpus x-slabs (whole cilium
 * spacings each) on devices p->device, p->device+1, ... (wrapping when the machine has fewer),
 * coupled by iblb_connect_local and advanced in lockstep by the calling thread; fields are gathered
 * into the same files, the flux comes from the slab that owns the flux plane, a checkpoint is one
 * file per slab (<path>.<slab>).  iblb_run takes n_gpus from the environment (IBLB_GPUS, default 1);
 * the reference is single-GPU (device selection: main.cu:395-396). */
int iblb_run_gpus(const iblb_params *p, const char *root, int n_gpus);
/* The partition iblb_run_gpus uses: columns [x0, x0+w) of slab `rank` of `n_slabs`, widths multiples
 * of `align` (c_space with cilia, so that no cilium base sits on a seam), remainder to the low ranks. */
int iblb_slab_range(int rank, int n_slabs, unsigned XDIM, unsigned align, unsigned *x0, unsigned *w);
/* CUDA devices visible to this process (0 when the runtime cannot be initialised). */
int iblb_device_count(void);

/* Development: device-side index violations counted by a library built with IBLB_CHECKED=1
 * (range checks on every population / band / marker / shared-table index); -1 in product builds. */
long long iblb_debug_violations(void);

const char *iblb_last_error(void);
const char *iblb_version(void);

#ifdef __cplusplus
}
#endif
#endif
