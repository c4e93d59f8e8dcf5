/* TEST INFRASTRUCTURE — not part of the product path.
 *
 * One small C API shared by the three checkers, so that tests drive them interchangeably:
 *   oracle/_ref/libiblb_ref_cpu.so   reference kernels compiled unmodified as host C++ (shim)
 *   oracle/_ref/libiblb_ref_cuda.so  reference kernels compiled unmodified by nvcc for sm_100a
 *   oracle/libiblb_oracle.so         plain-C restatement of the same algorithm (oracle port)
 *
 * All of them keep the REFERENCE's memory layout (main.cu:469-501): f/f0/f1/F AoS [9*j+i],
 * u/force plane-major [c*size+j], rho[j], j = y*XDIM+x; markers interleaved (x,y) floats.
 * One "step" is one iteration of the reference loop main.cu:817-934 in its launch order:
 * define_filament, boundary_check, equilibrium, collision, streaming, macro, interpolate, spread.
 * With c_num == 0 only the four LB kernels run (the reference kernels, driven as a pure-LB
 * channel; `force` planes hold the uniform body force; no half-force correction because
 * `spread` — which owns it, ImmersedBoundary.cu:249-255 — is not launched).
 */
#ifndef IBLB_REF_API_H
#define IBLB_REF_API_H

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ref_params {
    int XDIM, YDIM;   /* YDIM must be 192 when c_num > 0 (ImmersedBoundary.cu:146,261) */
    int c_num;        /* number of cilia; 0 = pure LB chain */
    int c_space;      /* lattice units between cilia bases (XDIM == c_num*c_space when c_num>0) */
    int T;            /* beat period in steps (main.cu:299) */
    int p_step;       /* phase lag between adjacent cilia in steps (main.cu:336) */
    double TAU, TAU2; /* relaxation times (main.cu:320-321) */
    double gx, gy;    /* uniform body force put in the force planes (0 in the reference) */
    double u0_amp;    /* initial u_x(y) = u0_amp*sin(2*pi*(y+0.5)/YDIM); 0 in the reference */
    int threads;      /* CPU builds: OpenMP threads (0 = all) */
    int flags;        /* must be 0 for the two reference builds; oracle port: REF_FLAG_* */
    /* geometry the reference hard-wires (0 = its value); honoured by the oracle port only, the two
     * reference builds require the defaults */
    int length;         /* markers per cilium, main.cu:83,180,279 (96) */
    int fine_points;    /* fine points per cilium, main.cu:81 (9600) */
    double shape_scale; /* filament length scale, main.cu:137-145 (111) */
    int flux_back;      /* flux plane x = XDIM - flux_back, ImmersedBoundary.cu:259 (5) */
} ref_params;

/* oracle-port extensions (documented deviations from the reference driver):
 * HALF_FORCE_BODY: run the spread stage every step even without cilia, with `force` starting from
 *   the uniform body force instead of 0 — i.e. Guo's half-force velocity for a forced channel;
 * PICK_LOWEST_K: resolve the define_filament write race (main.cu:158-172) in favour
 *   of the lower fine-point index instead of the higher one. */
#define REF_FLAG_HALF_FORCE_BODY 1
#define REF_FLAG_PICK_LOWEST_K 2
/* MUCUS: the reference's other beat-shape coefficient set ("WITH MUCUS PRESENT", main.cu:36-54,
 * commented out in the source); pinned against libiblb_ref_cpu_mucus.so, which is the reference
 * compiled with that block swapped in (oracle/Makefile, ref_cpu_mucus). */
#define REF_FLAG_MUCUS 4

void *ref_create(const ref_params *p);
int ref_step(void *h, int n_steps);
void ref_get_fields(void *h, double *rho /*size*/, double *u /*2*size*/);
void ref_get_force(void *h, double *force /*2*size*/);
void ref_get_f(void *h, double *f /*9*size, post-stream populations*/);
void ref_get_markers(void *h, float *s, float *u_s, float *F_s /*2*Ns each*/, int *eps /*Ns*/);
double ref_get_Q(void *h);
int ref_get_it(void *h);
void ref_destroy(void *h);
/* wall-clock seconds spent inside the last ref_step call (CUDA build: device time by events) */
double ref_last_step_seconds(void *h);
const char *ref_kind(void);

#ifdef __cplusplus
}
#endif
#endif
