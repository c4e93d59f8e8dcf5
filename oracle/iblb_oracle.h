/* TEST INFRASTRUCTURE — declarations of the oracle port (oracle/iblb_oracle.c). */
#ifndef IBLB_ORACLE_H
#define IBLB_ORACLE_H
#include "ref_api.h"

#ifdef __cplusplus
extern "C" {
#endif

/* same contract as the ref_* functions of ref_api.h */
void *oracle_create(const ref_params *p);
int oracle_step(void *h, int n_steps);
void oracle_get_fields(void *h, double *rho, double *u);
void oracle_get_force(void *h, double *force);
void oracle_get_f(void *h, double *f);
void oracle_get_markers(void *h, float *s, float *u_s, float *F_s, int *eps);
double oracle_get_Q(void *h);
int oracle_get_it(void *h);
void oracle_destroy(void *h);
double oracle_last_step_seconds(void *h);
const char *oracle_kind(void);
int oracle_marker_fine_index(void *h, int j);
/* replay which candidate won the reference's race on marker 32, per step and cilium */
void oracle_set_marker32_trace(void *h, const unsigned char *trace, int n_steps);
/* replay marker 32 of every cilium as recorded from the reference's CUDA build:
 * vals[(it*c_num+m)*4 + {0..3}] = s_x, s_y, u_s_x, u_s_y */
void oracle_set_marker32_replay(void *h, const float *vals, int n_steps);

/* derived run parameters of main.cu:291-321,336 */
typedef struct oracle_derived {
    unsigned XDIM, T, ITERATIONS, INTERVAL;
    int p_step;
    double SPEED, t_scale, x_scale, s_scale, TAU, TAU2, Ma;
} oracle_derived;
void oracle_derive(unsigned c_fraction, unsigned c_num, unsigned c_space, double Re, float T_num,
                   unsigned T_pow, float I_pow, unsigned P_num, oracle_derived *d);

#ifdef __cplusplus
}
#endif
#endif
