/* TEST INFRASTRUCTURE — the oracle.  Not part of the product path: only tests/, the smoke check in
 * __graft_entry__.py and bench.py's cpu_baseline / --impl reference legs may load this.
 *
 * Plain-C restatement of the time-step path of ptheywood/CUDA_IBLB_11, one function per reference
 * kernel, each citing the lines it follows.  It keeps the reference's memory layout, evaluation
 * order and rounding points (the quirk ledger of SURVEY.md §8: Q1-Q9, Q11, Q14-Q17, Q20), so that
 * on the host — where neither build can contract a*b+c — it is BIT-IDENTICAL to the reference's
 * own kernels compiled as host C++ (oracle/_ref/libiblb_ref_cpu.so); tests/test_oracle_pin.py
 * checks exactly that, and checks both against golden vectors produced by the reference's CUDA
 * build on a B200 (tests/golden/).  Parity status: PINNED against the reference's own code (the
 * reference ships no tests or golden vectors of its own for the current source).
 *
 * Two deliberate restructurings, both result-preserving:
 *  - define_filament evaluates only the fine points that can be picked as markers (the other
 *    ~9430 of 9600 per cilium never reach any output, main.cu:158-172);
 *  - spread loops markers-outer over each marker's delta support instead of nodes-outer over all
 *    markers (ImmersedBoundary.cu:178-231): every node still receives its non-zero terms in
 *    ascending marker order; the skipped terms are exact zeros.
 * Extensions beyond the reference driver (off unless asked for): YDIM other than 192 (flux divisor
 * and plane stride follow YDIM), REF_FLAG_* in ref_api.h.
 */
#define _GNU_SOURCE
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "iblb_oracle.h"

/* LatticeBoltzmann.cu:11 — device sound speed (differs from the host's 0.577, Q1) */
static const double C_S = 0.57735;
/* LatticeBoltzmann.cu:15-20, ImmersedBoundary.cu:14-19 */
static const double c_l[9 * 2] = {0., 0., 1., 0., 0., 1., -1., 0., 0., -1.,
                                  1., 1., -1., 1., -1., -1., 1., -1.};
/* LatticeBoltzmann.cu:22-27 */
static const double t_w[9] = {4. / 9,  1. / 9,  1. / 9,  1. / 9, 1. / 9,
                              1. / 36, 1. / 36, 1. / 36, 1. / 36};
/* main.cu:29, ImmersedBoundary.cu:10 */
#define PI_REF 3.14159

/* main.cu:56-74 (coefficient set "without mucus") */
static const double A_mn[7 * 2 * 3] = {
    -0.654, 0.393,  -0.097, 0.079,  0.119,  0.119,  0.009,  1.895,  -0.018, 0.158,  0.010,
    0.003,  0.013,  0.040,  0.787,  -1.516, 0.032,  -0.302, -0.252, -0.015, 0.035,  -0.552,
    -0.126, -0.341, 0.035,  0.006,  -0.029, -0.068, 0.202,  0.716,  -0.118, 0.142,  0.110,
    -0.013, -0.043, 0.096,  0.263,  0.186,  -0.067, -0.032, -0.002, 0.015};
static const double B_mn[7 * 2 * 3] = {
    0.0,    0.284, 0.006,  -0.059, 0.018,  0.053,  0.009,  0.0,    0.192,  -0.050, 0.012,
    -0.007, -0.014, -0.017, 0.0,   1.045,  0.317,  0.226,  0.004,  -0.082, -0.040, 0.0,
    -0.499, 0.423, 0.138,  0.125,  0.075,  0.067,  0.0,    -1.017, -0.276, -0.196, -0.037,
    0.025,  0.023, 0.0,    0.339,  -0.327, -0.114, -0.105, -0.057, -0.055};

/* main.cu:36-54 (coefficient set "WITH MUCUS PRESENT", commented out in the reference) */
static const double A_mn_mucus[7 * 2 * 3] = {
    -0.449, 0.130,  -0.169, 0.063,  -0.050, -0.040, -0.068, 2.076,  -0.003, 0.054,  0.007,
    0.026,  0.022,  0.010,  -0.072, -1.502, 0.260,  -0.123, 0.011,  -0.009, 0.196,  -1.074,
    -0.230, -0.305, -0.180, -0.069, 0.001,  -0.080, 0.658,  0.793,  -0.251, 0.049,  0.009,
    0.023,  -0.111, 0.381,  0.331,  0.193,  0.082,  0.029,  0.002,  0.048};
static const double B_mn_mucus[7 * 2 * 3] = {
    0.0,   -0.030, -0.093, 0.037,  0.062,  0.016,  -0.065, 0.0,   0.080,  -0.044, -0.017,
    0.052, 0.007,  0.051,  0.0,    1.285,  -0.036, -0.244, -0.093, -0.137, 0.095,  0.0,
    -0.298, 0.513, 0.004,  -0.222, 0.035,  -0.128, 0.0,    -1.034, 0.050,  0.143,  0.043,
    0.098, -0.054, 0.0,    0.210,  -0.367, 0.009,  0.120,  -0.024, 0.102};

/* the reference's hard-wired geometry (ref_params fields of the same names override them) */
#define F_LENGTH_REF 9600 /* main.cu:81 */
#define LENGTH_REF 96     /* main.cu:83 */
#define SCALE_REF 111.    /* main.cu:137-145 */
#define FLUX_BACK_REF 5   /* ImmersedBoundary.cu:259 */
/* run-time values */
#define F_LENGTH (o->f_length)
#define LENGTH (o->length)
#define SCALE (o->scale)

typedef struct Oracle {
    ref_params p;
    int size, Ns, it;
    int length, f_length, flux_back; /* markers / fine points per cilium, flux plane offset */
    double scale;                    /* 111 */
    const double *A, *B;             /* beat-shape coefficient set */
    double *u, *rho, *f0, *f, *f1, *force, *F;
    float *s, *u_s, *F_s, *b_points;
    int *epsilon;
    /* candidate fine points per cilium (same for every cilium): index k, marker j it feeds */
    int ncand;
    int *cand_k, *cand_j;
    float *fine;  /* [c_num][ncand][5]  the reference's d_boundary rows (main.cu:137-151) */
    float *lasts; /* [c_num][ncand][2]  main.cu:154-155 */
    double Q;
    double last_seconds;
    int threads;
    /* optional replay of the reference's timing-dependent race on marker 32 (its two candidate
     * fine points sit in different warps): trace[it*c_num+m] != 0 -> higher k wins */
    unsigned char *trace;
    int trace_steps;
    /* optional replay of marker 32 exactly as the reference produced it (covers torn writes of the
     * racing stores as well): replay[(it*c_num+m)*4 + {0,1,2,3}] = s_x, s_y, u_s_x, u_s_y after
     * boundary_check's translation (main.cu:193-209) */
    float *replay;
    int replay_steps;
} Oracle;

/* ------------------------------------------------------------------ LB */

/* LatticeBoltzmann.cu:30-62 */
static void o_equilibrium(Oracle *o) {
    const int size = o->size;
    const double TAU = o->p.TAU;
    const double *u = o->u, *rho = o->rho, *force = o->force;
#pragma omp parallel for num_threads(o->threads)
    for (int j = 0; j < size; j++) {
        double vec[2];
        for (int i = 0; i < 9; i++) {
            const double ux = u[0 * size + j], uy = u[1 * size + j];
            const double cx = c_l[2 * i + 0], cy = c_l[2 * i + 1];
            /* :47-50 */
            o->f0[9 * j + i] =
                rho[j] * t_w[i] *
                (1 + (ux * cx + uy * cy) / (C_S * C_S) +
                 (ux * cx + uy * cy) * (ux * cx + uy * cy) / (2 * C_S * C_S * C_S * C_S) -
                 (ux * ux + uy * uy) / (2 * C_S * C_S));
            /* :53-54 */
            vec[0] = (cx - ux) / (C_S * C_S) + (cx * ux + cy * uy) / (C_S * C_S * C_S * C_S) * cx;
            vec[1] = (cy - uy) / (C_S * C_S) + (cx * ux + cy * uy) / (C_S * C_S * C_S * C_S) * cy;
            /* :56 */
            o->F[9 * j + i] = (1. - 1. / (2. * TAU)) * t_w[i] *
                              (vec[0] * force[size * 0 + j] + vec[1] * force[size * 1 + j]);
        }
    }
}

/* LatticeBoltzmann.cu:64-136 — TRT; the rest population gets no force term (Q2) */
static void o_collision(Oracle *o) {
    const int size = o->size;
    const double omega_plus = 1 / o->p.TAU, omega_minus = 1 / o->p.TAU2;
    static const int pairs[4][2] = {{1, 3}, {2, 4}, {5, 7}, {6, 8}};
#pragma omp parallel for num_threads(o->threads)
    for (int j = 0; j < size; j++) {
        const double *f = o->f + 9 * (size_t)j, *f0 = o->f0 + 9 * (size_t)j,
                     *F = o->F + 9 * (size_t)j;
        double *f1 = o->f1 + 9 * (size_t)j;
        f1[0] = f[0] - omega_plus * (f[0] - f0[0]); /* :86 */
        for (int q = 0; q < 4; q++) {               /* :88-134 */
            const int a = pairs[q][0], b = pairs[q][1];
            double f_plus = (f[a] + f[b]) / 2.;
            double f_minus = (f[a] - f[b]) / 2.;
            double f0_plus = (f0[a] + f0[b]) / 2.;
            double f0_minus = (f0[a] - f0[b]) / 2.;
            f1[a] = f[a] - omega_plus * (f_plus - f0_plus) - omega_minus * (f_minus - f0_minus) +
                    F[a];
            f_minus *= -1.;
            f0_minus *= -1.;
            f1[b] = f[b] - omega_plus * (f_plus - f0_plus) - omega_minus * (f_minus - f0_minus) +
                    F[b];
        }
    }
}

/* LatticeBoltzmann.cu:173-373 — push streaming.  Rules (wall tests before periodic ones,
 * :243-252): y==0 reflects 4,7,8 in place to 2,5,6 (:231,287,309,328-340); y==YDIM-1 reflects
 * 2,5,6 in place to 4,8,7 without tangential shift (:225,243,265,341-353); x periodic (:354-359). */
static void o_streaming(Oracle *o) {
    const int XDIM = o->p.XDIM, YDIM = o->p.YDIM, size = o->size;
    static const int back_k[9] = {0, 3, 4, 1, 2, 7, 8, 5, 6};
    static const int slip_k[9] = {0, 1, 4, 3, 2, 8, 7, 6, 5};
#pragma omp parallel for num_threads(o->threads)
    for (int j = 0; j < size; j++) {
        const int x = j % XDIM, y = j / XDIM;
        const int up = (y == YDIM - 1), down = (y == 0), left = (x == 0), right = (x == XDIM - 1);
        for (int i = 0; i < 9; i++) {
            int back = 0, thru = 0, slip = 0;
            switch (i) {
                case 1: if (right) thru = 1; break;
                case 2: if (up) slip = 1; break;
                case 3: if (left) thru = 1; break;
                case 4: if (down) back = 1; break;
                case 5: if (up) slip = 1; else if (right) thru = 1; break;
                case 6: if (up) slip = 1; else if (left) thru = 1; break;
                case 7: if (down) back = 1; else if (left) thru = 1; break;
                case 8: if (down) back = 1; else if (right) thru = 1; break;
                default: break;
            }
            long jstream;
            int k;
            if (back) {
                jstream = j;
                k = back_k[i];
            } else if (slip) {
                jstream = j;
                k = slip_k[i];
            } else if (thru) {
                jstream = (long)(j - (XDIM - 1) * c_l[i * 2 + 0] + XDIM * c_l[i * 2 + 1]); /* :356 */
                k = i;
            } else {
                jstream = (long)(j + c_l[i * 2 + 0] + XDIM * c_l[i * 2 + 1]); /* :362 */
                k = i;
            }
            o->f[9 * jstream + k] = o->f1[9 * (size_t)j + i]; /* :367 */
        }
    }
}

/* LatticeBoltzmann.cu:375-411 — sequential i=0..8 sums, no force correction (Q20) */
static void o_macro(Oracle *o) {
    const int size = o->size;
#pragma omp parallel for num_threads(o->threads)
    for (int j = 0; j < size; j++) {
        double r = 0, m0 = 0, m1 = 0;
        for (int i = 0; i < 9; i++) {
            r += o->f[9 * (size_t)j + i];
            m0 += c_l[i * 2 + 0] * o->f[9 * (size_t)j + i];
            m1 += c_l[i * 2 + 1] * o->f[9 * (size_t)j + i];
        }
        o->rho[j] = r;
        o->u[0 * size + j] = m0 / r;
        o->u[1 * size + j] = m1 / r;
    }
}

/* ------------------------------------------------------------------ IB */

/* one factor of ImmersedBoundary.cu:21-81; `r` is the float |distance| (:25-26) */
static float delta_1d(float r) {
    double a = 0., b = 0., d = 0.;
    int c = 0;
    if (r <= 1.5) {
        if (r <= 0.5) { /* :33-40 */
            a = 0.33333;
            b = 1.;
            c = 1;
            d = r;
        } else { /* :41-47; `1 - dx` is a float subtraction */
            a = 0.16667;
            b = 5. - 3. * r;
            c = -1;
            d = 1 - r;
        }
    }
    return (float)(a * (b + c * sqrt(-3. * d * d + 1))); /* :50,76 */
}

/* ImmersedBoundary.cu:21-81 */
static float o_delta(float xs, float ys, int x, int y) {
    float dx = fabsf(x - xs);
    float dy = fabsf(y - ys);
    float deltax = delta_1d(dx);
    float deltay = delta_1d(dy);
    return deltax * deltay; /* :78 float product */
}

/* ImmersedBoundary.cu:94-133 — 3x3 stencil round nearbyint(s), unwrapped node index (Q4),
 * F_s re-rounded to float after each of the 9 terms (Q6).  Nodes whose unwrapped index falls
 * outside the arrays (the reference would read out of bounds) are skipped. */
static void o_interpolate(Oracle *o) {
    const int size = o->size, XDIM = o->p.XDIM;
    for (int k = 0; k < o->Ns; k++) {
        o->F_s[2 * k + 0] = 0.;
        o->F_s[2 * k + 1] = 0.;
        double xs = o->s[k * 2 + 0], ys = o->s[k * 2 + 1];
        int x0 = (int)nearbyint(xs), y0 = (int)nearbyint(ys);
        for (int i = 0; i < 9; i++) {
            int x = (int)nearbyint(x0 + c_l[i * 2 + 0]);
            int y = (int)nearbyint(y0 + c_l[i * 2 + 1]);
            long j = (long)y * XDIM + x;
            if (j < 0 || j >= size) continue;
            double del = o_delta((float)xs, (float)ys, x, y);
            o->F_s[2 * k + 0] += 2. * (1. * 1. * del) * o->rho[j] *
                                 (o->u_s[2 * k + 0] - o->u[0 * size + j]); /* :126 */
            o->F_s[2 * k + 1] += 2. * (1. * 1. * del) * o->rho[j] *
                                 (o->u_s[2 * k + 1] - o->u[1 * size + j]); /* :127 */
        }
    }
}

/* ImmersedBoundary.cu:138-267.  force = sum over markers in ascending k of
 * (double)(F_s*delta [float product]) * epsilon (:196-197, Q7), no periodic image (Q5);
 * u = (sum c f + force/2)/rho from the post-stream f (:249-255); flux through x == XDIM-5
 * (:259-264).  The reference's divisor 192 is YDIM (Q9). */
static void o_spread(Oracle *o, double gx, double gy) {
    const int size = o->size, XDIM = o->p.XDIM, YDIM = o->p.YDIM;
    for (int j = 0; j < size; j++) {
        o->force[0 * size + j] = gx; /* :166-167 (0 in the reference) */
        o->force[1 * size + j] = gy;
    }
    for (int k = 0; k < o->Ns; k++) {
        const float xs = o->s[2 * k + 0], ys = o->s[2 * k + 1];
        const int xc = (int)floorf(xs), yc = (int)floorf(ys);
        for (int y = yc - 2; y <= yc + 3; y++) {
            if (y < 0 || y >= YDIM) continue;
            for (int x = xc - 2; x <= xc + 3; x++) {
                if (x < 0 || x >= XDIM) continue;
                float del = o_delta(xs, ys, x, y);
                const int j = y * XDIM + x;
                o->force[0 * size + j] += o->F_s[2 * k + 0] * del * 1. * o->epsilon[k];
                o->force[1 * size + j] += o->F_s[2 * k + 1] * del * 1. * o->epsilon[k];
            }
        }
    }
#pragma omp parallel for num_threads(o->threads)
    for (int j = 0; j < size; j++) {
        const double *f = o->f + 9 * (size_t)j;
        /* :249-255, same operand order (the u_y line's first term repeats f_1 times zero, Q8) */
        o->u[0 * size + j] =
            (c_l[0 * 2 + 0] * f[0] + c_l[1 * 2 + 0] * f[1] + c_l[2 * 2 + 0] * f[2] +
             c_l[3 * 2 + 0] * f[3] + c_l[4 * 2 + 0] * f[4] + c_l[5 * 2 + 0] * f[5] +
             c_l[6 * 2 + 0] * f[6] + c_l[7 * 2 + 0] * f[7] + c_l[8 * 2 + 0] * f[8] +
             0.5 * o->force[0 * size + j]) /
            o->rho[j];
        o->u[1 * size + j] =
            (c_l[1 * 2 + 1] * f[1] + c_l[1 * 2 + 1] * f[1] + c_l[2 * 2 + 1] * f[2] +
             c_l[3 * 2 + 1] * f[3] + c_l[4 * 2 + 1] * f[4] + c_l[5 * 2 + 1] * f[5] +
             c_l[6 * 2 + 1] * f[6] + c_l[7 * 2 + 1] * f[7] + c_l[8 * 2 + 1] * f[8] +
             0.5 * o->force[1 * size + j]) /
            o->rho[j];
    }
    /* the reference's order of the YDIM CAS-adds is arbitrary (Q10); ascending y here */
    const int xq = XDIM - o->flux_back;
    for (int y = 0; y < YDIM; y++) o->Q += o->u[0 * size + y * XDIM + xq] / (double)YDIM;
}

/* ------------------------------------------------------------------ cilia */

/* marker pick of main.cu:158-172 evaluated for every k: which fine points can write marker j */
static void build_candidates(Oracle *o) {
    int cap = 4 * LENGTH, n = 0;
    o->cand_k = (int *)malloc(sizeof(int) * cap);
    o->cand_j = (int *)malloc(sizeof(int) * cap);
    for (int k = 0; k < F_LENGTH; k++) {
        float arcl = 1. * k / F_LENGTH; /* :100 */
        float s2 = SCALE * arcl;        /* :139 */
        for (int j = 0; j < LENGTH; j++) {
            float b_length = j % LENGTH;
            if (fabsf(s2 - b_length) < 0.01) { /* :162 */
                if (n == cap) {
                    cap *= 2;
                    o->cand_k = (int *)realloc(o->cand_k, sizeof(int) * cap);
                    o->cand_j = (int *)realloc(o->cand_j, sizeof(int) * cap);
                }
                o->cand_k[n] = k;
                o->cand_j[n] = j;
                n++;
            }
        }
    }
    o->ncand = n;
}

/* main.cu:77-173 for the candidate fine points; later writers (higher k) win the race on
 * b_points unless REF_FLAG_PICK_LOWEST_K (Q12) */
static void o_define_filament(Oracle *o, int it) {
    const int T = o->p.T, p_step = o->p.p_step;
    const double c_space = o->p.c_space, c_num = o->p.c_num;
    const int lowest = (o->p.flags & REF_FLAG_PICK_LOWEST_K) != 0;
#pragma omp parallel for num_threads(o->threads)
    for (int m = 0; m < o->p.c_num; m++) {
        char written[512];
        memset(written, 0, sizeof written);
        for (int q = 0; q < o->ncand; q++) {
            const int k = o->cand_k[q];
            float *s = o->fine + 5 * ((size_t)m * o->ncand + q);
            float *lasts = o->lasts + 2 * ((size_t)m * o->ncand + q);
            float a_n[2 * 7], b_n[2 * 7];
            float arcl = 1. * k / F_LENGTH; /* :100 */
            int phase;
            if (it + m * p_step == T) phase = T; /* :102-103 */
            else phase = (it + m * p_step) % T;
            float offset = 1. * (m - (c_num - 1) / 2.) * c_space; /* :105 */
            for (int n = 0; n < 7; n++) {                         /* :109-135 */
                for (int c = 0; c < 2; c++) {
                    a_n[2 * n + c] = 0.;
                    b_n[2 * n + c] = 0.;
                    for (int pw = 0; pw < 3; pw++) {
                        a_n[2 * n + c] += o->A[n + 14 * pw + 7 * c] * pow(arcl, pw + 1);
                        b_n[2 * n + c] += o->B[n + 14 * pw + 7 * c] * pow(arcl, pw + 1);
                    }
                }
            }
            s[0] = 1. * SCALE * a_n[2 * 0 + 0] * 0.5 + offset; /* :137-139 */
            s[1] = 1. * SCALE * a_n[2 * 0 + 1] * 0.5;
            s[2] = SCALE * arcl;
            for (int n = 1; n < 7; n++) { /* :141-145 */
                s[0] += 1. * SCALE *
                        (a_n[2 * n + 0] * cos(n * 2. * PI_REF * phase / T) +
                         b_n[2 * n + 0] * sin(n * 2. * PI_REF * phase / T));
                s[1] += 1. * SCALE *
                        (a_n[2 * n + 1] * cos(n * 2. * PI_REF * phase / T) +
                         b_n[2 * n + 1] * sin(n * 2. * PI_REF * phase / T));
            }
            if (it > 0) { /* :147-151 */
                s[3] = s[0] - lasts[0];
                s[4] = s[1] - lasts[1];
            }
            lasts[0] = s[0]; /* :154-155 */
            lasts[1] = s[1];
            const int j = m * LENGTH + o->cand_j[q]; /* :158-172 */
            int keep_first = lowest;
            if (o->trace && it < o->trace_steps && o->cand_j[q] == 32)
                keep_first = !o->trace[(size_t)it * o->p.c_num + m];
            if (keep_first && written[o->cand_j[q]]) continue;
            written[o->cand_j[q]] = 1;
            o->b_points[5 * j + 0] = s[0];
            o->b_points[5 * j + 1] = s[1];
            o->b_points[5 * j + 2] = s[3];
            o->b_points[5 * j + 3] = s[4];
        }
    }
}

/* main.cu:176-252; the block barrier at :214 is taken as a barrier over all markers (exact for
 * single-block launches; for 12 cilia the reference itself is racy there, Q13) */
static void o_boundary_check(Oracle *o, int it) {
    const double c_space = o->p.c_space;
    const int c_num = o->p.c_num, XDIM = o->p.XDIM;
    const int r_max = 2 * LENGTH / c_space; /* :185 */
    for (int j = 0; j < o->Ns; j++) {
        o->s[2 * j + 0] = (c_space * c_num) / 2. + o->b_points[5 * j + 0]; /* :193 */
        if (o->s[2 * j + 0] < 0) o->s[2 * j + 0] += XDIM;                  /* :195-196 */
        else if (o->s[2 * j + 0] > XDIM) o->s[2 * j + 0] -= XDIM;
        o->s[2 * j + 1] = o->b_points[5 * j + 1] + 1; /* :198 */
        if (it == 0) {                                /* :200-209 */
            o->u_s[2 * j + 0] = 0.;
            o->u_s[2 * j + 1] = 0.;
        } else {
            o->u_s[2 * j + 0] = o->b_points[5 * j + 2];
            o->u_s[2 * j + 1] = o->b_points[5 * j + 3];
        }
        o->epsilon[j] = 1;
    }
    if (o->replay && it < o->replay_steps) {
        for (int m = 0; m < c_num; m++) {
            const float *r = o->replay + 4 * ((size_t)it * c_num + m);
            const int j = m * LENGTH + 32;
            o->s[2 * j + 0] = r[0];
            o->s[2 * j + 1] = r[1];
            o->u_s[2 * j + 0] = r[2];
            o->u_s[2 * j + 1] = r[3];
        }
    }
    for (int j = 0; j < o->Ns; j++) { /* :217-248 */
        const int m = (j - j % LENGTH) / LENGTH;
        const float x_m = o->s[2 * j + 0], y_m = o->s[2 * j + 1];
        for (int r = 1; r < r_max; r++) {
            const int mm = (m - r < 0) ? (m - r + c_num) : (m - r);
            for (int l = 0; l < LENGTH; l++) {
                const float x_l = o->s[2 * (l + mm * LENGTH) + 0];
                const float y_l = o->s[2 * (l + mm * LENGTH) + 1];
                if (fabsf(x_l - x_m) < 1 && fabsf(y_l - y_m) < 1) o->epsilon[j] = 0;
            }
        }
    }
}

/* ------------------------------------------------------------------ driver */

static double now_s(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}

const char *oracle_kind(void) { return "oracle-port"; }

void *oracle_create(const ref_params *pp) {
    Oracle *o = (Oracle *)calloc(1, sizeof(Oracle));
    o->p = *pp;
    const ref_params *p = &o->p;
    const int size = o->size = p->XDIM * p->YDIM;
    o->length = p->length > 0 ? p->length : LENGTH_REF;
    o->f_length = p->fine_points > 0 ? p->fine_points : F_LENGTH_REF;
    o->scale = p->shape_scale > 0. ? p->shape_scale : SCALE_REF;
    o->flux_back = p->flux_back > 0 ? p->flux_back : FLUX_BACK_REF;
    o->A = (p->flags & REF_FLAG_MUCUS) ? A_mn_mucus : A_mn;
    o->B = (p->flags & REF_FLAG_MUCUS) ? B_mn_mucus : B_mn;
    if (o->length > 512) {
        free(o);
        return NULL;
    }
    o->Ns = LENGTH * p->c_num;
#ifdef _OPENMP
    o->threads = p->threads > 0 ? p->threads : omp_get_max_threads();
#else
    o->threads = 1;
#endif
    o->u = (double *)calloc(2 * (size_t)size, sizeof(double));
    o->rho = (double *)calloc(size, sizeof(double));
    o->f0 = (double *)calloc(9 * (size_t)size, sizeof(double));
    o->f = (double *)calloc(9 * (size_t)size, sizeof(double));
    o->f1 = (double *)calloc(9 * (size_t)size, sizeof(double));
    o->F = (double *)calloc(9 * (size_t)size, sizeof(double));
    o->force = (double *)calloc(2 * (size_t)size, sizeof(double));
    /* main.cu:636-654 plus the two synthetic-input knobs of ref_api.h */
    for (int j = 0; j < size; j++) {
        int y = j / p->XDIM;
        o->rho[j] = 1.;
        o->u[j] = p->u0_amp * sin(2. * M_PI * (y + 0.5) / p->YDIM);
        o->force[j] = p->gx;
        o->force[size + j] = p->gy;
    }
    o->s = (float *)calloc(2 * (size_t)o->Ns + 2, sizeof(float));
    o->u_s = (float *)calloc(2 * (size_t)o->Ns + 2, sizeof(float));
    o->F_s = (float *)calloc(2 * (size_t)o->Ns + 2, sizeof(float));
    o->b_points = (float *)calloc(5 * (size_t)o->Ns + 5, sizeof(float));
    o->epsilon = (int *)calloc((size_t)o->Ns + 1, sizeof(int));
    for (int k = 0; k < o->Ns; k++) o->epsilon[k] = 1; /* main.cu:461-464 */
    build_candidates(o);
    o->fine = (float *)calloc(5 * (size_t)p->c_num * o->ncand + 5, sizeof(float));
    o->lasts = (float *)calloc(2 * (size_t)p->c_num * o->ncand + 2, sizeof(float));
    /* main.cu:722-754: f0 from the initial state, f = f0 */
    o_equilibrium(o);
    memcpy(o->f, o->f0, sizeof(double) * 9 * (size_t)size);
    return o;
}

/* main.cu:817-934 in launch order */
int oracle_step(void *h, int n_steps) {
    Oracle *o = (Oracle *)h;
    const double t0 = now_s();
    for (int n = 0; n < n_steps; n++) {
        const int it = o->it;
        if (o->p.c_num > 0) {
            o_define_filament(o, it);
            o_boundary_check(o, it);
        }
        o_equilibrium(o);
        o_collision(o);
        o_streaming(o);
        o_macro(o);
        if (o->p.c_num > 0) {
            o_interpolate(o);
            o_spread(o, (o->p.flags & REF_FLAG_HALF_FORCE_BODY) ? o->p.gx : 0.,
                     (o->p.flags & REF_FLAG_HALF_FORCE_BODY) ? o->p.gy : 0.);
        } else if (o->p.flags & REF_FLAG_HALF_FORCE_BODY) {
            o_spread(o, o->p.gx, o->p.gy);
        }
        o->it++;
    }
    o->last_seconds = now_s() - t0;
    return 0;
}

void oracle_get_fields(void *h, double *rho, double *u) {
    Oracle *o = (Oracle *)h;
    if (rho) memcpy(rho, o->rho, sizeof(double) * o->size);
    if (u) memcpy(u, o->u, sizeof(double) * 2 * o->size);
}
void oracle_get_force(void *h, double *force) {
    Oracle *o = (Oracle *)h;
    memcpy(force, o->force, sizeof(double) * 2 * o->size);
}
void oracle_get_f(void *h, double *f) {
    Oracle *o = (Oracle *)h;
    memcpy(f, o->f, sizeof(double) * 9 * (size_t)o->size);
}
void oracle_get_markers(void *h, float *s, float *u_s, float *F_s, int *eps) {
    Oracle *o = (Oracle *)h;
    if (s) memcpy(s, o->s, sizeof(float) * 2 * o->Ns);
    if (u_s) memcpy(u_s, o->u_s, sizeof(float) * 2 * o->Ns);
    if (F_s) memcpy(F_s, o->F_s, sizeof(float) * 2 * o->Ns);
    if (eps) memcpy(eps, o->epsilon, sizeof(int) * o->Ns);
}
double oracle_get_Q(void *h) { return ((Oracle *)h)->Q; }
int oracle_get_it(void *h) { return ((Oracle *)h)->it; }
double oracle_last_step_seconds(void *h) { return ((Oracle *)h)->last_seconds; }

/* which fine point feeds marker j (for the product's kinematics table test) */
int oracle_marker_fine_index(void *h, int j) {
    Oracle *o = (Oracle *)h;
    const int lowest = (o->p.flags & REF_FLAG_PICK_LOWEST_K) != 0;
    int best = -1;
    for (int q = 0; q < o->ncand; q++)
        if (o->cand_j[q] == j && (best < 0 || !lowest)) best = o->cand_k[q];
    return best;
}

void oracle_set_marker32_trace(void *h, const unsigned char *trace, int n_steps) {
    Oracle *o = (Oracle *)h;
    free(o->trace);
    o->trace = NULL;
    o->trace_steps = 0;
    if (trace && n_steps > 0) {
        o->trace = (unsigned char *)malloc((size_t)n_steps * o->p.c_num);
        memcpy(o->trace, trace, (size_t)n_steps * o->p.c_num);
        o->trace_steps = n_steps;
    }
}

void oracle_set_marker32_replay(void *h, const float *vals, int n_steps) {
    Oracle *o = (Oracle *)h;
    free(o->replay);
    o->replay = NULL;
    o->replay_steps = 0;
    if (vals && n_steps > 0) {
        const size_t n = (size_t)n_steps * o->p.c_num * 4;
        o->replay = (float *)malloc(sizeof(float) * n);
        memcpy(o->replay, vals, sizeof(float) * n);
        o->replay_steps = n_steps;
    }
}

void oracle_destroy(void *h) {
    Oracle *o = (Oracle *)h;
    free(o->trace);
    free(o->replay);
    free(o->u); free(o->rho); free(o->f0); free(o->f); free(o->f1); free(o->F); free(o->force);
    free(o->s); free(o->u_s); free(o->F_s); free(o->b_points); free(o->epsilon);
    free(o->cand_k); free(o->cand_j); free(o->fine); free(o->lasts);
    free(o);
}

/* ---- derived run parameters, main.cu:291-321,336 (restated for the host-logic tests) ---- */
void oracle_derive(unsigned c_fraction, unsigned c_num, unsigned c_space, double Re, float T_num,
                   unsigned T_pow, float I_pow, unsigned P_num, oracle_derived *d) {
    const unsigned LEN = 96;
    d->XDIM = c_num * c_space;                                  /* :298 */
    d->T = (unsigned)nearbyint(T_num * pow(10, T_pow));         /* :299 */
    d->ITERATIONS = (unsigned)(d->T * I_pow);                   /* :300 (float product) */
    d->INTERVAL = P_num ? d->ITERATIONS / P_num : 0;            /* :301 */
    const double dx = 1. / LEN, dt = 1. / (d->T);               /* :312-313 */
    d->SPEED = 0.8 * 1000 / d->T;                               /* :314 */
    d->t_scale = 1000. * dt * 0.067;                            /* :316, t_0 :34 */
    d->x_scale = 1000000. * dx * 0.000006;                      /* :317, l_0 :33 */
    d->s_scale = d->x_scale / d->t_scale;                       /* :318 */
    d->TAU = (d->SPEED * LEN) / (Re * 0.577 * 0.577) + 1. / 2.; /* :320, host C_S :27 */
    d->TAU2 = 1. / (12. * (d->TAU - (1. / 2.))) + (1. / 2.);    /* :321 */
    d->Ma = 1. * d->SPEED / 0.577;                              /* :408 */
    d->p_step = c_num ? (int)(d->T * c_fraction / c_num) : 0;   /* :336 (unsigned arithmetic) */
}
