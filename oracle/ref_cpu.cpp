/* TEST INFRASTRUCTURE — not part of the product path.
 *
 * CPU harness around the reference's own kernels, compiled UNMODIFIED as host C++ through
 * oracle/shim (see oracle/Makefile: the .cu files are read in place from /root/reference; the
 * two cilia kernels come from the head of main.cu, cut before its host helpers, through a pipe —
 * no reference source is copied into this repository).
 *
 * This file only supplies what main() supplies in the reference: allocation + initial state
 * (main.cu:338-359, 414-464, 636-654, 722-754), launch dimensions (main.cu:366-387) and the
 * per-iteration launch order (main.cu:817-934).
 */
#include <ucontext.h>
#include <omp.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <vector>

#include "cuda_shim.h"
#include "LatticeBoltzmann.cuh"
#include "ImmersedBoundary.cuh"
#include "ref_api.h"

/* signatures of main.cu:77 and main.cu:176 */
void define_filament(const int T, const int it, const double c_space, const int p_step,
                     const double c_num, float *s, float *lasts, float *b_points);
void boundary_check(const double c_space, const int c_num, const int XDIM, const int it,
                    const float *b_points, float *s, float *u_s, int *epsilon);

thread_local shim_dim3 threadIdx, blockIdx, blockDim, gridDim;

/* ------------------------------------------------------------------ fibers */
namespace {

struct BlockRunner {
    static constexpr size_t kStack = 64 * 1024;
    std::vector<ucontext_t> ctx;
    std::vector<char> stacks;
    std::vector<char> done;
    ucontext_t sched;
    int current = -1;
    const std::function<void()> *body = nullptr;
};
thread_local BlockRunner *g_runner = nullptr;

void fiber_entry() {
    BlockRunner *r = g_runner;
    (*r->body)();
    r->done[r->current] = 1;
    swapcontext(&r->ctx[r->current], &r->sched);
}

/* run one block's threads as fibers; __syncthreads() == yield, round-robin ⇒ block barrier */
void run_block_fibers(unsigned nthreads, const std::function<void()> &body) {
    BlockRunner r;
    r.ctx.resize(nthreads);
    r.stacks.resize((size_t)nthreads * BlockRunner::kStack);
    r.done.assign(nthreads, 0);
    r.body = &body;
    g_runner = &r;
    for (unsigned t = 0; t < nthreads; t++) {
        getcontext(&r.ctx[t]);
        r.ctx[t].uc_stack.ss_sp = &r.stacks[(size_t)t * BlockRunner::kStack];
        r.ctx[t].uc_stack.ss_size = BlockRunner::kStack;
        r.ctx[t].uc_link = nullptr;
        makecontext(&r.ctx[t], fiber_entry, 0);
    }
    unsigned remaining = nthreads;
    while (remaining) {
        for (unsigned t = 0; t < nthreads; t++) {
            if (r.done[t]) continue;
            r.current = (int)t;
            threadIdx = shim_dim3(t);
            swapcontext(&r.sched, &r.ctx[t]);
            if (r.done[t]) remaining--;
        }
    }
    g_runner = nullptr;
}

template <class F>
void launch(unsigned grid, unsigned block, bool barrier, int threads, F &&f) {
    std::function<void()> body = f;
#pragma omp parallel for schedule(dynamic, 4) num_threads(threads)
    for (long b = 0; b < (long)grid; b++) {
        blockIdx = shim_dim3((unsigned)b);
        blockDim = shim_dim3(block);
        gridDim = shim_dim3(grid);
        if (barrier) {
            run_block_fibers(block, body);
        } else {
            for (unsigned t = 0; t < block; t++) {
                threadIdx = shim_dim3(t);
                body();
            }
        }
    }
}

}  // namespace

void __syncthreads() {
    BlockRunner *r = g_runner;
    if (r) swapcontext(&r->ctx[r->current], &r->sched);
}

/* ------------------------------------------------------------------ state */
struct RefCpu {
    ref_params p;
    int size, Ns, it;
    int gridsize, blocksize, gridsize2, blocksize2, gridsize3, blocksize3;
    std::vector<double> u, rho, f0, f, f1, force, F;
    std::vector<float> s, u_s, F_s, lasts, boundary, b_points;
    std::vector<int> epsilon;
    double Q;
    double last_seconds;
    int threads;
};

extern "C" {

const char *ref_kind(void) { return "reference-cpu-shim"; }

void *ref_create(const ref_params *pp) {
    RefCpu *h = new RefCpu;
    h->p = *pp;
    const ref_params &p = h->p;
    h->size = p.XDIM * p.YDIM;
    h->Ns = 96 * p.c_num;
    h->it = 0;
    h->Q = 0.;
    h->last_seconds = 0.;
    h->threads = p.threads > 0 ? p.threads : omp_get_max_threads();
    const int size = h->size;

    /* main.cu:366-387 */
    h->blocksize = 128;
    h->gridsize = size / h->blocksize;
    if (h->gridsize * h->blocksize != size) {
        fprintf(stderr, "ref_cpu: XDIM*YDIM must be a multiple of 128 (main.cu:366-368)\n");
        delete h;
        return nullptr;
    }
    h->blocksize2 = p.c_num * 96;
    h->gridsize2 = 1;
    if (h->blocksize2 > 1024) {
        for (h->blocksize2 = 1024; h->blocksize2 > 0; h->blocksize2 -= 96) {
            if ((p.c_num * 96) % h->blocksize2 == 0) {
                h->gridsize2 = (p.c_num * 96) / h->blocksize2;
                break;
            }
        }
    }
    h->blocksize3 = 48;
    h->gridsize3 = 9600 / h->blocksize3 * p.c_num;

    h->u.assign(2 * (size_t)size, 0.);
    h->rho.assign(size, 1.);
    h->f0.assign(9 * (size_t)size, 0.);
    h->f.assign(9 * (size_t)size, 0.);
    h->f1.assign(9 * (size_t)size, 0.);
    h->F.assign(9 * (size_t)size, 0.);
    h->force.assign(2 * (size_t)size, 0.);
    for (int j = 0; j < size; j++) {
        int y = j / p.XDIM;
        h->u[j] = p.u0_amp * sin(2. * M_PI * (y + 0.5) / p.YDIM);
        h->force[j] = p.gx;
        h->force[size + j] = p.gy;
    }
    h->s.assign(2 * (size_t)h->Ns + 2, 0.f);
    h->u_s.assign(2 * (size_t)h->Ns + 2, 0.f);
    h->F_s.assign(2 * (size_t)h->Ns + 2, 0.f);
    h->epsilon.assign((size_t)h->Ns + 64, 1); /* +64: spread reads past the end, Q19 */
    h->lasts.assign(2 * (size_t)p.c_num * 9600 + 2, 0.f);
    h->boundary.assign(5 * (size_t)p.c_num * 9600 + 5, 0.f);
    h->b_points.assign(5 * (size_t)h->Ns + 5, 0.f);

    /* main.cu:722-754: initial equilibrium, f = f0 */
    RefCpu *q = h;
    launch(h->gridsize, h->blocksize, false, h->threads, [q]() {
        equilibrium(q->u.data(), q->rho.data(), q->f0.data(), q->force.data(), q->F.data(),
                    q->p.XDIM, q->p.YDIM, q->p.TAU);
    });
    h->f = h->f0;
    return h;
}

int ref_step(void *hh, int n_steps) {
    RefCpu *h = (RefCpu *)hh;
    const ref_params &p = h->p;
    auto t0 = std::chrono::steady_clock::now();
    for (int n = 0; n < n_steps; n++) {
        const int it = h->it;
        if (p.c_num > 0) {
            launch(h->gridsize3, h->blocksize3, false, h->threads, [h, it]() {
                define_filament(h->p.T, it, h->p.c_space, h->p.p_step, h->p.c_num,
                                h->boundary.data(), h->lasts.data(), h->b_points.data());
            });
            launch(h->gridsize2, h->blocksize2, true, h->gridsize2 > 1 ? 1 : h->threads,
                   [h, it]() {
                       boundary_check(h->p.c_space, h->p.c_num, h->p.XDIM, it, h->b_points.data(),
                                      h->s.data(), h->u_s.data(), h->epsilon.data());
                   });
        }
        launch(h->gridsize, h->blocksize, false, h->threads, [h]() {
            equilibrium(h->u.data(), h->rho.data(), h->f0.data(), h->force.data(), h->F.data(),
                        h->p.XDIM, h->p.YDIM, h->p.TAU);
        });
        launch(h->gridsize, h->blocksize, false, h->threads, [h, it]() {
            collision(h->f0.data(), h->f.data(), h->f1.data(), h->F.data(), h->p.TAU, h->p.TAU2,
                      h->p.XDIM, h->p.YDIM, it);
        });
        launch(h->gridsize, h->blocksize, false, h->threads, [h]() {
            streaming(h->f1.data(), h->f.data(), h->p.XDIM, h->p.YDIM);
        });
        launch(h->gridsize, h->blocksize, false, h->threads, [h]() {
            macro(h->f.data(), h->u.data(), h->rho.data(), h->p.XDIM, h->p.YDIM);
        });
        if (p.c_num > 0) {
            launch(h->gridsize2, h->blocksize2, false, h->threads, [h]() {
                interpolate(h->rho.data(), h->u.data(), h->Ns, h->u_s.data(), h->F_s.data(),
                            h->s.data(), h->p.XDIM, h->p.YDIM);
            });
            launch(h->gridsize, h->blocksize, true, h->threads, [h]() {
                spread(h->rho.data(), h->u.data(), h->f.data(), h->Ns, h->u_s.data(),
                       h->F_s.data(), h->force.data(), h->s.data(), h->p.XDIM, &h->Q,
                       h->epsilon.data());
            });
        }
        h->it++;
    }
    h->last_seconds =
        std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return 0;
}

void ref_get_fields(void *hh, double *rho, double *u) {
    RefCpu *h = (RefCpu *)hh;
    if (rho) memcpy(rho, h->rho.data(), sizeof(double) * h->size);
    if (u) memcpy(u, h->u.data(), sizeof(double) * 2 * h->size);
}
void ref_get_force(void *hh, double *force) {
    RefCpu *h = (RefCpu *)hh;
    memcpy(force, h->force.data(), sizeof(double) * 2 * h->size);
}
void ref_get_f(void *hh, double *f) {
    RefCpu *h = (RefCpu *)hh;
    memcpy(f, h->f.data(), sizeof(double) * 9 * h->size);
}
void ref_get_markers(void *hh, float *s, float *u_s, float *F_s, int *eps) {
    RefCpu *h = (RefCpu *)hh;
    if (s) memcpy(s, h->s.data(), sizeof(float) * 2 * h->Ns);
    if (u_s) memcpy(u_s, h->u_s.data(), sizeof(float) * 2 * h->Ns);
    if (F_s) memcpy(F_s, h->F_s.data(), sizeof(float) * 2 * h->Ns);
    if (eps) memcpy(eps, h->epsilon.data(), sizeof(int) * h->Ns);
}
double ref_get_Q(void *hh) { return ((RefCpu *)hh)->Q; }
int ref_get_it(void *hh) { return ((RefCpu *)hh)->it; }
double ref_last_step_seconds(void *hh) { return ((RefCpu *)hh)->last_seconds; }
void ref_destroy(void *hh) { delete (RefCpu *)hh; }

}  // extern "C"
