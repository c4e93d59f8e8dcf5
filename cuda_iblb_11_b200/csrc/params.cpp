// Driver surface of the reference's main(): argument parsing and derived run parameters.
// Replaces main.cu:267-321 and main.cu:336.  Pure host C++ (no CUDA).
#include <cmath>
#include <cstdio>
#include <cstring>
#include <sstream>
#include <string>

#include "iblb.h"

namespace iblb {
static thread_local std::string g_error;
void set_error(const std::string &msg) { g_error = msg; }
}  // namespace iblb

extern "C" {

const char *iblb_last_error(void) { return iblb::g_error.c_str(); }
const char *iblb_version(void) { return "libiblb 0.3 (sm_100a)"; }

void iblb_params_default(iblb_params *p) {
    memset(p, 0, sizeof *p);
    // main.cu:267-282
    p->c_fraction = 1;
    p->c_num = 6;
    p->Re = 1.0;
    p->XDIM = 288;
    p->YDIM = 192;
    p->T = 100000;
    p->T_pow = 1;
    p->T_num = 1.0f;
    p->ITERATIONS = p->T;
    p->P_num = 100;
    p->I_pow = 1.0f;
    p->INTERVAL = 500;
    p->LENGTH = 96;
    p->c_space = 48;
    p->ShARC = 0;
    p->BigData = 0;
    p->device = -1;
}

int iblb_params_derive(iblb_params *p) {
    if (p->YDIM == 0) p->YDIM = 192;   // main.cu:271
    if (p->LENGTH == 0) p->LENGTH = 96;  // main.cu:279
    if (p->c_num > 0) p->XDIM = p->c_num * p->c_space;             // main.cu:298
    p->T = (unsigned)nearbyint(p->T_num * pow(10, p->T_pow));      // main.cu:299
    p->ITERATIONS = (unsigned)(p->T * p->I_pow);                   // main.cu:300 (float product)
    p->INTERVAL = p->P_num ? p->ITERATIONS / p->P_num : 0;         // main.cu:301
    if (p->c_num > 0 && p->XDIM < 2 * p->LENGTH) {                 // main.cu:303-308
        std::ostringstream m;
        m << "not enough cilia in simulation! Cilia spacing of " << p->c_space
          << " requires at least " << (p->c_space ? 2 * p->LENGTH / p->c_space : 0) << " cilia";
        iblb::set_error(m.str());
        return IBLB_ESIZE;
    }
    if (p->T == 0) {
        iblb::set_error("beat period T is zero");
        return IBLB_EINVAL;
    }
    const double l_0 = 0.000006, t_0 = 0.067;  // main.cu:33-34
    const double C_S_HOST = 0.577;             // main.cu:27 (differs from the device's 0.57735)
    const double dx = 1. / p->LENGTH;          // main.cu:312
    const double dt = 1. / (p->T);             // main.cu:313
    p->SPEED = 0.8 * 1000 / p->T;              // main.cu:314
    p->t_scale = 1000. * dt * t_0;             // main.cu:316
    p->x_scale = 1000000. * dx * l_0;          // main.cu:317
    p->s_scale = p->x_scale / p->t_scale;      // main.cu:318
    p->TAU = (p->SPEED * p->LENGTH) / (p->Re * C_S_HOST * C_S_HOST) + 1. / 2.;  // main.cu:320
    p->TAU2 = 1. / (12. * (p->TAU - (1. / 2.))) + (1. / 2.);                    // main.cu:321
    p->Ma = 1. * p->SPEED / C_S_HOST;                                           // main.cu:408
    p->p_step = p->c_num ? (int)(p->T * p->c_fraction / p->c_num) : 0;          // main.cu:336
    return IBLB_OK;
}

int iblb_params_from_args(int argc, const char *const *argv, iblb_params *p) {
    iblb_params_default(p);
    const int given = argc - 1;
    if (given == 8) {
        // legacy script line (app.sh:7): c_fraction c_num Re T ITERATIONS P_num ShARC BigData —
        // the declaration order of the non-derived variables of main.cu:267-282
        std::stringstream arg;
        for (int i = 1; i <= 8; i++) arg << argv[i] << ' ';
        unsigned T = 0, ITER = 0;
        bool sh = 0, big = 0;
        arg >> p->c_fraction >> p->c_num >> p->Re >> T >> ITER >> p->P_num >> sh >> big;
        if (T == 0) {
            iblb::set_error("legacy argument line: beat period T must be positive");
            return IBLB_EINVAL;
        }
        p->ShARC = sh;
        p->BigData = big;
        p->c_space = 48;  // main.cu:280
        unsigned pw = 0;
        double num = T;
        while (num >= 10.) {
            num /= 10.;
            pw++;
        }
        p->T_num = (float)num;
        p->T_pow = pw;
        p->I_pow = (float)((double)ITER / (double)T);
        return iblb_params_derive(p);
    }
    if (argc < 11) {  // main.cu:284-289
        std::ostringstream m;
        m << "Too few arguments! " << argc - 1 << " entered of 10 required. ";
        iblb::set_error(m.str());
        return IBLB_EARGS;
    }
    std::stringstream arg;  // main.cu:291-296: one stream, extraction by type
    for (int i = 1; i <= 10; i++) arg << argv[i] << ' ';
    bool sh = 0, big = 0;
    arg >> p->c_fraction >> p->c_num >> p->c_space >> p->Re >> p->T_num >> p->T_pow >> p->I_pow >>
        p->P_num >> sh >> big;
    p->ShARC = sh;
    p->BigData = big;
    return iblb_params_derive(p);
}

}  // extern "C"
