// `app` — the reference's executable surface (main.cu:263-296): same positional arguments, same
// messages and exit codes for bad input, same output files; the time-step loop is libiblb.
//
//   ./app c_fraction c_num c_space Re T_num T_pow I_pow P_num ShARC BigData      (main.cu:291-296)
//   ./app c_fraction c_num Re T ITERATIONS P_num ShARC BigData                   (app.sh:7 legacy)
//
// Extensions come only through the environment so that the reference's lines run unchanged:
//   IBLB_ROOT   output root (default "Data/"; the reference hard-codes a Windows/cluster path)
//   IBLB_DEVICE CUDA device (default 0; the reference uses 3 when ShARC=1, main.cu:395-396)
//   IBLB_YDIM   channel height (default 192)        IBLB_GX  uniform body force
//   IBLB_ITERATIONS  override the iteration count (smoke runs)
//   IBLB_GPUS   x-slabs over that many GPUs of this machine (default 1), first device IBLB_DEVICE
//   IBLB_CHECKPOINT / IBLB_RESTART   checkpoint file written at every output interval / read at start
#include <cstdlib>
#include <iostream>

#include "iblb.h"

int main(int argc, char *argv[]) {
    iblb_params p;
    int rc = iblb_params_from_args(argc, argv, &p);
    if (rc == IBLB_EARGS || rc == IBLB_ESIZE) {
        std::cout << iblb_last_error() << std::endl;  // main.cu:286,305
        return 1;
    }
    if (rc != IBLB_OK) {
        std::cerr << iblb_last_error() << std::endl;
        return 1;
    }
    if (const char *e = getenv("IBLB_YDIM")) p.YDIM = (unsigned)atoi(e);
    if (const char *e = getenv("IBLB_GX")) p.gx = atof(e);
    if (const char *e = getenv("IBLB_DEVICE")) p.device = atoi(e);
    else p.device = 0;
    if (const char *e = getenv("IBLB_ITERATIONS")) {
        p.ITERATIONS = (unsigned)atoi(e);
        p.INTERVAL = p.P_num ? p.ITERATIONS / p.P_num : 0;
    }
    const char *root = getenv("IBLB_ROOT");
    rc = iblb_run(&p, root);
    if (rc != IBLB_OK) {
        std::cerr << "app: " << iblb_last_error() << std::endl;
        return 1;
    }
    return 0;
}
