// Checkpoint / restart (absent in the reference, whose 100 000-step runs
// restart from scratch).  A checkpoint is the complete time-stepping state of one solver:
//
//   * both population buffers, raw (ghost columns included): the buffer the next step pulls from
//     is all the time loop needs; the other one is what the getters recompute the macroscopic
//     fields, F_s and the force of the last completed iteration from (step.cu: run_moments,
//     materialise_ib), so iblb_get_* right after a restart return what they returned when the
//     checkpoint was written;
//   * the iteration counter, the buffer parity, the per-row-tile flux partial sums (so that Q
//     continues bit for bit: main.cu:391-393, ImmersedBoundary.cu:259-264 never reset it).
//
// Nothing about the markers is stored: the cilia are prescribed (main.cu:77-252 is a function of
// `it` alone; the velocity of iteration `it` is the difference of the positions of `it` and
// `it - 1`, both evaluated directly), so the marker slots are rebuilt on the next step.
//
// File layout (little endian, no padding between the fields below):
//   char magic[8] = "IBLBCKP1"; u64 sizeof(iblb_params); iblb_params; u32 it; i32 cur; i32 nq;
//   i32 reserved; u64 pop_bytes; double qacc[nq]; byte buf0[pop_bytes]; byte buf1[pop_bytes]
// Loading requires a solver created with the same geometry, cilia, relaxation times, body force and
// slab (same_params below); run length, output switches and launch-mode flags may differ.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include <unistd.h>

#include "iblb_internal.h"

using namespace iblb;

namespace {

const char kMagic[8] = {'I', 'B', 'L', 'B', 'C', 'K', 'P', '1'};
constexpr size_t kChunk = (size_t)64 << 20;  // staging buffer for the device <-> file copies

size_t pop_bytes(const iblb_solver *s) { return sizeof(double) * 9 * (size_t)s->L.plane; }

bool coupled(const iblb_solver *s) { return s->connected && !(s->self_left && s->self_right); }

// everything the state depends on; run length, output switches, device and the launch-mode flags
// (graph / fusion / PDL) may differ
bool same_params(const iblb_params &a, const iblb_params &b) {
    const unsigned physics = IBLB_FLAG_HALF_FORCE_BODY | IBLB_FLAG_PICK_HIGHEST_K | IBLB_FLAG_MUCUS;
    return a.c_fraction == b.c_fraction && a.c_num == b.c_num && a.c_space == b.c_space &&
           a.XDIM == b.XDIM && a.YDIM == b.YDIM && a.T == b.T && a.LENGTH == b.LENGTH &&
           a.p_step == b.p_step && a.TAU == b.TAU && a.TAU2 == b.TAU2 && a.gx == b.gx &&
           a.gy == b.gy && a.slab_x0 == b.slab_x0 && a.slab_w == b.slab_w &&
           a.fine_points == b.fine_points && a.shape_scale == b.shape_scale &&
           a.flux_back == b.flux_back &&
           (a.flags & physics) == (b.flags & physics);
}

}  // namespace

extern "C" {

int iblb_save_checkpoint(iblb_solver *s, const char *path) {
    if (!s || !path) {
        set_error("iblb_save_checkpoint: null argument");
        return IBLB_EINVAL;
    }
    IBLB_CUDA(cudaSetDevice(s->device));
    IBLB_CUDA(cudaStreamSynchronize(s->stream));
    // written beside the target and renamed over it once complete and on disk: a crash or kill
    // during the (multi-GB, seconds-long) write leaves the previous checkpoint intact
    const std::string tmp = std::string(path) + ".tmp";
    FILE *fp = fopen(tmp.c_str(), "wb");
    if (!fp) {
        set_error(std::string("iblb_save_checkpoint: cannot open ") + tmp);
        return IBLB_EIO;
    }
    const uint64_t psz = sizeof(iblb_params), pb = pop_bytes(s);
    const uint32_t it = s->it;
    const int32_t cur = s->cur, nq = s->nq, reserved = 0;
    std::vector<double> q(s->nq);
    cudaError_t e = cudaMemcpy(q.data(), s->qacc, sizeof(double) * s->nq, cudaMemcpyDeviceToHost);
    bool ok = e == cudaSuccess;
    ok = ok && fwrite(kMagic, 1, 8, fp) == 8 && fwrite(&psz, 8, 1, fp) == 1 &&
         fwrite(&s->p, sizeof(iblb_params), 1, fp) == 1 && fwrite(&it, 4, 1, fp) == 1 &&
         fwrite(&cur, 4, 1, fp) == 1 && fwrite(&nq, 4, 1, fp) == 1 &&
         fwrite(&reserved, 4, 1, fp) == 1 && fwrite(&pb, 8, 1, fp) == 1 &&
         fwrite(q.data(), sizeof(double), q.size(), fp) == q.size();
    std::vector<char> stage(pb < kChunk ? (size_t)pb : kChunk);
    for (int b = 0; b < 2 && ok; b++) {
        const char *src = reinterpret_cast<const char *>(s->buf[b]);
        for (uint64_t o = 0; o < pb && ok; o += stage.size()) {
            const size_t n = (size_t)((pb - o) < stage.size() ? (pb - o) : stage.size());
            e = cudaMemcpy(stage.data(), src + o, n, cudaMemcpyDeviceToHost);
            ok = e == cudaSuccess && fwrite(stage.data(), 1, n, fp) == n;
        }
    }
    ok = ok && fflush(fp) == 0 && fsync(fileno(fp)) == 0;
    ok = (fclose(fp) == 0) && ok;
    if (e != cudaSuccess) {
        remove(tmp.c_str());
        return cuda_fail(e, "iblb_save_checkpoint");
    }
    if (!ok) {
        remove(tmp.c_str());
        set_error(std::string("iblb_save_checkpoint: short write to ") + tmp);
        return IBLB_EIO;
    }
    if (rename(tmp.c_str(), path) != 0) {
        remove(tmp.c_str());
        set_error(std::string("iblb_save_checkpoint: cannot rename onto ") + path);
        return IBLB_EIO;
    }
    return IBLB_OK;
}

int iblb_load_checkpoint(iblb_solver *s, const char *path) {
    if (!s || !path) {
        set_error("iblb_load_checkpoint: null argument");
        return IBLB_EINVAL;
    }
    if (coupled(s)) {
        // the neighbours' step flags would have to be rewound together with this slab's
        set_error("iblb_load_checkpoint: load every slab before iblb_connect");
        return IBLB_EINVAL;
    }
    IBLB_CUDA(cudaSetDevice(s->device));
    IBLB_CUDA(cudaStreamSynchronize(s->stream));
    FILE *fp = fopen(path, "rb");
    if (!fp) {
        set_error(std::string("iblb_load_checkpoint: cannot open ") + path);
        return IBLB_EIO;
    }
    char magic[8];
    uint64_t psz = 0, pb = 0;
    iblb_params fp_params;
    uint32_t it = 0;
    int32_t cur = 0, nq = 0, reserved = 0;
    bool ok = fread(magic, 1, 8, fp) == 8 && memcmp(magic, kMagic, 8) == 0 &&
              fread(&psz, 8, 1, fp) == 1 && psz == sizeof(iblb_params) &&
              fread(&fp_params, sizeof(iblb_params), 1, fp) == 1 && fread(&it, 4, 1, fp) == 1 &&
              fread(&cur, 4, 1, fp) == 1 && fread(&nq, 4, 1, fp) == 1 &&
              fread(&reserved, 4, 1, fp) == 1 && fread(&pb, 8, 1, fp) == 1;
    if (!ok) {
        fclose(fp);
        if (memcmp(magic, kMagic, 8) == 0 && psz != 0 && psz != sizeof(iblb_params))
            // the layout after the magic starts with sizeof(iblb_params): a library whose parameter
            // block has other fields (an older or newer libiblb) wrote this file
            set_error(std::string("iblb_load_checkpoint: ") + path + " was written by another version of "
                      "libiblb (its iblb_params has " + std::to_string(psz) + " bytes, this build's " +
                      std::to_string(sizeof(iblb_params)) + "); checkpoints do not carry over between them");
        else
            set_error(std::string("iblb_load_checkpoint: ") + path + " is not an IBLBCKP1 checkpoint file");
        return IBLB_EINVAL;
    }
    if (!same_params(s->p, fp_params) || nq != s->nq || pb != pop_bytes(s) || (cur & ~1)) {
        fclose(fp);
        set_error(std::string("iblb_load_checkpoint: ") + path + " was written by a solver with other parameters");
        return IBLB_EINVAL;
    }
    std::vector<double> q(nq);
    ok = fread(q.data(), sizeof(double), q.size(), fp) == q.size();
    cudaError_t e = cudaSuccess;
    if (ok) e = cudaMemcpy(s->qacc, q.data(), sizeof(double) * nq, cudaMemcpyHostToDevice);
    std::vector<char> stage(pb < kChunk ? (size_t)pb : kChunk);
    for (int b = 0; b < 2 && ok && e == cudaSuccess; b++) {
        char *dst = reinterpret_cast<char *>(s->buf[b]);
        for (uint64_t o = 0; o < pb && ok && e == cudaSuccess; o += stage.size()) {
            const size_t n = (size_t)((pb - o) < stage.size() ? (pb - o) : stage.size());
            ok = fread(stage.data(), 1, n, fp) == n;
            if (ok) e = cudaMemcpy(dst + o, stage.data(), n, cudaMemcpyHostToDevice);
        }
    }
    fclose(fp);
    if (e != cudaSuccess) return cuda_fail(e, "iblb_load_checkpoint");
    if (!ok) {
        set_error(std::string("iblb_load_checkpoint: ") + path + " is truncated");
        return IBLB_EIO;  // the solver's state is undefined now: load again or iblb_set_fields
    }
    // step flags of a slab that gets connected later carry flag_base + it (solver.cu: init_state)
    s->flag_base += (int)s->it - (int)it;
    s->it = it;
    s->hist_from = it;  // the flux history is not part of a checkpoint
    s->cur = cur;
    if (s->it_dev) IBLB_CUDA(cudaMemcpy(s->it_dev, &it, sizeof(int), cudaMemcpyHostToDevice));
    // marker slots and tile plans are rebuilt by the next step; F_s and the force band of the last
    // completed iteration by the next getter (from the buffer that iteration pulled from)
    s->kin_base = -1;
    s->lists_base = -1;
    s->pdl_chain = false;
    s->ib_stale = s->p.c_num > 0 && it > 0;
    // a loaded state clears an earlier device-side failure, as iblb_set_fields does
    if (s->err_dev) IBLB_CUDA(cudaMemset(s->err_dev, 0, sizeof(int)));
    return IBLB_OK;
}

}  // extern "C"
