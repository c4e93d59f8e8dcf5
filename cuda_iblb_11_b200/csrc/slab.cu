// libiblb: x-slab coupling between GPUs (one solver per process).
//
// Each slab exports CUDA IPC handles of its two population buffers and of its step flags; after
// iblb_connect the fused LB kernel of a boundary column stores its nine results a second time
// straight into the neighbour's ghost columns through the peer mapping (NVLink / NVSwitch), and
// steps are ordered by counters in peer-visible memory (slab_wait / slab_signal).  No staging
// buffers, no host involvement per step, no collective.
#include <unistd.h>

#include <cstring>
#include <string>

#include "iblb_internal.h"

using namespace iblb;

static_assert(sizeof(cudaIpcMemHandle_t) <= IBLB_IPC_HANDLE_BYTES, "IPC handle size");

// the GPU behind ordinal `device` of this process
static cudaError_t device_uuid(int device, unsigned char (&out)[16]) {
    cudaDeviceProp prop;
    const cudaError_t e = cudaGetDeviceProperties(&prop, device);
    if (e == cudaSuccess) memcpy(out, prop.uuid.bytes, 16);
    return e;
}

extern "C" {

int iblb_export_halo(iblb_solver *s, iblb_halo_handle *h) {
    if (!s || !h) return IBLB_EINVAL;
    IBLB_CUDA(cudaSetDevice(s->device));
    memset(h, 0, sizeof *h);
    cudaIpcMemHandle_t m;
    for (int b = 0; b < 2; b++) {
        IBLB_CUDA(cudaIpcGetMemHandle(&m, s->buf[b]));
        memcpy(h->mem[b], &m, sizeof m);
    }
    IBLB_CUDA(cudaIpcGetMemHandle(&m, s->flags));
    memcpy(h->flag, &m, sizeof m);
    h->plane = (unsigned long long)s->L.plane;
    h->ld = (unsigned)s->L.ld;
    h->ghost = (unsigned)s->L.G;
    h->width = (unsigned)s->L.W;
    h->device = s->device;
    h->pid = (int)getpid();
    IBLB_CUDA(device_uuid(s->device, h->uuid));
    return IBLB_OK;
}

static int map_neighbour(iblb_solver *s, int face, const iblb_halo_handle *h) {
    // same GPU?  By identity: the exporter's ordinal is local to its process (CUDA_VISIBLE_DEVICES
    // may list the GPUs in another order there, or show the same GPU under another number)
    unsigned char mine[16];
    IBLB_CUDA(device_uuid(s->device, mine));
    const bool same_gpu = memcmp(mine, h->uuid, 16) == 0;
    const bool self = (h->pid == (int)getpid() && same_gpu &&
                       h->plane == (unsigned long long)s->L.plane && h->width == (unsigned)s->L.W);
    // a handle exported by this very solver cannot be opened in its own process: use the pointers
    cudaIpcMemHandle_t own;
    bool is_own = false;
    if (self) {
        if (cudaIpcGetMemHandle(&own, s->buf[0]) == cudaSuccess)
            is_own = memcmp(&own, h->mem[0], sizeof own) == 0;
    }
    if (!is_own && same_gpu && h->pid != (int)getpid()) {
        // slab_wait spins until the neighbour's kernel has run: two processes time-slicing one GPU
        // can starve each other for ever
        set_error("iblb_connect: the neighbour is another process on the same GPU "
                  "(one slab per GPU; use iblb_connect_local within one process)");
        return IBLB_EINVAL;
    }
    if ((int)h->ghost != s->L.G || (int)h->ld != s->L.ld) {
        set_error("iblb_connect: neighbour has a different ghost width or column pitch");
        return IBLB_EINVAL;
    }
    for (int b = 0; b < 2; b++) {
        double *base = nullptr;
        if (is_own) {
            base = s->buf[b];
        } else {
            cudaIpcMemHandle_t m;
            memcpy(&m, h->mem[b], sizeof m);
            void *ptr = nullptr;
            IBLB_CUDA(cudaIpcOpenMemHandle(&ptr, m, cudaIpcMemLazyEnablePeerAccess));
            s->nb_mapped[face][b] = ptr;
            base = (double *)ptr;
        }
        s->nb[face][b] = Remote{base, (long long)h->plane, (int)h->ld, (int)h->ghost, (int)h->width};
    }
    int *flags = nullptr;
    if (is_own) {
        flags = s->flags;
    } else {
        cudaIpcMemHandle_t m;
        memcpy(&m, h->flag, sizeof m);
        void *ptr = nullptr;
        IBLB_CUDA(cudaIpcOpenMemHandle(&ptr, m, cudaIpcMemLazyEnablePeerAccess));
        s->nb_flag_mapped[face] = ptr;
        flags = (int *)ptr;
    }
    // I am the RIGHT neighbour of my left neighbour: I write its flags[1] / [3]; and vice versa
    s->nb_flag[face] = &flags[face == 0 ? 1 : 0];
    s->nb_rel[face] = &flags[face == 0 ? 3 : 2];
    if (face == 0) s->self_left = is_own;
    else s->self_right = is_own;
    return IBLB_OK;
}

int iblb_connect(iblb_solver *s, const iblb_halo_handle *left, const iblb_halo_handle *right) {
    if (!s || !left || !right) return IBLB_EINVAL;
    IBLB_CUDA(cudaSetDevice(s->device));
    IBLB_CUDA(cudaStreamSynchronize(s->stream));
    int rc = map_neighbour(s, 0, left);
    if (rc) return rc;
    rc = map_neighbour(s, 1, right);
    if (rc) return rc;
    s->connected = true;
    if (!s->stream_b) {
        // the boundary stream outranks the interior one: its small kernels must not queue behind
        // the interior kernel's thread blocks
        int lo = 0, hi = 0;
        IBLB_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        IBLB_CUDA(cudaStreamCreateWithPriority(&s->stream_b, cudaStreamNonBlocking, hi));
        IBLB_CUDA(cudaEventCreateWithFlags(&s->ev_b, cudaEventDisableTiming));
        IBLB_CUDA(cudaEventCreateWithFlags(&s->ev_i, cudaEventDisableTiming));
        IBLB_CUDA(cudaEventCreateWithFlags(&s->ev_ib, cudaEventDisableTiming));
    }
    if (s->graph_exec) {
        cudaGraphExecDestroy(s->graph_exec);
        s->graph_exec = nullptr;
    }
    // refresh the neighbours' ghost columns of the current source buffer with our boundary columns
    dim3 g((s->L.YDIM + 255) / 256, s->L.G, 9);
    halo_push<<<g, 256, 0, s->stream>>>(s->L, s->buf[s->cur], s->nb[0][s->cur], s->nb[1][s->cur]);
    IBLB_CUDA(cudaGetLastError());
    IBLB_CUDA(cudaStreamSynchronize(s->stream));
    return IBLB_OK;
}

int iblb_connect_local(iblb_solver *s, iblb_solver *left, iblb_solver *right) {
    if (!s || !left || !right) return IBLB_EINVAL;
    IBLB_CUDA(cudaSetDevice(s->device));
    IBLB_CUDA(cudaStreamSynchronize(s->stream));
    iblb_solver *nbs[2] = {left, right};
    for (int f = 0; f < 2; f++) {
        iblb_solver *n = nbs[f];
        if (n->L.G != s->L.G || n->L.ld != s->L.ld || n->L.YDIM != s->L.YDIM) {
            set_error("iblb_connect_local: neighbour geometry differs");
            return IBLB_EINVAL;
        }
        if (n->device != s->device) {
            cudaError_t e = cudaDeviceEnablePeerAccess(n->device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled)
                return cuda_fail(e, "cudaDeviceEnablePeerAccess");
            cudaGetLastError();
        }
        for (int b = 0; b < 2; b++)
            s->nb[f][b] = Remote{n->buf[b], n->L.plane, n->L.ld, n->L.G, n->L.W};
        s->peer[f] = n;
    }
    s->self_left = (left == s);
    s->self_right = (right == s);
    s->connected = true;
    if (!s->ev_boundary) IBLB_CUDA(cudaEventCreateWithFlags(&s->ev_boundary, cudaEventDisableTiming));
    if (s->graph_exec) {
        cudaGraphExecDestroy(s->graph_exec);
        s->graph_exec = nullptr;
    }
    dim3 g((s->L.YDIM + 255) / 256, s->L.G, 9);
    halo_push<<<g, 256, 0, s->stream>>>(s->L, s->buf[s->cur], s->nb[0][s->cur], s->nb[1][s->cur]);
    IBLB_CUDA(cudaGetLastError());
    IBLB_CUDA(cudaStreamSynchronize(s->stream));
    return IBLB_OK;
}

}  // extern "C"
