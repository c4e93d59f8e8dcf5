// Device-side data layout and kernel parameter blocks of libiblb (sm_100a only).
//
// Population storage (replaces the reference's AoS d_f/d_f0/d_f1/d_F of main.cu:469-483):
//   ONE set of post-collision populations, double-buffered (A/B), structure-of-arrays,
//   y-fastest:   f_i(xl, y)  at  base + i*plane + (xl + G)*ld + y,   xl in [-G, W+G)
//   G ghost columns on each side hold the neighbouring slab's boundary columns (the slab's own
//   opposite edge when one GPU owns the whole periodic channel), so the pull stencil never
//   wraps an index.  ld (column pitch) is a multiple of 16 doubles, so every column is 128-byte
//   aligned and a pair of rows (y0, y0+1), y0 even, is one 16-byte vector.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

// Development build (-DIBLB_CHECKED, `python -m cuda_iblb_11_b200.build --checked`): every
// index the kernels form into the population buffers, the force band, the marker arrays and the
// shared-memory tables is range-checked on the device; violations are counted per translation unit
// and summed by iblb_debug_violations().  Stands in for compute-sanitizer, which is closed on the
// GPU pool (profiles/r02_sanitizer_unavailable.md).  Product builds compile the checks away.
#ifdef IBLB_CHECKED
static __device__ unsigned long long g_iblb_bad;
#define IBLB_CHK(cond)                                  \
    do {                                                \
        if (!(cond)) atomicAdd(&g_iblb_bad, 1ull);      \
    } while (0)
#define IBLB_CHECKED_COUNTER(fn)                                            \
    unsigned long long fn() {                                               \
        unsigned long long v = 0;                                           \
        cudaMemcpyFromSymbol(&v, g_iblb_bad, sizeof v);                     \
        return v;                                                           \
    }
#else
#define IBLB_CHK(cond) do {} while (0)
#define IBLB_CHECKED_COUNTER(fn) unsigned long long fn() { return 0; }
#endif

namespace iblb {

// D2Q9 lattice of LatticeBoltzmann.cu:15-27, i = 0..8
__device__ __constant__ const int kCx[9] = {0, 1, 0, -1, 0, 1, -1, -1, 1};
__device__ __constant__ const int kCy[9] = {0, 0, 1, 0, -1, 1, 1, -1, -1};

constexpr double kCs = 0.57735;  // device sound speed, LatticeBoltzmann.cu:11 (not 1/sqrt(3))
constexpr int kMarkersPerCilium = 96;   // main.cu:83: the reference's value (run-time: IbParams::mpc)
constexpr int kMaxMarkersPerCilium = 128;  // capacity of the per-cilium shared-memory tables
constexpr int kFinePoints = 9600;       // main.cu:81
constexpr double kShapeScale = 111.;    // main.cu:137-145
constexpr int kSpreadTileX = 8, kSpreadTileY = 32;
constexpr int kKinBatch = 32;      // iterations of marker kinematics computed per launch
constexpr int kFluxHist = 4096;  // iterations of per-step flux the solver keeps (iblb_get_flux_history)
constexpr int kSlabFlags = 4;  // slab step counters: [0],[1] data ready from the left / right neighbour; [2],[3] their release
constexpr int kSlabPoison = 0x7fffffff;  // a flag word of a slab whose neighbour failed (above every step number)
constexpr int kMaxLeftCilia = 16;  // r_max = 2*96/c_space (main.cu:185) must not exceed this

struct Lattice {
    int W;          // owned columns
    int G;          // ghost columns per side
    int YDIM;       // rows
    int ld;         // column pitch in doubles
    int XDIM;       // global channel width
    int x0;         // global x of local column 0
    long long plane;  // doubles per population plane = (W + 2G)*ld
    __host__ __device__ long long off(int xl, int y) const {
#ifdef __CUDA_ARCH__
        IBLB_CHK(xl >= -G && xl < W + G && y >= 0 && y < ld);
#endif
        return (long long)(xl + G) * ld + y;
    }
};

// where a slab's outgoing boundary columns go: the neighbour's destination buffer
struct Remote {
    double *base;       // neighbour's population buffer (same A/B parity as our dst); may be ours
    long long plane;    // its plane size
    int ld, G, W;       // its pitch, ghost width and owned width
};

struct LbParams {
    Lattice L;
    const double *src;
    double *dst;
    Remote left, right;          // left neighbour's right ghosts / right neighbour's left ghosts
    const double *band;          // IB force band [2][W][band_ld] (x then y fastest) or nullptr
    int band_rows, band_ld;
    double omega_plus, omega_minus;  // 1/TAU, 1/TAU2 (LatticeBoltzmann.cu:80-81)
    double kforce;                   // 1 - 1/(2 TAU)  (LatticeBoltzmann.cu:56)
    double k1I2, k1I4, k2I2, k2I4;   // kforce * w / cs^2, / cs^4 for the axis (w=1/9) and diagonal (1/36) pairs
    double gx, gy;                   // uniform body force
    double half;                     // 0.5 when the half-force velocity applies, else 0
    double *qacc;                    // per-warp running flux sums [YDIM/64] or nullptr
    int xq;                          // local column of the flux plane (XDIM-5) or -1000
    // flux history: row (iteration % kFluxHist) of qhist [kFluxHist][nq] receives the running sums
    // as they stand after this iteration; iteration = (it_ptr ? *it_ptr : 0) + it
    double *qhist;
    const int *it_ptr;
    int it, nq;
    // columns this launch covers: c in [0, col_count) -> x = col_begin + c, and x += split_skip
    // when x >= split_at (boundary launch = the first and last G columns in one grid)
    int col_begin, col_count, split_at, split_skip;
    int row_begin;  // rows [row_begin, row_end) this launch covers (multiples of 64, of 128 for
    int row_end;    // lb_fused_v4; vector kernels only); row_end == 0 means YDIM
};

struct MarkerArrays {
    float2 *s, *u_s, *F_s;   // position, velocity, force per marker
    int *eps;                // on/off switch
    float2 *bbox;            // per cilium (min x, max x) of the wrapped positions
    // spreading stencil of each marker: the (up to) four columns bx..bx+3 and rows by..by+3 within
    // 1.5 of it and their 1-D delta factors (0 beyond the reach), so that the per-node loops of
    // the spreading only multiply two table entries
    float4 *wx, *wy;
    int2 *base;
};

struct IbParams {
    Lattice L;
    const double *src;       // populations the step pulls from
    MarkerArrays M;
    const float *coef;       // [mpc][28] a_n(x,y), b_n(x,y), n = 0..6 per marker (main.cu:109-135)
    int mpc;                 // markers per cilium (main.cu:83: 96)
    double scale;            // filament length scale (main.cu:137-145: 111)
    int c_num, c_space, T, p_step;
    int it, it_off;          // iteration = (it_ptr ? *it_ptr : it) + it_off
    const int *it_ptr;       // when set, the iteration number lives on the device (CUDA graphs)
    int Ns;
    int r_max;               // main.cu:185
    double *band;            // force band, see LbParams
    int band_rows, band_ld;
    // test hook (iblb_set_marker_replay): marker replay_j of cilium m at iteration it takes
    // (s_x, s_y, u_s_x, u_s_y) = replay[it*c_num + m] for it < replay_steps
    const float4 *replay;
    int replay_steps, replay_j;
};

// single-launch IB+LB iteration (fused_step.cu)
constexpr int kBandMaxRel = 128;  // cilia whose bounding box may overlap one 8-column tile
constexpr int kTileCap = 384;     // markers (epsilon != 0) whose support can touch one 8x64 tile
constexpr int kHitWords = 4;      // precomputed 32-marker hit-mask words per thread
constexpr int kTabX = 12, kTabY = 68;  // a tile's 8x64 nodes plus two rings around them
constexpr int kTabCells = kTabX * kTabY;

// Everything about one band tile at one iteration that does not depend on the fluid (the cilia are
// prescribed), written by ib_tile_lists for a batch of iterations, read by lb_ib_tile_step.
constexpr int kRingNodes = 304;   // nodes of the two rings around an 8x64 tile (the widest)
struct TileProgram {
    int n;                        // markers (epsilon != 0) whose 3x3 support can touch the tile
    int nring;                    // ring nodes that are support nodes of these markers
    int pad[2];
    int list[kTileCap];           // the markers' indices, ascending (the reference's summation order)
    unsigned hit[kHitWords][256]; // per thread (node pair): which of list[32w..32w+31] force its nodes
    short ring[kRingNodes];       // the needed ring nodes
    unsigned char own[256];       // per thread: bit 0/1 = node y0 / y0+1 is a support node
};
static_assert(sizeof(TileProgram) % 16 == 0, "TileProgram alignment");

struct BandParams {
    LbParams lb;
    MarkerArrays M;          // ib_tile_lists: slot 0; lb_ib_tile_step: the iteration's slot
    int c_num, Ns, mpc;
    int band_rows;           // rows that can receive IB force (multiple of 64)
    int tw;                  // columns per tile (6 or 8) = warps per CTA of lb_ib_tile_step
    int tiles_x;             // column tiles = ceil(W/tw)
    TileProgram *prog;       // [slot][band row tile][column tile]
    int *err;                // set non-zero when a tile exceeds kBandMaxRel / kTileCap
#ifdef IBLB_TILE_TIMING
    long long *dbg;          // development: clock64 stamps of lb_ib_tile_step, 16 per CTA
#endif
    int rows_lb;             // row tiles the launch covers (all of them, or the band's on tall grids)
    int band_first;          // lb_ib_tile_step: block order = the band's tiles, then the rows above
};

}  // namespace iblb
