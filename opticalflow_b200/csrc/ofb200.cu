// ofb200 - host side of the C ABI (include/ofb200.h): workspace planning, multigrid hierarchy set-up,
// the BiCGStab driver loop and the launch wrappers.  Compiled for sm_100a only.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cstdarg>
#include <vector>
#include <algorithm>
#include <atomic>
#include <mutex>
#include "../../include/ofb200.h"
#include "misc_kernels.cuh"
#include "stream_sweep.cuh"
#include "comm.h"
#include "halo_exchange.cuh"
#include "coarse_fused.cuh"
#include "legacy_kernels.cuh"
#include "direct_solver.cuh"

using namespace ofb;

namespace {

thread_local char g_error[512] = "";

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
    return code;
}

#define OFB_CHECK_CUDA(expr)                                                                        \
    do {                                                                                            \
        cudaError_t e__ = (expr);                                                                   \
        if (e__ != cudaSuccess) return fail(-1000 - (int)e__, "%s: %s", #expr, cudaGetErrorString(e__)); \
    } while (0)

std::atomic<long long> g_launches{0};     // kernels launched by this process (diagnostic counter only)
#define OFB_CHECK_LAUNCH()                  \
    do {                                    \
        g_launches.fetch_add(1);            \
        OFB_CHECK_CUDA(cudaGetLastError()); \
    } while (0)

#define OFB_CHECK_NCCL(expr)                                                                        \
    do {                                                                                            \
        ncclResult_t r__ = (expr);                                                                  \
        if (r__ != ncclSuccess) return fail(-2000 - (int)r__, "%s: %s", #expr, nccl_api().GetErrorString(r__)); \
    } while (0)

// Launch with the programmatic-stream-serialisation attribute (see pdl_enter in common.cuh) when `pdl` is set.
template <typename... KArgs, typename... Args>
cudaError_t launch_kernel(bool pdl, void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args&&... args) {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, std::forward<Args>(args)...);
}
#define OFB_LAUNCH(pdl, kern, grid, block, smem, stream, ...)                                      \
    do {                                                                                            \
        g_launches.fetch_add(1);                                                                    \
        OFB_CHECK_CUDA(launch_kernel(pdl, kern, grid, block, smem, stream, __VA_ARGS__));           \
    } while (0)

constexpr int MAX_LEVELS = 16;
constexpr int MIN_COARSE = 6;          // keep coarsening while both coarse sizes stay >= this
constexpr int MAX_DENSE_N = 768;       // largest coarsest system solved with the dense inverse

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }


constexpr int MAX_DEVICES = 64;

struct DeviceInfo { int sms; };
inline int device_sms() {
    static std::atomic<int> sms[MAX_DEVICES];
    int dev = 0;
    cudaGetDevice(&dev);
    dev = std::min(std::max(dev, 0), MAX_DEVICES - 1);
    if (sms[dev] == 0) {
        int v = 0;
        cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
        sms[dev] = v > 0 ? v : 148;
    }
    return sms[dev];
}

// Streamed (TMA bulk copy + mbarrier ring) sweep: persistent grid of (CTAs per SM) x SMs.
template <typename T, bool FROM_IMAGE, typename Op>
int launch_stream(LevelArgs<T> L, const T* x, const Op& op, int B, int rows_per_strip, bool skip_last_ghost, cudaStream_t stream,
                  bool pdl = false) {
    using C = StreamCfg<T>;
    using Lay = StreamLayout<T, FROM_IMAGE, Op>;
    auto kern = k_stream<T, FROM_IMAGE, Op>;
    static std::atomic<int> ctas_per_sm[MAX_DEVICES];        // zero-initialised; filled once per device
    int dev = 0;
    cudaGetDevice(&dev);
    dev = std::min(std::max(dev, 0), MAX_DEVICES - 1);
    if (ctas_per_sm[dev] == 0) {
        OFB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Lay::SMEM_BYTES));
        int nb = 0;
        OFB_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kern, C::THREADS, Lay::SMEM_BYTES));
        if (nb < 1) return fail(-9, "streamed sweep does not fit on an SM (%zu bytes of shared memory)", (size_t)Lay::SMEM_BYTES);
        ctas_per_sm[dev] = nb;
    }
    const int slots = device_sms() * ctas_per_sm[dev];
    // On coarse levels the last column is a ghost column that no kernel ever reads from memory (every reader
    // mirrors it from column Ny-3), so tiles need not cover it: 513 columns are one 512-wide tile, not two.
    L.tiles_x = ((skip_last_ghost ? L.g.Ny - 1 : L.g.Ny) + C::TW - 1) / C::TW;
    L.n_pairs = B;
    // wide frames: let concurrently running CTAs read neighbouring column tiles of the same rows (whole DRAM pages);
    // 8192 columns: 0.85 -> 0.90 of the HBM peak, neutral at 1024 columns
    if (L.tiles_x >= 4) L.tile_fastest = 1;
    int R = rows_per_strip;
    if (R <= 0 && L.tiles_x < 4 && !(L.tile_fastest) && !L.balanced) {
        // balanced walk: every persistent CTA gets the same number of rows of the end-to-end (pair, tile) columns
        const long long columns = (long long)L.tiles_x * B;
        const long long total_rows = columns * L.g.Nx;
        if (columns > 0x7fffffffLL) return fail(-10, "too many work items");
        L.balanced = 1;
        L.rows_per_strip = L.g.Nx; L.strips = 1;
        const int grid = (int)std::max<long long>(1, std::min<long long>(slots, total_rows / 8));
        OFB_LAUNCH(pdl, kern, dim3(grid), dim3(C::THREADS), Lay::SMEM_BYTES, stream, L, x, op, (int)columns);
        return 0;
    }
    L.balanced = 0;
    if (R <= 0) {
        // Persistent CTAs take work items round-robin, so a launch lasts ceil(items / slots) rounds of one item
        // each; an item costs its R rows plus the two extra window rows (only the windowed streams re-read them).
        // Pick the strip length that minimises rounds x item cost.
        const double halo = 2.0 * (double)Lay::NWIN / (double)Lay::NROWS;
        double best = 1e300;
        for (int r = 4; r <= 64; ++r) {
            const long long strips = (L.g.Nx + r - 1) / r;
            const long long items = strips * L.tiles_x * B;
            const long long rounds = (items + slots - 1) / slots;
            const double cost = (double)rounds * ((double)r + halo + 0.5);      // +0.5: per-item bookkeeping
            if (cost <= best) { best = cost; R = r; }
        }
    }
    L.rows_per_strip = R;
    L.strips = (L.g.Nx + R - 1) / R;
    const long long n_items = (long long)L.tiles_x * L.strips * B;
    if (n_items > 0x7fffffffLL) return fail(-10, "too many work items");
    const int grid = (int)std::min<long long>(n_items, slots);
    OFB_LAUNCH(pdl, kern, dim3(grid), dim3(C::THREADS), Lay::SMEM_BYTES, stream, L, x, op, (int)n_items);
    return 0;
}

struct LevelPlan {
    int Nx, Ny, pitch;
    long long plane;                    // elements per plane
    int far_i, far_j;                   // far-side ghost rule (Geom::far_i/far_j)
    double delta_i, delta_j;            // distance of the true mirror axis beyond the last interior node, in cells
};

struct Plan {
    int nlevels;                        // 1 = no coarse levels
    LevelPlan lv[MAX_LEVELS];
    bool dense;                         // coarsest level solved with the dense inverse
    int dense_n;
    int fused_from;                     // first level of the shared-memory coarse tail (0 = none)
    size_t fused_smem;                  // bytes of shared memory the tail needs
    // slab decomposition (ofb_solve_slab): levels < gather_level hold this rank's rows (+2 halo/ghost rows), levels
    // >= gather_level hold the whole grid, replicated on every rank
    int gather_level;                   // 0 = not a slab solve
    int stage_rows;                     // rows of the local staging planes used at the gather level
};

struct SlabInfo {
    bool on;
    int rank, world, R, Lg;
    Comm* comm;
};

constexpr size_t FUSED_SMEM_BUDGET = 220 * 1024;

Plan make_plan(int Nx, int Ny, int precond, int max_levels = 0, size_t elem_size = 4, bool legacy_boundaries = false) {
    Plan p;
    p.fused_from = 0;
    p.fused_smem = 0;
    p.gather_level = 0;
    p.stage_rows = 0;
    p.nlevels = 1;
    p.lv[0] = {Nx, Ny, Ny, (long long)Nx * Ny, 2, 2, 0.0, 0.0};
    p.dense = false;
    p.dense_n = 0;
    if (precond != OFB_PRECOND_MG) return p;
    const int level_cap = (max_levels > 0) ? std::min(max_levels, MAX_LEVELS) : MAX_LEVELS;
    while (p.nlevels < level_cap) {
        const LevelPlan& f = p.lv[p.nlevels - 1];
        const int cx = coarse_size(f.Nx), cy = coarse_size(f.Ny);
        if (cx < MIN_COARSE || cy < MIN_COARSE || cx >= f.Nx || cy >= f.Ny) break;
        LevelPlan c;
        c.Nx = cx; c.Ny = cy; c.pitch = (cy + 3) / 4 * 4; c.plane = (long long)cx * c.pitch;
        // Coarse node K sits on fine node 2K-1, so the near mirror axes (node 1) coincide on every level.  The far
        // axis (fine node M = N-2, plus delta) lands on a coarse node only when M is odd; otherwise it lies half a
        // coarse cell beyond the last coarse interior node and the coarse ghost mirrors about that half point.
        // Without this the coarse problem is half a cell short and the two-grid convergence factor drops from
        // 0.17 to 0.54 for even sizes such as 1024 (numpy study).
        c.delta_i = ((f.Nx - 2) % 2 == 1) ? f.delta_i * 0.5 : (f.delta_i + 1.0) * 0.5;
        c.delta_j = ((f.Ny - 2) % 2 == 1) ? f.delta_j * 0.5 : (f.delta_j + 1.0) * 0.5;
        c.far_i = (!legacy_boundaries && c.delta_i >= 0.25) ? 1 : 2;
        c.far_j = (!legacy_boundaries && c.delta_j >= 0.25) ? 1 : 2;
        p.lv[p.nlevels++] = c;
    }
    if (p.nlevels > 1) {
        LevelPlan& c = p.lv[p.nlevels - 1];
        const int n = 3 * c.Nx * c.Ny;
        if (n <= MAX_DENSE_N) {
            p.dense = true; p.dense_n = n;
            c.pitch = c.Ny; c.plane = (long long)c.Nx * c.Ny;      // unpadded: flat index == unknown number
        }
    }
    // coarse tail: the longest suffix of levels (>= 2 levels, level >= 1) whose vectors fit in shared memory
    size_t bytes = 0;
    for (int l = p.nlevels - 1; l >= 1 && p.nlevels - l <= FUSED_MAX_LEVELS; --l) {
        bytes += (size_t)9 * p.lv[l].Nx * p.lv[l].Ny * elem_size;
        if (bytes > FUSED_SMEM_BUDGET) break;
        if (p.nlevels - l >= 2) { p.fused_from = l; p.fused_smem = bytes; }
    }
    return p;
}

// ---- slab decomposition (config d-5: one very large pair over several GPUs) --------------------------------
// Rank k owns the interior rows [1 + k R, 1 + (k+1) R) of level 0 (the last rank whatever is left).  Coarse node K
// sits on fine node 2K-1, so with R a multiple of 2^Lg the same rule holds on every level l <= Lg with R_l = R >> l:
// local fine row i and local coarse row K are related by i = 2K-1 exactly as on the whole grid, and a slab looks
// like a smaller frame whose first/last row is the neighbour's boundary row instead of a mirror ghost.
inline int slab_owned(int N_level, int R_level, int rank) {
    return std::min(std::max(N_level - 2 - R_level * rank, 0), R_level);
}

// Chooses the gather level and the rows per rank.  Returns 0 or an error code.
int slab_partition(int NxG, int Ny, int world, int hint, int* rows_per_rank, int* gather_level) {
    if (world < 2) return fail(-20, "slab decomposition needs at least 2 ranks (got %d)", world);
    const Plan pg = make_plan(NxG, Ny, OFB_PRECOND_MG);
    int max_lg = pg.nlevels - 1;
    if (pg.dense) max_lg -= 1;                         // the dense coarsest level keeps its unpadded layout
    if (max_lg < 1) return fail(-21, "frame %dx%d is too small for a slab-decomposed multigrid solve", NxG, Ny);
    const int per = (NxG - 2 + world - 1) / world;     // interior rows per rank, before rounding
    int Lg = hint;
    if (Lg <= 0) {
        // distribute a level while a rank keeps >= 128 rows of it: below that the sweeps are launch-latency bound
        // and solving the level redundantly on every rank is cheaper than its four halo exchanges per V-cycle
        // (8192^2 over 8 GPUs: gather level 4 / 5 / 6 = 14.8 / 15.2 / 15.9 ms)
        Lg = 1;
        while (Lg < max_lg && (per >> Lg) >= 128) ++Lg;
    }
    Lg = std::min(Lg, max_lg);
    for (; Lg >= 1; --Lg) {
        const int unit = 1 << Lg;
        const int R = (per + unit - 1) / unit * unit;
        bool ok = true;
        for (int l = 0; l <= Lg && ok; ++l) ok = slab_owned(pg.lv[l].Nx, R >> l, world - 1) >= 1;
        if (ok) { *rows_per_rank = R; *gather_level = Lg; return 0; }
    }
    return fail(-22, "cannot split %d rows over %d ranks so that every rank owns rows on every distributed level", NxG, world);
}

Plan make_plan_slab(const ofb_slab& sl, int Ny, int max_levels, size_t elem_size, bool legacy_boundaries) {
    Plan p = make_plan(sl.Nx_global, Ny, OFB_PRECOND_MG, max_levels, elem_size, legacy_boundaries);
    const int Lg = sl.gather_level;
    p.gather_level = Lg;
    for (int l = 0; l < Lg; ++l) {
        const int owned = slab_owned(p.lv[l].Nx, sl.rows_per_rank >> l, sl.rank);
        p.lv[l].Nx = owned + 2;
        p.lv[l].plane = (long long)(owned + 2) * p.lv[l].pitch;
    }
    // the gathered level receives world x R_Lg rows per plane (the last rank pads): leave room behind the plane
    const int gathered_rows = sl.world * (sl.rows_per_rank >> Lg) + 2;
    p.lv[Lg].plane = (long long)std::max(p.lv[Lg].Nx, gathered_rows) * p.lv[Lg].pitch;
    p.stage_rows = (sl.rows_per_rank >> Lg) + 2;
    if (p.fused_from > 0 && p.fused_from < Lg) {       // the shared-memory tail must start on a replicated level
        p.fused_from = (p.nlevels - Lg >= 2) ? Lg : 0;
        p.fused_smem = 0;
        for (int l = p.fused_from; p.fused_from > 0 && l < p.nlevels; ++l) p.fused_smem += (size_t)9 * p.lv[l].Nx * p.lv[l].Ny * elem_size;
    }
    return p;
}

// Workspace carving (all offsets 256-byte aligned).
template <typename T> struct Workspace {
    PairState* st; int* active; int* n_active;
    T *r, *rhat, *p, *v, *s, *t, *ph, *sh, *zalt;              // level-0 vectors
    T* coef[MAX_LEVELS];                                      // level >= 1: 7 coefficient planes per pair
    unsigned short* coef16[MAX_LEVELS];                       // float solves, levels wide enough for the streamed sweep: bfloat16 copies
    int pitch16[MAX_LEVELS];
    T *lb[MAX_LEVELS], *xa[MAX_LEVELS], *xb[MAX_LEVELS], *res[MAX_LEVELS];
    double* dense_M; T* dense_Minv; int* dense_piv;
    T* stage;                                                 // slab solves: local planes at the gather level (7 per pair)
    size_t bytes;
};

template <typename T>
Workspace<T> carve(void* base, int B, const Plan& plan) {
    Workspace<T> w;
    memset(&w, 0, sizeof(w));
    size_t off = 0;
    auto take = [&](size_t nbytes) { void* p = base ? (char*)base + off : nullptr; off = align_up(off + nbytes, 256); return p; };
    w.st = (PairState*)take(sizeof(PairState) * B);
    w.active = (int*)take(sizeof(int) * B);
    w.n_active = (int*)take(sizeof(int) * 4);
    const size_t vec0 = (size_t)3 * plan.lv[0].plane * B * sizeof(T);
    T** v0[] = {&w.r, &w.rhat, &w.p, &w.v, &w.s, &w.t, &w.ph, &w.sh, &w.zalt};
    for (T** q : v0) *q = (T*)take(vec0);
    for (int l = 1; l < plan.nlevels; ++l) {
        const size_t planeb = (size_t)plan.lv[l].plane * B * sizeof(T);
        w.coef[l] = (T*)take(NCOEF * planeb);
        if (sizeof(T) == 4 && plan.gather_level == 0 && plan.lv[l].Ny >= 2 * WARP * VecOf<T>::V) {
            w.pitch16[l] = (plan.lv[l].Ny + 7) / 8 * 8;
            w.coef16[l] = (unsigned short*)take((size_t)NCOEF * plan.lv[l].Nx * w.pitch16[l] * B * sizeof(unsigned short));
        }
        w.lb[l] = (T*)take(3 * planeb);
        w.xa[l] = (T*)take(3 * planeb);
        w.xb[l] = (T*)take(3 * planeb);
        w.res[l] = (T*)take(3 * planeb);
    }
    if (plan.gather_level > 0)
        w.stage = (T*)take((size_t)NCOEF * plan.stage_rows * plan.lv[plan.gather_level].pitch * B * sizeof(T));
    if (plan.dense) {
        const size_t nn = (size_t)plan.dense_n * plan.dense_n;
        w.dense_M = (double*)take(nn * B * sizeof(double));
        w.dense_Minv = (T*)take(nn * B * sizeof(T));
        w.dense_piv = (int*)take((size_t)plan.dense_n * B * sizeof(int));
    }
    w.bytes = off;
    return w;
}

template <typename T> Geom make_geom(const LevelPlan& lp, bool aligned_ok) {
    Geom g;
    g.Nx = lp.Nx; g.Ny = lp.Ny; g.pitch = lp.pitch; g.plane = lp.plane;
    g.pair = 3 * lp.plane; g.img_pair = lp.plane;
    g.aligned = (aligned_ok && lp.pitch % VecOf<T>::V == 0) ? 1 : 0;
    g.far_i = lp.far_i; g.far_j = lp.far_j;
    g.halo_lo = g.halo_hi = 0;
    return g;
}

inline bool ptr_aligned16(const void* p) { return ((uintptr_t)p & 15) == 0; }

int check_shape(int B, int Nx, int Ny, int dtype) {
    if (B < 1) return fail(-1, "batch size must be >= 1 (got %d)", B);
    if (Nx < 5 || Ny < 5) return fail(-2, "frames must be at least 5x5 (got %dx%d)", Nx, Ny);
    if (dtype != OFB_F32 && dtype != OFB_F64) return fail(-3, "dtype must be OFB_F32 or OFB_F64 (got %d)", dtype);
    return 0;
}

}  // namespace

// =====================================================================================================
// Solver driver
// =====================================================================================================
namespace {

// Side stream per device for work that is off a solve's critical path (the dense coarsest-level inverse, which is a
// latency chain on 16 SMs: it runs beside the first fine-level sweeps instead of in front of them).  Shared by the
// solves of a device; ordering between a solve and the side stream is by per-solve events.
inline cudaStream_t side_stream() {
    static cudaStream_t streams[MAX_DEVICES] = {nullptr};
    static std::mutex mu;
    int dev = 0;
    cudaGetDevice(&dev);
    dev = std::min(std::max(dev, 0), MAX_DEVICES - 1);
    std::lock_guard<std::mutex> lock(mu);
    if (streams[dev] == nullptr && cudaStreamCreateWithFlags(&streams[dev], cudaStreamNonBlocking) != cudaSuccess) streams[dev] = nullptr;
    return streams[dev];
}

// Mapped pinned words through which the solve loop reads its counters (k_publish): per host thread and device.
struct PollBox {
    int* host = nullptr; int* dev = nullptr;
    ~PollBox() { if (host != nullptr) cudaFreeHost(host); }      // thread_local: released when the host thread ends
};
inline PollBox* poll_box() {
    thread_local PollBox boxes[MAX_DEVICES];
    int dev = 0;
    cudaGetDevice(&dev);
    dev = std::min(std::max(dev, 0), MAX_DEVICES - 1);
    PollBox& b = boxes[dev];
    if (b.host == nullptr) {
        void *h = nullptr, *d = nullptr;
        if (cudaHostAlloc(&h, 64, cudaHostAllocMapped | cudaHostAllocPortable) != cudaSuccess) { cudaGetLastError(); return nullptr; }
        if (cudaHostGetDevicePointer(&d, h, 0) != cudaSuccess) { cudaGetLastError(); cudaFreeHost(h); return nullptr; }
        b.host = (int*)h;
        b.dev = (int*)d;
    }
    return &b;
}
thread_local int g_last_unconverged = -1;
thread_local int g_last_restarts = 0;         // pairs the calling thread's last solve restarted from their true residual

template <typename T> struct Solver {
    int B;
    cudaEvent_t setup_done = nullptr;    // recorded on the side stream after the dense inverse; waited for before its first use
    ~Solver() {
        if (setup_done != nullptr) {
            cudaStreamWaitEvent(stream, setup_done, 0);      // nothing of this solve may outlive the caller's workspace
            cudaEventDestroy(setup_done);
        }
    }
    int join_setup() {
        if (setup_done == nullptr) return 0;
        OFB_CHECK_CUDA(cudaStreamWaitEvent(stream, setup_done, 0));
        cudaEventDestroy(setup_done);
        setup_done = nullptr;
        return 0;
    }
    Plan plan;
    Workspace<T> w;
    Geom g[MAX_LEVELS];
    const T* P;                      // previous frames (level-0 coefficient source)
    const double *alpha, *beta;
    ofb_solve_options opt;
    cudaStream_t stream;
    const int* active;               // nullptr while setting up / for single applications
    SlabInfo slab;                   // slab.on: levels < slab.Lg are distributed over the ranks of slab.comm
    Geom g_stage;                    // local geometry of the staging planes at the gather level

    // ---- slab communication (no-ops for ordinary solves) ------------------------------------------------
    bool pdl() const { return (opt.flags & OFB_FLAG_PDL) != 0; }
    bool distributed(int l) const { return slab.on && l < slab.Lg; }
    // refresh the halo rows of `nplanes` planes per pair on level l from the neighbouring slabs
    int exchange(int l, T* base, int nplanes, long long pair_stride) {
        if (!distributed(l)) return 0;
        const Geom& G = g[l];
        if (!G.halo_lo && !G.halo_hi) return 0;
        PeerBox& box = slab.comm->box;
        if (box.ok && (opt.flags & OFB_FLAG_SLAB_NCCL_HALO) == 0 &&
            (size_t)B * nplanes * G.pitch * sizeof(T) <= box.slot_bytes) {
            // rows written straight into the neighbour's mailbox by peer stores (halo_exchange.cuh)
            HaloArgs<T> A;
            A.base = base; A.plane = G.plane; A.pair_stride = pair_stride; A.nplanes = nplanes; A.B = B; A.Nx = G.Nx; A.pitch = G.pitch;
            A.local = box.local;
            A.peer[0] = G.halo_lo ? box.peer[0] : nullptr;
            A.peer[1] = G.halo_hi ? box.peer[1] : nullptr;
            A.seq = ++box.seq; A.header = PEERBOX_HEADER; A.slot_bytes = box.slot_bytes;
            A.timeout_ns = 4000000000ull;
            k_halo_exchange<T><<<2, 1024, 0, stream>>>(A);
            OFB_CHECK_LAUNCH();
            return 0;
        }
        NcclApi& api = nccl_api();
        const ncclDataType_t dt = NcclType<T>::value;
        const size_t cnt = (size_t)G.pitch;
        OFB_CHECK_NCCL(api.GroupStart());
        for (int pair = 0; pair < B; ++pair)
            for (int q = 0; q < nplanes; ++q) {
                T* pl = base + (long long)pair * pair_stride + (long long)q * G.plane;
                if (G.halo_lo) {
                    OFB_CHECK_NCCL(api.Send(pl + G.pitch, cnt, dt, slab.rank - 1, slab.comm->comm, stream));
                    OFB_CHECK_NCCL(api.Recv(pl, cnt, dt, slab.rank - 1, slab.comm->comm, stream));
                }
                if (G.halo_hi) {
                    OFB_CHECK_NCCL(api.Send(pl + (long long)(G.Nx - 2) * G.pitch, cnt, dt, slab.rank + 1, slab.comm->comm, stream));
                    OFB_CHECK_NCCL(api.Recv(pl + (long long)(G.Nx - 1) * G.pitch, cnt, dt, slab.rank + 1, slab.comm->comm, stream));
                }
            }
        OFB_CHECK_NCCL(api.GroupEnd());
        return 0;
    }
    int exchange_vec(int l, T* v) { return exchange(l, v, 3, g[l].pair); }
    // sum the partial dot products of all slabs (4 doubles per pair) before a scalar stage
    int allreduce_acc() {
        if (!slab.on) return 0;
        NcclApi& api = nccl_api();
        OFB_CHECK_NCCL(api.GroupStart());
        for (int pair = 0; pair < B; ++pair)
            OFB_CHECK_NCCL(api.AllReduce(w.st[pair].acc, w.st[pair].acc, 4, ncclFloat64, ncclSum, slab.comm->comm, stream));
        OFB_CHECK_NCCL(api.GroupEnd());
        return 0;
    }
    // staging planes (local rows of the gather level) -> whole-grid planes on every rank
    int gather_planes(const T* stage, long long stage_pair, T* whole, long long whole_pair, int nplanes) {
        NcclApi& api = nccl_api();
        const Geom& G = g[slab.Lg];
        const size_t cnt = (size_t)(slab.R >> slab.Lg) * G.pitch;
        OFB_CHECK_NCCL(api.GroupStart());
        for (int pair = 0; pair < B; ++pair)
            for (int q = 0; q < nplanes; ++q)
                OFB_CHECK_NCCL(api.AllGather(stage + (long long)pair * stage_pair + (long long)q * g_stage.plane + G.pitch,
                                             whole + (long long)pair * whole_pair + (long long)q * G.plane + G.pitch, cnt,
                                             NcclType<T>::value, slab.comm->comm, stream));
        OFB_CHECK_NCCL(api.GroupEnd());
        return 0;
    }
    // this rank's window (owned rows + one row either side) of a replicated gather-level vector
    const T* window(const T* whole) const { return whole + (long long)(slab.R >> slab.Lg) * slab.rank * g[slab.Lg].pitch; }
    Geom window_geom() const {
        Geom v = g[slab.Lg];
        v.Nx = g_stage.Nx;
        v.halo_lo = g_stage.halo_lo; v.halo_hi = g_stage.halo_hi;
        return v;
    }

    LevelArgs<T> level_args(int l) const {
        LevelArgs<T> L;
        L.g = g[l];
        if (l == 0) { L.coef = P; L.coef_pair = g[0].img_pair; }
        else { L.coef = w.coef[l]; L.coef_pair = (long long)NCOEF * g[l].plane; }
        L.alpha = alpha; L.beta = beta; L.active = active;
        L.h2 = 1.0 / ((double)(1 << l) * (double)(1 << l));
        L.coef16 = nullptr; L.pitch16 = 0; L.plane16 = 0; L.pair16 = 0;
        if (l >= 1 && w.coef16[l] != nullptr && use_stream(l) && !slab.on && (opt.flags & OFB_FLAG_FP32_COEF) == 0) {
            L.coef16 = w.coef16[l]; L.pitch16 = w.pitch16[l];
            L.plane16 = (long long)g[l].Nx * w.pitch16[l]; L.pair16 = (long long)NCOEF * L.plane16;
        }
        constexpr int V = VecOf<T>::V;
        L.tiles_x = (g[l].Ny + WARP * V - 1) / (WARP * V);
        int R = opt.rows_per_strip;
        if (R <= 0) {
            // one wave of warps when the level is small (latency bound: short strips = more rows in flight),
            // longer strips (fewer halo rows re-read) once the level fills the machine several times over
            const long long target_warps = (long long)device_sms() * 16;
            const long long total_rows = (long long)g[l].Nx * L.tiles_x * B;
            R = (int)std::min<long long>(32, std::max<long long>(1, total_rows / target_warps));
        }
        L.rows_per_strip = R;
        L.strips = (g[l].Nx + R - 1) / R;
        L.tile_fastest = (opt.flags & OFB_FLAG_TILE_FASTEST) ? 1 : 0;
        L.balanced = (opt.flags & OFB_FLAG_STRIP_ITEMS) ? 1 : 0;      // launch_stream: 1 here = balanced walk disabled
        L.n_pairs = B;
        return L;
    }

    // streamed sweep for levels wide enough to fill a tile, 16-byte aligned rows required
    bool use_stream(int l) const {
        return (opt.flags & OFB_FLAG_NO_STREAM) == 0 && g[l].aligned != 0 && g[l].Ny >= 2 * WARP * VecOf<T>::V;
    }
    template <typename Op> int sweep(int l, const T* x, Op op) {
        const LevelArgs<T> L = level_args(l);
        if (use_stream(l)) {
            if (l == 0) return launch_stream<T, true, Op>(L, x, op, B, opt.rows_per_strip, false, stream, pdl());
            return launch_stream<T, false, Op>(L, x, op, B, opt.rows_per_strip, true, stream, pdl());
        }
        const int warps_per_block = 4;
        LevelArgs<T> Lr = L;
        if (opt.rows_per_strip <= 0) {
            // One wave: a small level is latency bound, and a grid that exceeds the resident blocks by a few per cent runs a second,
            // nearly empty wave (130^2 x 64 pairs: 640 blocks on 592 slots).  Lengthen the strips until the launch fits.
            static std::atomic<int> blocks_per_sm[MAX_DEVICES][2];
            int dev = 0;
            cudaGetDevice(&dev);
            dev = std::min(std::max(dev, 0), MAX_DEVICES - 1);
            int nb = blocks_per_sm[dev][l == 0 ? 0 : 1].load();
            if (nb == 0) {
                cudaError_t e = (l == 0) ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_sweep<T, true, Op>, warps_per_block * 32, 0)
                                         : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, k_sweep<T, false, Op>, warps_per_block * 32, 0);
                if (e != cudaSuccess || nb < 1) { cudaGetLastError(); nb = 4; }
                blocks_per_sm[dev][l == 0 ? 0 : 1].store(nb);
            }
            const long long slots = (long long)device_sms() * nb;
            auto blocks_for = [&](int R) {
                const long long strips = (g[l].Nx + R - 1) / R;
                return ((long long)Lr.tiles_x * strips + warps_per_block - 1) / warps_per_block * B;
            };
            int R = Lr.rows_per_strip;
            if (blocks_for(R) > slots && blocks_for(R) <= 2 * slots)
                while (R < 64 && blocks_for(R) > slots) ++R;
            Lr.rows_per_strip = R;
            Lr.strips = (g[l].Nx + R - 1) / R;
        }
        const int items = Lr.tiles_x * Lr.strips;
        dim3 grid((items + warps_per_block - 1) / warps_per_block, B);
        if (l == 0) OFB_LAUNCH(pdl(), (k_sweep<T, true, Op>), grid, dim3(warps_per_block * 32), 0, stream, Lr, x, op);
        else        OFB_LAUNCH(pdl(), (k_sweep<T, false, Op>), grid, dim3(warps_per_block * 32), 0, stream, Lr, x, op);
        return 0;
    }

    double w_pre(int s) const { return (s < 4 && opt.omega_pre[s] > 0.0) ? opt.omega_pre[s] : opt.omega; }
    double w_post(int s) const { return (s < 4 && opt.omega_post[s] > 0.0) ? opt.omega_post[s] : opt.omega; }
    int smooth(int l, const T* b, const T* xin, T* xout, double wj) {
        OpSmooth<T> op; op.g = g[l]; op.cen[0] = b; op.xo = xout; op.w = T(wj);
        return sweep(l, xin, op);
    }
    // first post-smoothing sweep with the prolongation folded in (fp32 image level, streamed sweep only)
    int smooth_prolong(int l, const T* b, const T* xin, T* xout, double wj, const T* ec, const Geom& gc) {
        if constexpr (sizeof(T) == 4) {
            OpSmoothProlong<T> op; op.g = g[l]; op.cen[0] = b; op.xo = xout; op.w = T(wj); op.ec = ec; op.gc = gc;
            const LevelArgs<T> L = level_args(l);
            return launch_stream<T, true, OpSmoothProlong<T>>(L, xin, op, B, opt.rows_per_strip, false, stream, pdl());
        } else {
            return fail(-9, "fused prolongation is float32 only");
        }
    }
    bool fuse_prolong(int l) const {
        return sizeof(T) == 4 && l == 0 && use_stream(0) && opt.nu_post >= 1 && !slab.on && g[1].aligned != 0 &&
               (opt.flags & OFB_FLAG_FUSED_PROLONG) != 0;
    }
    // OFB_FLAG_BF16_RESIDUAL: the V-cycle's residual goes to the restriction and nowhere else, so float solves may store it as
    // bfloat16 (6 of the 13.75 words per pixel the residual + restriction pair moves become 3).  Measured on the B200 (64 pairs of
    // 1024^2): 68.1-68.5 ms per step against 67.3-67.7 with the float32 residual - the residual sweep is issue bound, the restriction
    // reads two-byte scalars, and the rounding costs 0.05 iterations per pair.  Off by default.
    bool half_residual(int l) const {
        return sizeof(T) == 4 && !slab.on && g[l].aligned != 0 && (opt.flags & OFB_FLAG_BF16_RESIDUAL) != 0;
    }
    int residual(int l, const T* b, const T* x, T* r) {
        OpResidual<T, false> op; op.g = g[l]; op.cen[0] = b; op.r = r; op.r2 = nullptr; op.st = w.st; op.acc[0] = 0.0;
        op.half_out = half_residual(l) ? 1 : 0;
        return sweep(l, x, op);
    }
    int jacobi_zero(int l, const T* b, T* z, double wj) {
        OpDirJacobi<T, 2> op; op.g = g[l]; op.cen[0] = b; op.d = nullptr; op.z = z; op.w = T(wj); op.st = w.st; op.acc[0] = 0.0;
        return sweep(l, nullptr, op);
    }

    int build_hierarchy() {
        int rc;
        for (int l = 1; l < plan.nlevels; ++l) {
            // on the gather level of a slab solve the local rows go to the staging planes and are then all-gathered
            const bool to_stage = slab.on && l == slab.Lg;
            const Geom gc = to_stage ? g_stage : g[l];
            T* out = to_stage ? w.stage : w.coef[l];
            const long long out_pair = (long long)NCOEF * gc.plane;
            dim3 block(128), grid((gc.Ny + 127) / 128, gc.Nx, B);
            if (l == 1) k_coef_from_image<T><<<grid, block, 0, stream>>>(g[0], P, gc, out);
            else k_restrict<T, true><<<grid, block, 0, stream>>>(g[l - 1], w.coef[l - 1], (long long)NCOEF * g[l - 1].plane, gc,
                                                                 out, out_pair, NCOEF, nullptr);
            OFB_CHECK_LAUNCH();
            if (to_stage) { if ((rc = gather_planes(w.stage, out_pair, w.coef[l], (long long)NCOEF * g[l].plane, NCOEF))) return rc; }
            else if ((rc = exchange(l, w.coef[l], NCOEF, out_pair))) return rc;
            if constexpr (sizeof(T) == 4) {
                // bfloat16 copies of the planes for the streamed sweeps of this level (the next level is still restricted from fp32)
                if (w.coef16[l] != nullptr && use_stream(l) && !slab.on && (opt.flags & OFB_FLAG_FP32_COEF) == 0) {
                    dim3 pgrid((unsigned)(((long long)g[l].Nx * (w.pitch16[l] / 8) + 255) / 256), 1, B * NCOEF);
                    k_pack_bf16<<<pgrid, 256, 0, stream>>>(g[l], (const float*)w.coef[l], w.coef16[l], w.pitch16[l],
                                                          (long long)g[l].Nx * w.pitch16[l]);
                    OFB_CHECK_LAUNCH();
                }
            }
        }
        if (plan.dense) {
            const int l = plan.nlevels - 1;
            const double h2 = 1.0 / ((double)(1 << l) * (double)(1 << l));
            const size_t mbytes = (size_t)plan.dense_n * plan.dense_n * sizeof(double);
            const int use_smem = mbytes <= 200 * 1024 ? 1 : 0;
            auto kern = k_coarse_setup<T>;
            if (use_smem) {
                static std::atomic<bool> configured[MAX_DEVICES];      // zero-initialised; set once per device
                int dev = 0;
                cudaGetDevice(&dev);
                dev = std::min(std::max(dev, 0), MAX_DEVICES - 1);
                if (!configured[dev]) {
                    OFB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
                    configured[dev] = true;
                }
            }
            cudaStream_t setup_stream = stream;
            cudaStream_t side = (slab.on || (opt.flags & OFB_FLAG_SERIAL_SETUP)) ? nullptr : side_stream();
            cudaEvent_t coef_ready = nullptr;
            if (side != nullptr && cudaEventCreateWithFlags(&coef_ready, cudaEventDisableTiming) == cudaSuccess &&
                cudaEventCreateWithFlags(&setup_done, cudaEventDisableTiming) == cudaSuccess) {
                OFB_CHECK_CUDA(cudaEventRecord(coef_ready, stream));
                OFB_CHECK_CUDA(cudaStreamWaitEvent(side, coef_ready, 0));
                setup_stream = side;
            } else {
                setup_done = nullptr;
                side = nullptr;
            }
            kern<<<B, 1024, use_smem ? mbytes : 0, setup_stream>>>(g[l], w.coef[l], alpha, beta, h2, w.dense_M, w.dense_Minv, w.dense_piv,
                                                                   use_smem);
            OFB_CHECK_LAUNCH();
            if (side != nullptr) OFB_CHECK_CUDA(cudaEventRecord(setup_done, side));
            if (coef_ready != nullptr) cudaEventDestroy(coef_ready);
        }
        return 0;
    }

    // levels lc .. coarsest in one launch, one CTA per pair: w.lb[lc] -> w.xa[lc]
    int fused_tail(int lc) {
        if (int rc = join_setup()) return rc;
        FusedArgs<T> A;
        memset(&A, 0, sizeof(A));
        A.nlev = plan.nlevels - lc;
        int off = 0;
        for (int k = 0; k < A.nlev; ++k) {
            const int l = lc + k;
            FusedLevel<T>& L = A.lv[k];
            L.Nx = g[l].Nx; L.Ny = g[l].Ny; L.pitch = g[l].pitch; L.plane = g[l].plane;
            L.coef = w.coef[l];
            L.h2 = 1.0 / ((double)(1 << l) * (double)(1 << l));
            L.npix = g[l].Nx * g[l].Ny;
            L.far_i = g[l].far_i; L.far_j = g[l].far_j;
            L.smem_off = off;
            off += 9 * L.npix;
        }
        A.b_in = w.lb[lc]; A.x_out = w.xa[lc];
        A.alpha = alpha; A.beta = beta; A.active = active;
        A.nu_pre = opt.nu_pre; A.nu_post = opt.nu_post;
        for (int s = 0; s < 4; ++s) { A.w_pre[s] = w_pre(s); A.w_post[s] = w_post(s); }
        A.dense = plan.dense ? 1 : 0; A.dense_n = plan.dense_n; A.Minv = w.dense_Minv;
        A.coarse_sweeps = opt.coarse_sweeps; A.w_coarse = opt.omega;
        auto kern = k_coarse_vcycle<T>;
        static std::atomic<bool> configured[MAX_DEVICES];      // zero-initialised; set once per device
        int dev = 0;
        cudaGetDevice(&dev);
        dev = std::min(std::max(dev, 0), MAX_DEVICES - 1);
        if (!configured[dev]) {
            OFB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FUSED_SMEM_BUDGET));
            configured[dev] = true;
        }
        const size_t smem = (size_t)off * sizeof(T);
        OFB_LAUNCH(pdl(), kern, dim3(B), dim3(1024), smem, stream, A);
        return 0;
    }

    // number of buffer swaps a V-cycle on level l performs after the first (zero-guess) sweep
    int swaps() const { return (opt.nu_pre - 1) + opt.nu_post; }

    // V-cycle on level l.  `b` = right-hand side, `cur` holds the result of the first sweep w D^-1 b,
    // `alt` is the ping-pong partner.  Returns the buffer holding the result in *out.
    int vcycle(int l, const T* b, T* cur, T* alt, T** out) {
        int rc;
        if ((rc = exchange_vec(l, cur))) return rc;
        for (int s = 1; s < opt.nu_pre; ++s) {
            if ((rc = smooth(l, b, cur, alt, w_pre(s)))) return rc;
            std::swap(cur, alt);
            if ((rc = exchange_vec(l, cur))) return rc;
        }
        const int lc = l + 1;
        const bool gather = slab.on && lc == slab.Lg;          // fine level distributed, coarse level replicated
        T* res = (l == 0) ? w.t : w.res[l];
        if ((rc = residual(l, b, cur, res))) return rc;
        if ((rc = exchange_vec(l, res))) return rc;
        {
            const Geom gc = gather ? g_stage : g[lc];
            T* lb = gather ? w.stage : w.lb[lc];
            dim3 block(128), grid((gc.Ny + 127) / 128, gc.Nx, B);
            if (half_residual(l)) OFB_LAUNCH(pdl(), (k_restrict<T, false, true>), grid, block, 0, stream, g[l], (const T*)res, g[l].pair, gc, lb, gc.pair, 3, active);
            else OFB_LAUNCH(pdl(), (k_restrict<T, false>), grid, block, 0, stream, g[l], (const T*)res, g[l].pair, gc, lb, gc.pair, 3, active);
            if (gather && (rc = gather_planes(w.stage, gc.pair, w.lb[lc], g[lc].pair, 3))) return rc;
        }
        T* ec = nullptr;
        if (plan.fused_from > 0 && lc == plan.fused_from) {
            if ((rc = fused_tail(lc))) return rc;
            ec = w.xa[lc];
        } else if (lc == plan.nlevels - 1) {
            if (plan.dense) {
                if ((rc = join_setup())) return rc;
                const int n = plan.dense_n;
                k_coarse_solve<T><<<B, 128, n * sizeof(T), stream>>>(n, w.dense_Minv, w.lb[lc], w.xa[lc], g[lc].pair, active);
                OFB_CHECK_LAUNCH();
                ec = w.xa[lc];
            } else {
                if ((rc = jacobi_zero(lc, w.lb[lc], w.xa[lc], opt.omega))) return rc;
                T *c = w.xa[lc], *a = w.xb[lc];
                for (int s = 1; s < opt.coarse_sweeps; ++s) { if ((rc = smooth(lc, w.lb[lc], c, a, opt.omega))) return rc; std::swap(c, a); }
                ec = c;
            }
        } else {
            if ((rc = jacobi_zero(lc, w.lb[lc], w.xa[lc], w_pre(0)))) return rc;
            if ((rc = vcycle(lc, w.lb[lc], w.xa[lc], w.xb[lc], &ec))) return rc;
        }
        const bool fused = fuse_prolong(l);      // the correction is added inside the first post-smoothing sweep
        if (!fused) {
            // a distributed fine level below a replicated coarse level interpolates from its window of the coarse grid
            const Geom gcv = gather ? window_geom() : g[lc];
            const T* ecv = gather ? window(ec) : ec;
            bool done = false;
            if constexpr (sizeof(T) == 4) {
                if (g[l].aligned && g[l].pitch % 4 == 0 && g[l].Ny >= 8) {
                    dim3 block(128), grid(((g[l].Ny + 3) / 4 + 127) / 128, g[l].Nx, B);
                    OFB_LAUNCH(pdl(), k_prolong_add4, grid, block, 0, stream, gcv, ecv, g[l], cur, active);
                    done = true;
                }
            }
            if (!done) {
                dim3 block(128), grid((g[l].Ny + 127) / 128, g[l].Nx, B);
                OFB_LAUNCH(pdl(), k_prolong_add<T>, grid, block, 0, stream, gcv, ecv, g[l], cur, active);
            }
        }
        for (int s = 0; s < opt.nu_post; ++s) {
            if ((rc = (fused && s == 0) ? smooth_prolong(l, b, cur, alt, w_post(s), ec, g[lc]) : smooth(l, b, cur, alt, w_post(s)))) return rc;
            std::swap(cur, alt);
            if ((rc = exchange_vec(l, cur))) return rc;
        }
        *out = cur;
        return 0;
    }

    // z = M^-1 d where the first sweep (fused with the direction update) already produced `first`.
    // For block-Jacobi the first sweep (w = 1) is the whole preconditioner.
    int finish_precondition(const T* d, T* first, T* partner, T** out) {
        if (opt.precond != OFB_PRECOND_MG || plan.nlevels == 1) { *out = first; return 0; }
        return vcycle(0, d, first, partner, out);
    }
    double first_sweep_weight() const { return (opt.precond == OFB_PRECOND_MG && plan.nlevels > 1) ? w_pre(0) : 1.0; }
    // buffer the first sweep must write to so that the V-cycle result lands in `target`
    T* first_buffer(T* target) const {
        if (opt.precond != OFB_PRECOND_MG || plan.nlevels == 1) return target;
        return (swaps() % 2 == 0) ? target : w.zalt;
    }
    T* partner_of(T* first, T* target) const { return first == target ? w.zalt : target; }
};

template <typename T>
int init_solver(Solver<T>& S, const void* P, const double* alpha, const double* beta, const ofb_solve_options* opt,
                void* workspace, size_t workspace_bytes, int B, int Nx, int Ny, void* stream, const void* a1, const void* a2,
                const ofb_slab* sl = nullptr) {
    if (opt) S.opt = *opt; else ofb_default_options(&S.opt);
    if (S.opt.nu_pre < 1) S.opt.nu_pre = 1;
    if (S.opt.nu_post < 0) S.opt.nu_post = 0;
    if (S.opt.coarse_sweeps < 1) S.opt.coarse_sweeps = 1;
    if (S.opt.check_every < 1) S.opt.check_every = 1;
    S.B = B;
    memset(&S.slab, 0, sizeof(S.slab));
    const bool legacy = (S.opt.flags & OFB_FLAG_LEGACY_COARSE_BOUNDARY) != 0;
    if (sl != nullptr) {
        if (S.opt.precond != OFB_PRECOND_MG) return fail(-23, "slab solves use the multigrid preconditioner");
        if (sl->comm == nullptr || sl->world < 2 || sl->rank < 0 || sl->rank >= sl->world) return fail(-24, "bad slab descriptor");
        S.plan = make_plan_slab(*sl, Ny, S.opt.max_levels, sizeof(T), legacy);
        const int Lg = sl->gather_level;
        if (Lg < 1 || Lg > S.plan.nlevels - 1 - (S.plan.dense ? 1 : 0) || sl->rows_per_rank % (1 << Lg) != 0)
            return fail(-25, "bad slab descriptor: gather level %d, rows per rank %d", Lg, sl->rows_per_rank);
        if (S.plan.lv[0].Nx != Nx)
            return fail(-26, "rank %d owns %d rows (+2) of the %d-row frame, but Nx = %d was passed", sl->rank, S.plan.lv[0].Nx - 2,
                        sl->Nx_global, Nx);
        S.slab.on = true; S.slab.rank = sl->rank; S.slab.world = sl->world; S.slab.R = sl->rows_per_rank; S.slab.Lg = Lg;
        S.slab.comm = (Comm*)sl->comm;
    } else {
        S.plan = make_plan(Nx, Ny, S.opt.precond, S.opt.max_levels, sizeof(T), legacy);
    }
    if (S.opt.flags & OFB_FLAG_NO_FUSED_TAIL) S.plan.fused_from = 0;
    S.w = carve<T>(workspace, B, S.plan);
    if (workspace == nullptr || workspace_bytes < S.w.bytes)
        return fail(-4, "workspace too small: need %zu bytes, got %zu", S.w.bytes, workspace_bytes);
    const bool al = ptr_aligned16(P) && ptr_aligned16(a1) && ptr_aligned16(a2) && ptr_aligned16(workspace);
    for (int l = 0; l < S.plan.nlevels; ++l) S.g[l] = make_geom<T>(S.plan.lv[l], al);
    if (S.slab.on) {
        const int lo = S.slab.rank > 0 ? 1 : 0, hi = S.slab.rank < S.slab.world - 1 ? 1 : 0;
        for (int l = 0; l < S.slab.Lg; ++l) { S.g[l].halo_lo = lo; S.g[l].halo_hi = hi; }
        LevelPlan st = S.plan.lv[S.slab.Lg];
        const Plan whole = make_plan(sl->Nx_global, Ny, OFB_PRECOND_MG, S.opt.max_levels, sizeof(T), legacy);
        st.Nx = slab_owned(whole.lv[S.slab.Lg].Nx, S.slab.R >> S.slab.Lg, S.slab.rank) + 2;
        st.plane = (long long)S.plan.stage_rows * st.pitch;
        S.g_stage = make_geom<T>(st, al);
        S.g_stage.halo_lo = lo; S.g_stage.halo_hi = hi;
    }
    S.P = (const T*)P; S.alpha = alpha; S.beta = beta; S.stream = (cudaStream_t)stream; S.active = nullptr;
    return 0;
}

template <typename T>
int solve_impl(const void* P, const void* b3, void* x3, const double* alpha, const double* beta, double rtol, int max_it,
               const ofb_solve_options* opt, void* workspace, size_t workspace_bytes, int* iters_out, double* relres_out,
               int* converged_out, double* recurrence_out, int B, int Nx, int Ny, void* stream_, const ofb_slab* sl = nullptr) {
    Solver<T> S;
    memset(&S.slab, 0, sizeof(S.slab));
    int rc = init_solver<T>(S, P, alpha, beta, opt, workspace, workspace_bytes, B, Nx, Ny, stream_, b3, x3, sl);
    if (rc) return rc;
    cudaStream_t stream = S.stream;
    Workspace<T>& w = S.w;
    const Geom g0 = S.g[0];
    const T* b = (const T*)b3;
    T* x = (T*)x3;
    const long long nvec = g0.pair;                        // elements per pair of a level-0 vector
    const int tb = 256;
    const dim3 blas_grid((unsigned)std::min<long long>((nvec + tb * 4 - 1) / (tb * 4), 148 * 8), B);
    const int sb = (B + 127) / 128;

    k_state_init<<<sb, 128, 0, stream>>>(w.st, w.active, B);
    OFB_CHECK_LAUNCH();
    k_norm2<T><<<blas_grid, tb, 0, stream>>>(nvec, nvec, b, w.st);
    OFB_CHECK_LAUNCH();
    if ((rc = S.allreduce_acc())) return rc;
    k_state_bnorm<<<sb, 128, 0, stream>>>(w.st, B);
    OFB_CHECK_LAUNCH();
    if (!S.slab.on) {   // b = 0 (two identical frames): the solution is x = 0 whatever the initial guess says
        k_zero_trivial<T><<<blas_grid, tb, 0, stream>>>(w.st, nvec, nvec, x);
        OFB_CHECK_LAUNCH();
    }
    // p and v enter the first direction update multiplied by beta_k = 0; they must be finite
    OFB_CHECK_CUDA(cudaMemsetAsync(w.p, 0, sizeof(T) * nvec * B, stream));
    OFB_CHECK_CUDA(cudaMemsetAsync(w.v, 0, sizeof(T) * nvec * B, stream));
    if (S.slab.on) OFB_CHECK_CUDA(cudaMemsetAsync(w.stage, 0, (size_t)NCOEF * S.g_stage.plane * B * sizeof(T), stream));
    if (S.opt.precond == OFB_PRECOND_MG && S.plan.nlevels > 1) { if ((rc = S.build_hierarchy())) return rc; }
    if ((rc = S.exchange_vec(0, x))) return rc;

    {   // r = b - S x0, rhat = r, acc[2] = acc[3] = r.r
        OpResidual<T, true> op; op.g = g0; op.cen[0] = b; op.r = w.r; op.r2 = w.rhat; op.st = w.st; op.acc[0] = 0.0; op.half_out = 0;
        if ((rc = S.sweep(0, x, op))) return rc;
    }
    if ((rc = S.allreduce_acc())) return rc;
    k_scalars<<<sb, 128, 0, stream>>>(w.st, w.active, B, 0, rtol, max_it, nullptr);
    OFB_CHECK_LAUNCH();
    S.active = w.active;

    T* ph_first = S.first_buffer(w.ph);
    T* sh_first = S.first_buffer(w.sh);
    const T wj = T(S.first_sweep_weight());
    int host_active = B;
    PollBox* box = poll_box();
    // read a device counter on the host: through the mapped pinned word when there is one (a 4-byte device->host copy
    // would queue on the copy engine behind another stream's bulk transfers)
    auto read_counter = [&](const int* dev_counter, int slot, int* out) -> int {
        if (box != nullptr) {
            k_publish<<<1, 1, 0, stream>>>(dev_counter, box->dev + slot);
            OFB_CHECK_LAUNCH();
            OFB_CHECK_CUDA(cudaStreamSynchronize(stream));
            *out = *(volatile int*)(box->host + slot);
        } else {
            OFB_CHECK_CUDA(cudaMemcpyAsync(out, dev_counter, sizeof(int), cudaMemcpyDeviceToHost, stream));
            OFB_CHECK_CUDA(cudaStreamSynchronize(stream));
        }
        return 0;
    };
    // OFB_FLAG_POLL_AHEAD: the host stays one iteration ahead of the device - iteration k+1 is enqueued BEFORE iteration
    // k's count of running pairs is read (from the mapped word its k_publish wrote, after an event, no stream
    // synchronisation), so the device never waits for the host at an iteration boundary; every kernel honours the
    // per-pair activity mask, so the surplus iteration that is already queued when the count reads 0 runs as no-ops.
    // Measured on the B200 (64 pairs of 1024^2, profiles/r02_poll_ahead_ab.txt): 68.8 vs 67.7 ms per step on one GPU,
    // 80.1 vs 78.2 ms on two - the ~58 empty launches of the surplus iteration (~17 us each: 444 persistent CTAs set up
    // their rings) cost more than the bubble they remove, which is ~0.3 ms per iteration because the host enqueues an
    // iteration 30x faster than the device runs it.  Off by default.
    struct PollEvents {
        cudaEvent_t ev[2] = {nullptr, nullptr};
        ~PollEvents() { for (cudaEvent_t e : ev) if (e != nullptr) cudaEventDestroy(e); }
    } pe;
    bool ahead = box != nullptr && (S.opt.flags & OFB_FLAG_POLL_AHEAD) != 0;
    if (ahead)
        for (int k = 0; k < 2; ++k)
            if (cudaEventCreateWithFlags(&pe.ev[k], cudaEventDisableTiming) != cudaSuccess) { cudaGetLastError(); pe.ev[k] = nullptr; ahead = false; }
    int pending = -1;                   // iteration whose count has been published but not read yet
    auto read_pending = [&]() -> int {
        if (pending < 0) return 0;
        OFB_CHECK_CUDA(cudaEventSynchronize(pe.ev[pending & 1]));
        host_active = *(volatile int*)(box->host + 4 + (pending & 1));
        pending = -1;
        return 0;
    };
    int it = 0;
    g_last_restarts = 0;
    for (int round = 0;; ++round) {
    for (; it < max_it && host_active > 0; ++it) {
        T *ph = nullptr, *sh = nullptr;
        {   // p = r + beta_k (p - omega v) ; z = w D^-1 p
            OpDirJacobi<T, 0> op; op.g = g0; op.cen[0] = w.r; op.cen[1] = w.v; op.cen[2] = w.p; op.d = w.p; op.z = ph_first; op.w = wj; op.st = w.st; op.acc[0] = 0.0;
            if ((rc = S.sweep(0, nullptr, op))) return rc;
        }
        if ((rc = S.finish_precondition(w.p, ph_first, S.partner_of(ph_first, w.ph), &ph))) return rc;
        {   // v = S ph ; acc[0] = rhat.v
            OpApplyDot<T, 1> op; op.g = g0; op.y = w.v; op.cen[0] = w.rhat; op.st = w.st; op.acc[0] = 0.0;
            if ((rc = S.sweep(0, ph, op))) return rc;
        }
        if ((rc = S.allreduce_acc())) return rc;
        OFB_LAUNCH(S.pdl(), k_scalars, dim3(sb), dim3(128), 0, stream, w.st, w.active, B, 1, rtol, max_it, (int*)nullptr);
        {   // s = r - alpha v ; z = w D^-1 s
            OpDirJacobi<T, 1> op; op.g = g0; op.cen[0] = w.r; op.cen[1] = w.v; op.d = w.s; op.z = sh_first; op.w = wj; op.st = w.st;
            op.acc[0] = 0.0;
            if ((rc = S.sweep(0, nullptr, op))) return rc;
        }
        if ((S.opt.flags & OFB_FLAG_NO_HALF_STEP_EXIT) == 0) {
            // pairs whose half-step residual already meets rtol skip the second preconditioner application
            if ((rc = S.allreduce_acc())) return rc;
            OFB_LAUNCH(S.pdl(), k_scalars, dim3(sb), dim3(128), 0, stream, w.st, w.active, B, 15, rtol, max_it, (int*)nullptr);
        }
        if ((rc = S.finish_precondition(w.s, sh_first, S.partner_of(sh_first, w.sh), &sh))) return rc;
        {   // t = S sh ; acc[0] = t.s, acc[1] = t.t
            OpApplyDot<T, 2> op; op.g = g0; op.y = w.t; op.cen[0] = w.s; op.st = w.st; op.acc[0] = op.acc[1] = 0.0;
            if ((rc = S.sweep(0, sh, op))) return rc;
        }
        if ((rc = S.allreduce_acc())) return rc;
        OFB_LAUNCH(S.pdl(), k_scalars, dim3(sb), dim3(128), 0, stream, w.st, w.active, B, 2, rtol, max_it, (int*)nullptr);
        OFB_LAUNCH(S.pdl(), k_update_x_r<T>, blas_grid, dim3(tb), 0, stream, nvec, nvec, x, (const T*)ph, (const T*)sh, w.r, (const T*)w.s,
                   (const T*)w.t, (const T*)w.rhat, w.st);
        const bool poll = ((it + 1) % S.opt.check_every == 0) || (it + 1 == max_it);
        if (poll) OFB_CHECK_CUDA(cudaMemsetAsync(w.n_active, 0, sizeof(int), stream));
        if ((rc = S.allreduce_acc())) return rc;
        OFB_LAUNCH(S.pdl(), k_scalars, dim3(sb), dim3(128), 0, stream, w.st, w.active, B, 3, rtol, max_it, poll ? w.n_active : (int*)nullptr);
        if (!ahead) {
            if (poll && (rc = read_counter(w.n_active, 0, &host_active))) return rc;
            continue;
        }
        if (poll) {
            k_publish<<<1, 1, 0, stream>>>(w.n_active, box->dev + 4 + (it & 1));
            OFB_CHECK_LAUNCH();
            OFB_CHECK_CUDA(cudaEventRecord(pe.ev[it & 1], stream));
        }
        if ((rc = read_pending())) return rc;            // the previous poll: this iteration is already queued behind it
        if (poll) pending = it;
    }
    if (host_active > 0 && (rc = read_pending())) return rc;
    pending = -1;
    {   // complete the ghost ring: x now solves the reference's full system, boundary rows included
        const int ring = 2 * Nx + 2 * Ny;
        k_fill_ghosts<T><<<dim3((ring + 127) / 128, B), 128, 0, stream>>>(g0, x);
        OFB_CHECK_LAUNCH();
        if ((rc = S.exchange_vec(0, x))) return rc;
    }
    {   // the independent residual the reference prints (:1150-1151), in fp64 arithmetic, and the verdict on it
        dim3 grid((Ny + 127) / 128, std::min(std::max(Nx - 2, 1), 64), B);
        k_true_residual<T><<<grid, 128, 0, stream>>>(g0, S.P, (const T*)x, b, alpha, beta, w.st);
        OFB_CHECK_LAUNCH();
        if ((rc = S.allreduce_acc())) return rc;
        OFB_CHECK_CUDA(cudaMemsetAsync(w.n_active + 2, 0, sizeof(int), stream));
        const double u = sizeof(T) == 4 ? 5.9604644775390625e-08 : 1.1102230246251565e-16;
        k_verify<<<sb, 128, 0, stream>>>(w.st, w.active, B, rtol, max_it, u, w.n_active + 2);
        OFB_CHECK_LAUNCH();
        int reactivated = 0;
        if (max_it > 0 && it < max_it && round < MAX_RESTARTS) { if ((rc = read_counter(w.n_active + 2, 2, &reactivated))) return rc; }
        if (reactivated == 0) break;
        g_last_restarts += reactivated;
        // restart BiCGStab for the re-activated pairs from the true residual: r = rhat = b - S x
        S.active = w.active;
        OpResidual<T, true> op; op.g = g0; op.cen[0] = b; op.r = w.r; op.r2 = w.rhat; op.st = w.st; op.acc[0] = 0.0; op.half_out = 0;
        if ((rc = S.sweep(0, x, op))) return rc;
        if ((rc = S.allreduce_acc())) return rc;
        k_restart<<<sb, 128, 0, stream>>>(w.st, B);
        OFB_CHECK_LAUNCH();
        host_active = reactivated;
    }
    }
    // the number of pairs that did not converge goes to the host the same way (ofb_last_unconverged)
    OFB_CHECK_CUDA(cudaMemsetAsync(w.n_active + 1, 0, sizeof(int), stream));
    k_state_export<<<sb, 128, 0, stream>>>(w.st, B, iters_out, relres_out, converged_out, w.n_active + 1, recurrence_out);
    OFB_CHECK_LAUNCH();
    g_last_unconverged = -1;
    if (box != nullptr) {
        int unconv = -1;
        if ((rc = read_counter(w.n_active + 1, 1, &unconv))) return rc;
        g_last_unconverged = unconv;
    }
    if (S.slab.on && S.slab.comm->box.ok) {
        int err = 0;
        OFB_CHECK_CUDA(cudaMemcpyAsync(&err, S.slab.comm->box.local + 4 * sizeof(unsigned long long), sizeof(int),
                                       cudaMemcpyDeviceToHost, stream));
        OFB_CHECK_CUDA(cudaStreamSynchronize(stream));
        if (err) return fail(-40, "halo exchange timed out waiting for a neighbouring rank");
    }
    return 0;
}

template <typename T>
int precondition_impl(const void* P, const void* r3, void* z3, const double* alpha, const double* beta, const ofb_solve_options* opt,
                      void* workspace, size_t workspace_bytes, int B, int Nx, int Ny, void* stream_) {
    Solver<T> S;
    memset(&S.slab, 0, sizeof(S.slab));
    int rc = init_solver<T>(S, P, alpha, beta, opt, workspace, workspace_bytes, B, Nx, Ny, stream_, r3, z3);
    if (rc) return rc;
    const bool mg = (S.opt.precond == OFB_PRECOND_MG && S.plan.nlevels > 1);
    if (mg && (rc = S.build_hierarchy())) return rc;
    T* z = (T*)z3;
    T* first = S.first_buffer(z);
    if ((rc = S.jacobi_zero(0, (const T*)r3, first, S.first_sweep_weight()))) return rc;
    T* out = nullptr;
    if (!mg) return 0;
    if ((rc = S.vcycle(0, (const T*)r3, first, first == z ? S.w.zalt : z, &out))) return rc;
    if (out != z) return fail(-5, "internal: V-cycle result landed in the wrong buffer");
    return 0;
}

}  // namespace

// =====================================================================================================
// C ABI
// =====================================================================================================
namespace {

template <typename T>
int blur_impl(const void* in, void* out, void* tmp, const double* weights, int radius, int F, int Nx, int Ny, cudaStream_t stream) {
    const long long fs = (long long)Nx * Ny;
    if (radius <= 64) {
        // tiled two-pass kernel: pick the tallest tile whose buffers stay under ~96 KB of shared memory
        int TH = 64;
        auto smem_for = [&](int th) {
            return ((size_t)(th + 2 * radius) * BLUR_BW + (size_t)BLUR_WARPS * (BLUR_BW + 2 * radius) + (2 * radius + 1)) * sizeof(T);
        };
        while (TH > 8 && smem_for(TH) > 96 * 1024) TH /= 2;
        TH = std::min(TH, std::max(8, (Nx + 7) / 8 * 8));
        auto kern = k_blur_tile<T>;
        static std::atomic<bool> configured[MAX_DEVICES];      // zero-initialised; set once per device
        int dev = 0;
        cudaGetDevice(&dev);
        dev = std::min(std::max(dev, 0), MAX_DEVICES - 1);
        if (!configured[dev]) {
            OFB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
            configured[dev] = true;
        }
        const int vec_ok = (Ny % 4 == 0 && ptr_aligned16(out)) ? 1 : 0;
        dim3 grid((Ny + BLUR_BW - 1) / BLUR_BW, (Nx + TH - 1) / TH, F);
        kern<<<grid, BLUR_WARPS * 32, smem_for(TH), stream>>>((const T*)in, (T*)out, Nx, Ny, fs, weights, radius, TH, vec_ok);
        OFB_CHECK_LAUNCH();
        return 0;
    }
    dim3 block(128), grid((Ny + 127) / 128, Nx, F);
    k_blur_axis<T, 0><<<grid, block, 0, stream>>>((const T*)in, (T*)tmp, Nx, Ny, fs, weights, radius);
    OFB_CHECK_LAUNCH();
    k_blur_axis<T, 1><<<grid, block, 0, stream>>>((const T*)tmp, (T*)out, Nx, Ny, fs, weights, radius);
    OFB_CHECK_LAUNCH();
    return 0;
}

template <typename T>
int rhs_impl(const void* P, const void* C, void* b3, int B, int Nx, int Ny, cudaStream_t stream) {
    LevelPlan lp{Nx, Ny, Ny, (long long)Nx * Ny, 2, 2, 0.0, 0.0};
    const Geom g = make_geom<T>(lp, false);
    dim3 block(128), grid((Ny + 127) / 128, Nx, B);
    k_rhs<T><<<grid, block, 0, stream>>>(g, (const T*)P, (const T*)C, g.img_pair, (T*)b3);
    OFB_CHECK_LAUNCH();
    return 0;
}

template <typename T>
int apply_full_impl(const void* P, const void* x3, void* y3, const double* alpha, const double* beta, int B, int Nx, int Ny,
                    cudaStream_t stream) {
    LevelPlan lp{Nx, Ny, Ny, (long long)Nx * Ny, 2, 2, 0.0, 0.0};
    const Geom g = make_geom<T>(lp, false);
    dim3 block(128), grid((Ny + 127) / 128, Nx, B);
    k_apply_full<T><<<grid, block, 0, stream>>>(g, (const T*)P, (const T*)x3, (T*)y3, alpha, beta);
    OFB_CHECK_LAUNCH();
    return 0;
}

template <typename T>
int apply_elim_impl(const void* P, const void* x3, void* y3, const double* alpha, const double* beta, int B, int Nx, int Ny,
                    void* stream) {
    Solver<T> S;
    memset(&S.slab, 0, sizeof(S.slab));
    ofb_default_options(&S.opt);
    S.opt.precond = OFB_PRECOND_BJACOBI;
    S.B = B;
    S.plan = make_plan(Nx, Ny, OFB_PRECOND_BJACOBI);
    memset(&S.w, 0, sizeof(S.w));
    S.g[0] = make_geom<T>(S.plan.lv[0], ptr_aligned16(P) && ptr_aligned16(x3) && ptr_aligned16(y3));
    S.P = (const T*)P; S.alpha = alpha; S.beta = beta; S.stream = (cudaStream_t)stream; S.active = nullptr;
    OpApply<T> op; op.g = S.g[0]; op.y = (T*)y3; op.cen[0] = nullptr;
    return S.sweep(0, (const T*)x3, op);
}

template <typename T>
int smooth_impl(const void* P, const void* b3, const void* x_in, void* x_out, const double* alpha, const double* beta,
                const ofb_solve_options* opt, int B, int Nx, int Ny, void* stream) {
    Solver<T> S;
    memset(&S.slab, 0, sizeof(S.slab));
    if (opt) S.opt = *opt; else ofb_default_options(&S.opt);
    S.opt.precond = OFB_PRECOND_BJACOBI;
    S.B = B;
    S.plan = make_plan(Nx, Ny, OFB_PRECOND_BJACOBI);
    memset(&S.w, 0, sizeof(S.w));
    S.g[0] = make_geom<T>(S.plan.lv[0], ptr_aligned16(P) && ptr_aligned16(b3) && ptr_aligned16(x_in) && ptr_aligned16(x_out));
    S.P = (const T*)P; S.alpha = alpha; S.beta = beta; S.stream = (cudaStream_t)stream; S.active = nullptr;
    return S.smooth(0, (const T*)b3, (const T*)x_in, (T*)x_out, S.opt.omega);
}

template <typename T>
int epilogue_impl(const void* P, const void* C, const void* x3, void* vx, void* vy, void* speed, void* gamma, double* functionals,
                  double* moments, const double* alpha, const double* beta, double scale, int B, int Nx, int Ny, cudaStream_t stream,
                  int halo_lo = 0, int halo_hi = 0) {
    LevelPlan lp{Nx, Ny, Ny, (long long)Nx * Ny, 2, 2, 0.0, 0.0};
    Geom g = make_geom<T>(lp, false);
    g.halo_lo = halo_lo; g.halo_hi = halo_hi;
    OFB_CHECK_CUDA(cudaMemsetAsync(functionals, 0, sizeof(double) * 3 * B, stream));
    if (moments) OFB_CHECK_CUDA(cudaMemsetAsync(moments, 0, sizeof(double) * 4 * B, stream));
    dim3 block(128), grid((Ny + 127) / 128, std::min(Nx, 32), B);
    k_epilogue<T><<<grid, block, 0, stream>>>(g, (const T*)P, (const T*)C, g.img_pair, (const T*)x3, (T*)vx, (T*)vy, (T*)speed,
                                              (T*)gamma, functionals, moments, alpha, beta, scale);
    OFB_CHECK_LAUNCH();
    return 0;
}

}  // namespace

extern "C" {

int ofb_version(void) { return OFB_ABI_VERSION; }

const char* ofb_last_error(void) { return g_error; }

long long ofb_launch_count(void) { return g_launches.load(); }

int ofb_last_unconverged(void) { return g_last_unconverged; }

int ofb_last_restarts(void) { return g_last_restarts; }

void ofb_default_options(ofb_solve_options* opt) {
    if (!opt) return;
    // V(1,2) cycle: plain block-Jacobi first sweep on the zero guess, two post-sweeps with Chebyshev-like
    // damping (1.3, 0.5).  Scanned on the B200 against V(2,2)/0.8 (tools/solver_scan.py, profiles/): 30% less
    // time per solve at the headline regime.  OFB_CYCLE_ROBUST is the slower, more forgiving V(2,2)/0.8.
    opt->precond = OFB_PRECOND_MG;
    opt->nu_pre = 1;
    opt->nu_post = 2;
    opt->omega = 0.8;
    opt->check_every = 1;
    opt->rows_per_strip = 0;
    opt->coarse_sweeps = 30;
    opt->flags = 0;
    for (int s = 0; s < 4; ++s) opt->omega_pre[s] = opt->omega_post[s] = 0.0;
    opt->omega_pre[0] = 1.0;
    opt->omega_post[0] = 1.3;
    opt->omega_post[1] = 0.5;
    opt->max_levels = 0;
    opt->reserved = 0;
}

size_t ofb_workspace_bytes(int B, int Nx, int Ny, int dtype, int precond) {
    if (check_shape(B, Nx, Ny, dtype)) return 0;
    const Plan plan = make_plan(Nx, Ny, precond);
    return dtype == OFB_F32 ? carve<float>(nullptr, B, plan).bytes : carve<double>(nullptr, B, plan).bytes;
}

int ofb_blur(const void* in, void* out, void* tmp, const double* weights, int radius, int F, int Nx, int Ny, int dtype, void* stream) {
    if (F < 1 || Nx < 1 || Ny < 1) return fail(-1, "ofb_blur: bad shape %d x %d x %d", F, Nx, Ny);
    if (dtype != OFB_F32 && dtype != OFB_F64) return fail(-3, "dtype must be OFB_F32 or OFB_F64 (got %d)", dtype);
    if (radius < 0 || !weights) return fail(-6, "ofb_blur: weights/radius invalid");
    if (in == out || in == tmp || out == tmp) return fail(-7, "ofb_blur: in/out/tmp must not alias");
    return dtype == OFB_F32 ? blur_impl<float>(in, out, tmp, weights, radius, F, Nx, Ny, (cudaStream_t)stream)
                            : blur_impl<double>(in, out, tmp, weights, radius, F, Nx, Ny, (cudaStream_t)stream);
}

int ofb_build_rhs(const void* P, const void* C, void* b3, int B, int Nx, int Ny, int dtype, void* stream) {
    if (int rc = check_shape(B, Nx, Ny, dtype)) return rc;
    return dtype == OFB_F32 ? rhs_impl<float>(P, C, b3, B, Nx, Ny, (cudaStream_t)stream)
                            : rhs_impl<double>(P, C, b3, B, Nx, Ny, (cudaStream_t)stream);
}

int ofb_apply(const void* P, const void* x3, void* y3, const double* alpha, const double* beta, int B, int Nx, int Ny, int dtype,
              void* stream) {
    if (int rc = check_shape(B, Nx, Ny, dtype)) return rc;
    if (x3 == y3) return fail(-7, "ofb_apply: x and y must not alias");
    return dtype == OFB_F32 ? apply_full_impl<float>(P, x3, y3, alpha, beta, B, Nx, Ny, (cudaStream_t)stream)
                            : apply_full_impl<double>(P, x3, y3, alpha, beta, B, Nx, Ny, (cudaStream_t)stream);
}

int ofb_apply_eliminated(const void* P, const void* x3, void* y3, const double* alpha, const double* beta, int B, int Nx, int Ny,
                         int dtype, void* stream) {
    if (int rc = check_shape(B, Nx, Ny, dtype)) return rc;
    if (x3 == y3) return fail(-7, "ofb_apply_eliminated: x and y must not alias");
    return dtype == OFB_F32 ? apply_elim_impl<float>(P, x3, y3, alpha, beta, B, Nx, Ny, stream)
                            : apply_elim_impl<double>(P, x3, y3, alpha, beta, B, Nx, Ny, stream);
}

int ofb_smooth(const void* P, const void* b3, const void* x_in, void* x_out, const double* alpha, const double* beta,
               const ofb_solve_options* opt, int B, int Nx, int Ny, int dtype, void* stream) {
    if (int rc = check_shape(B, Nx, Ny, dtype)) return rc;
    if (x_in == x_out) return fail(-7, "ofb_smooth: x_in and x_out must not alias (Jacobi reads neighbours of x_in)");
    return dtype == OFB_F32 ? smooth_impl<float>(P, b3, x_in, x_out, alpha, beta, opt, B, Nx, Ny, stream)
                            : smooth_impl<double>(P, b3, x_in, x_out, alpha, beta, opt, B, Nx, Ny, stream);
}

int ofb_solve(const void* P, const void* b3, void* x3, const double* alpha, const double* beta, double rtol, int max_it,
              const ofb_solve_options* opt, void* workspace, size_t workspace_bytes, int* iters_out, double* relres_out,
              int* converged_out, double* recurrence_out, int B, int Nx, int Ny, int dtype, void* stream) {
    if (int rc = check_shape(B, Nx, Ny, dtype)) return rc;
    if (max_it < 0) return fail(-8, "max_it must be >= 0");
    return dtype == OFB_F32
               ? solve_impl<float>(P, b3, x3, alpha, beta, rtol, max_it, opt, workspace, workspace_bytes, iters_out, relres_out,
                                   converged_out, recurrence_out, B, Nx, Ny, stream)
               : solve_impl<double>(P, b3, x3, alpha, beta, rtol, max_it, opt, workspace, workspace_bytes, iters_out, relres_out,
                                    converged_out, recurrence_out, B, Nx, Ny, stream);
}

int ofb_precondition(const void* P, const void* r3, void* z3, const double* alpha, const double* beta, const ofb_solve_options* opt,
                     void* workspace, size_t workspace_bytes, int B, int Nx, int Ny, int dtype, void* stream) {
    if (int rc = check_shape(B, Nx, Ny, dtype)) return rc;
    return dtype == OFB_F32 ? precondition_impl<float>(P, r3, z3, alpha, beta, opt, workspace, workspace_bytes, B, Nx, Ny, stream)
                            : precondition_impl<double>(P, r3, z3, alpha, beta, opt, workspace, workspace_bytes, B, Nx, Ny, stream);
}

int ofb_epilogue(const void* P, const void* C, const void* x3, void* vx, void* vy, void* speed, void* gamma, double* functionals,
                 double* moments, const double* alpha, const double* beta, double scale, int B, int Nx, int Ny, int dtype, void* stream) {
    if (int rc = check_shape(B, Nx, Ny, dtype)) return rc;
    return dtype == OFB_F32 ? epilogue_impl<float>(P, C, x3, vx, vy, speed, gamma, functionals, moments, alpha, beta, scale, B, Nx, Ny,
                                                   (cudaStream_t)stream)
                            : epilogue_impl<double>(P, C, x3, vx, vy, speed, gamma, functionals, moments, alpha, beta, scale, B, Nx, Ny,
                                                    (cudaStream_t)stream);
}

// ---------------------------------------------------------------------------------------- slab decomposition
int ofb_slab_partition(int Nx_global, int Ny, int world, int gather_level_hint, int* rows_per_rank, int* gather_level) {
    if (Nx_global < 5 || Ny < 5) return fail(-2, "frames must be at least 5x5 (got %dx%d)", Nx_global, Ny);
    if (rows_per_rank == nullptr || gather_level == nullptr) return fail(-9, "null output pointer");
    return slab_partition(Nx_global, Ny, world, gather_level_hint, rows_per_rank, gather_level);
}

int ofb_slab_local_rows(const ofb_slab* slab, int* first_row, int* owned_rows) {
    if (slab == nullptr || slab->world < 1 || slab->rank < 0 || slab->rank >= slab->world) return fail(-24, "bad slab descriptor");
    const int owned = slab_owned(slab->Nx_global, slab->rows_per_rank, slab->rank);
    if (first_row) *first_row = 1 + slab->rows_per_rank * slab->rank;
    if (owned_rows) *owned_rows = owned;
    return 0;
}

int ofb_comm_unique_id(void* id128) {
    NcclApi& api = nccl_api();
    if (!api.ok) return fail(-30, "NCCL is not available in this process (libnccl.so.2 could not be resolved)");
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    ncclUniqueId id;
    OFB_CHECK_NCCL(api.GetUniqueId(&id));
    memcpy(id128, &id, sizeof(id));
    return 0;
}

int ofb_comm_create(const void* id128, int rank, int world, void** comm_out) {
    NcclApi& api = nccl_api();
    if (!api.ok) return fail(-30, "NCCL is not available in this process (libnccl.so.2 could not be resolved)");
    if (comm_out == nullptr || id128 == nullptr || world < 1 || rank < 0 || rank >= world) return fail(-31, "bad communicator arguments");
    ncclUniqueId id;
    memcpy(&id, id128, sizeof(id));
    Comm* c = new Comm;
    memset(c, 0, sizeof(*c));
    c->rank = rank; c->world = world;
    ncclResult_t r = api.CommInitRank(&c->comm, world, id, rank);
    if (r != ncclSuccess) { delete c; return fail(-2000 - (int)r, "ncclCommInitRank: %s", api.GetErrorString(r)); }
    // Peer mailboxes for the halo rows: allocate, publish the IPC handle through the communicator, map the neighbours'.
    // Any failure on any rank leaves box.ok false everywhere and the exchanges fall back to ncclSend/ncclRecv.
    PeerBox& box = c->box;
    box.slot_bytes = PEERBOX_SLOT_BYTES;
    const size_t total = PEERBOX_HEADER + 4 * box.slot_bytes;
    int good = 1;
    cudaIpcMemHandle_t mine;
    memset(&mine, 0, sizeof(mine));
    if (cudaMalloc((void**)&box.local, total) != cudaSuccess) { box.local = nullptr; good = 0; }
    if (good && cudaMemset(box.local, 0, total) != cudaSuccess) good = 0;
    if (good && cudaIpcGetMemHandle(&mine, box.local) != cudaSuccess) good = 0;
    cudaGetLastError();
    unsigned char* dbuf = nullptr;
    const size_t hb = sizeof(cudaIpcMemHandle_t);
    std::vector<unsigned char> all(hb * world);
    if (cudaMalloc((void**)&dbuf, hb * world + 2 * sizeof(int)) != cudaSuccess) {
        api.CommDestroy(c->comm); delete c;
        return fail(-32, "cudaMalloc failed while setting up the communicator");
    }
    OFB_CHECK_CUDA(cudaMemcpy(dbuf + hb * rank, &mine, hb, cudaMemcpyHostToDevice));
    OFB_CHECK_NCCL(api.AllGather(dbuf + hb * rank, dbuf, hb, ncclChar, c->comm, 0));
    OFB_CHECK_CUDA(cudaStreamSynchronize(0));
    OFB_CHECK_CUDA(cudaMemcpy(all.data(), dbuf, hb * world, cudaMemcpyDeviceToHost));
    for (int side = 0; side < 2 && good; ++side) {
        const int nb = side == 0 ? rank - 1 : rank + 1;
        if (nb < 0 || nb >= world) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, all.data() + hb * nb, hb);
        void* mapped = nullptr;
        if (cudaIpcOpenMemHandle(&mapped, h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { good = 0; cudaGetLastError(); }
        else box.peer[side] = (unsigned char*)mapped;
    }
    int* dflag = (int*)(dbuf + hb * world);
    OFB_CHECK_CUDA(cudaMemcpy(dflag, &good, sizeof(int), cudaMemcpyHostToDevice));
    OFB_CHECK_NCCL(api.AllReduce(dflag, dflag + 1, 1, ncclInt32, ncclMin, c->comm, 0));
    OFB_CHECK_CUDA(cudaStreamSynchronize(0));
    int all_good = 0;
    OFB_CHECK_CUDA(cudaMemcpy(&all_good, dflag + 1, sizeof(int), cudaMemcpyDeviceToHost));
    cudaFree(dbuf);
    box.ok = all_good != 0;
    box.seq = 0;
    *comm_out = c;
    return 0;
}

int ofb_comm_peer_halo(void* comm) { return (comm != nullptr && ((Comm*)comm)->box.ok) ? 1 : 0; }

int ofb_comm_destroy(void* comm) {
    if (comm == nullptr) return 0;
    Comm* c = (Comm*)comm;
    NcclApi& api = nccl_api();
    cudaDeviceSynchronize();
    for (int side = 0; side < 2; ++side) if (c->box.peer[side]) cudaIpcCloseMemHandle(c->box.peer[side]);
    if (c->box.local) cudaFree(c->box.local);
    if (api.ok && c->comm) api.CommDestroy(c->comm);
    delete c;
    return 0;
}

size_t ofb_slab_workspace_bytes(int B, int Ny, int dtype, const ofb_slab* slab) {
    if (slab == nullptr || dtype < OFB_F32 || dtype > OFB_F64 || B < 1) return 0;
    const size_t es = dtype == OFB_F32 ? 4 : 8;
    const Plan plan = make_plan_slab(*slab, Ny, 0, es, false);
    return dtype == OFB_F32 ? carve<float>(nullptr, B, plan).bytes : carve<double>(nullptr, B, plan).bytes;
}

int ofb_solve_slab(const void* P, const void* b3, void* x3, const double* alpha, const double* beta, double rtol, int max_it,
                   const ofb_solve_options* opt, void* workspace, size_t workspace_bytes, int* iters_out, double* relres_out,
                   int* converged_out, double* recurrence_out, int B, int Nx_local, int Ny, int dtype, const ofb_slab* slab, void* stream) {
    if (int rc = check_shape(B, std::max(Nx_local, 5), Ny, dtype)) return rc;
    if (slab == nullptr) return fail(-24, "bad slab descriptor");
    if (max_it < 0) return fail(-8, "max_it must be >= 0");
    return dtype == OFB_F32
               ? solve_impl<float>(P, b3, x3, alpha, beta, rtol, max_it, opt, workspace, workspace_bytes, iters_out, relres_out,
                                   converged_out, recurrence_out, B, Nx_local, Ny, stream, slab)
               : solve_impl<double>(P, b3, x3, alpha, beta, rtol, max_it, opt, workspace, workspace_bytes, iters_out, relres_out,
                                    converged_out, recurrence_out, B, Nx_local, Ny, stream, slab);
}

int ofb_epilogue_slab(const void* P, const void* C, const void* x3, void* vx, void* vy, void* speed, void* gamma, double* functionals,
                      double* moments, const double* alpha, const double* beta, double scale, int B, int Nx_local, int Ny, int dtype,
                      const ofb_slab* slab, void* stream) {
    if (int rc = check_shape(B, std::max(Nx_local, 5), Ny, dtype)) return rc;
    if (slab == nullptr) return fail(-24, "bad slab descriptor");
    const int lo = slab->rank > 0 ? 1 : 0, hi = slab->rank < slab->world - 1 ? 1 : 0;
    return dtype == OFB_F32 ? epilogue_impl<float>(P, C, x3, vx, vy, speed, gamma, functionals, moments, alpha, beta, scale, B, Nx_local,
                                                   Ny, (cudaStream_t)stream, lo, hi)
                            : epilogue_impl<double>(P, C, x3, vx, vy, speed, gamma, functionals, moments, alpha, beta, scale, B, Nx_local,
                                                    Ny, (cudaStream_t)stream, lo, hi);
}

}  // extern "C"

// =====================================================================================================
// Callers either side of the hot path (legacy_kernels.cuh)
// =====================================================================================================
namespace {

inline int device_slot() {
    int dev = 0;
    cudaGetDevice(&dev);
    return std::min(std::max(dev, 0), MAX_DEVICES - 1);
}
inline int stream_grid(long long n, int threads) {
    const long long blocks = (n + threads - 1) / threads;
    return (int)std::max(1LL, std::min(blocks, (long long)device_sms() * 8));
}

template <typename T>
int liu_shen_impl(const void* P, const void* C, void* vx, void* vy, void* speed, void* tmp_vx, void* tmp_vy, double alpha,
                  int iterations, double scale, int B, int Nx, int Ny, cudaStream_t stream) {
    using Cfg = LsCfg<T>;
    auto kern = k_liu_shen_step<T>;
    static std::atomic<bool> configured[MAX_DEVICES];      // zero-initialised; set once per device
    const int dev = device_slot();
    if (!configured[dev]) {
        OFB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::SMEM));
        configured[dev] = true;
    }
    const long long plane = (long long)Nx * Ny;
    const int vec_ok = (Ny % 4 == 0 && ptr_aligned16(P) && ptr_aligned16(C) && ptr_aligned16(vx) && ptr_aligned16(vy) &&
                        ptr_aligned16(tmp_vx) && ptr_aligned16(tmp_vy)) ? 1 : 0;
    dim3 grid((Ny + Cfg::TW - 1) / Cfg::TW, (Nx + Cfg::TH - 1) / Cfg::TH, B);
    // rows that are 16-byte multiples go through the streamed sweep (bulk copies into a ring, persistent CTAs)
    using S = LsStream<T>;
    auto skern = k_liu_shen_stream<T>;
    const bool streamed = vec_ok && Ny >= 64 && !getenv("OFB_LIU_SHEN_TILED");
    static std::atomic<int> stream_ctas[MAX_DEVICES];
    int sgrid = 0, tiles_x = 0;
    if (streamed) {
        if (stream_ctas[dev] == 0) {
            OFB_CHECK_CUDA(cudaFuncSetAttribute(skern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::SMEM));
            int nb = 0;
            OFB_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, skern, S::THREADS, S::SMEM));
            if (nb < 1) return fail(-9, "streamed Liu-Shen sweep does not fit on an SM");
            stream_ctas[dev] = nb;
        }
        tiles_x = (Ny + S::TW - 1) / S::TW;
        const long long rows = (long long)B * tiles_x * Nx;
        sgrid = (int)std::min<long long>((long long)device_sms() * stream_ctas[dev], std::max<long long>(1, rows / 8));
    }
    T *ax = (T*)vx, *ay = (T*)vy, *bx = (T*)tmp_vx, *by = (T*)tmp_vy;
    for (int it = 0; it < iterations; ++it) {
        if (streamed)
            skern<<<sgrid, S::THREADS, S::SMEM, stream>>>((const T*)P, (const T*)C, ax, ay, bx, by, Nx, Ny, plane, (T)alpha, B, tiles_x);
        else
            kern<<<grid, Cfg::THREADS, Cfg::SMEM, stream>>>((const T*)P, (const T*)C, ax, ay, bx, by, Nx, Ny, plane, (T)alpha, vec_ok);
        OFB_CHECK_LAUNCH();
        std::swap(ax, bx);
        std::swap(ay, by);
    }
    if (ax != (T*)vx) {     // odd number of sweeps: the result sits in the scratch planes
        OFB_CHECK_CUDA(cudaMemcpyAsync(vx, ax, sizeof(T) * plane * B, cudaMemcpyDeviceToDevice, stream));
        OFB_CHECK_CUDA(cudaMemcpyAsync(vy, ay, sizeof(T) * plane * B, cudaMemcpyDeviceToDevice, stream));
    }
    const long long n = plane * B;
    k_liu_shen_finish<T><<<stream_grid(n, 256), 256, 0, stream>>>((T*)vx, (T*)vy, (T*)speed, n, (T)scale);
    OFB_CHECK_LAUNCH();
    return 0;
}

template <typename T>
int movie_max_impl(const void* in, long long n, double* out, cudaStream_t stream) {
    k_fill_double<<<1, 1, 0, stream>>>(out, -INFINITY);
    OFB_CHECK_LAUNCH();
    k_movie_max<T><<<stream_grid(n, 256 * 8), 256, 0, stream>>>((const T*)in, n, out);
    OFB_CHECK_LAUNCH();
    return 0;
}

template <typename T>
int adaptive_threshold_impl(const void* in, const double* movie_max, unsigned char* tmp, unsigned char* out, int window, int idelta,
                            int F, int Nx, int Ny, cudaStream_t stream) {
    const long long plane = (long long)Nx * Ny, n = plane * F, n_rows = (long long)F * Nx;
    unsigned short* hs = reinterpret_cast<unsigned short*>(tmp);          // scratch: 16-bit row sums, then the byte plane
    unsigned char* u8 = tmp + 2 * n;
    auto kern = k_threshold_rows<T>;
    const size_t per_warp = (size_t)(Ny + 1) * sizeof(unsigned int);
    const int warps = (int)std::max<size_t>(1, std::min<size_t>(AT_ROW_WARPS, (96 * 1024) / per_warp));
    const size_t smem = per_warp * warps;
    if (smem > 200 * 1024) return fail(-6, "ofb_adaptive_threshold: rows of %d columns do not fit the prefix-sum buffer", Ny);
    static std::atomic<bool> configured[MAX_DEVICES];      // zero-initialised; set once per device
    const int dev = device_slot();
    if (!configured[dev]) {
        OFB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        configured[dev] = true;
    }
    const int grid = (int)std::min<long long>((n_rows + warps - 1) / warps, (long long)device_sms() * 8);
    kern<<<grid, AT_ROW_WARPS * 32, smem, stream>>>((const T*)in, movie_max, u8, hs, n_rows, Ny, window / 2, warps);
    OFB_CHECK_LAUNCH();
    dim3 cgrid((Ny + 127) / 128, (Nx + AT_STRIP - 1) / AT_STRIP, F);
    k_threshold_cols<<<cgrid, 128, 0, stream>>>(u8, hs, out, Nx, Ny, plane, window, idelta);
    OFB_CHECK_LAUNCH();
    return 0;
}

inline size_t clahe_frame_bytes(int tiles_x, int tiles_y) {
    return (size_t)tiles_x * tiles_y * CLAHE_BINS * (sizeof(int) + sizeof(unsigned short));
}

template <typename T>
int clahe_impl(const void* in, void* out, const ClaheGeom& g, void* workspace, int frames_in_flight, int F, cudaStream_t stream) {
    const long long plane = (long long)g.Nx * g.Ny;
    const size_t tiles = (size_t)g.tiles_x * g.tiles_y;
    int* hist = (int*)workspace;
    unsigned short* lut = (unsigned short*)((char*)workspace + (size_t)frames_in_flight * tiles * CLAHE_BINS * sizeof(int));
    for (int f0 = 0; f0 < F; f0 += frames_in_flight) {
        const int nf = std::min(frames_in_flight, F - f0);
        const T* src = (const T*)in + plane * f0;
        OFB_CHECK_CUDA(cudaMemsetAsync(hist, 0, (size_t)nf * tiles * CLAHE_BINS * sizeof(int), stream));
        k_clahe_hist<T><<<dim3((g.ext_y + 255) / 256, g.ext_x, nf), 256, 0, stream>>>(src, hist, g, plane);
        OFB_CHECK_LAUNCH();
        k_clahe_lut<<<dim3((unsigned)tiles, nf), 1024, 0, stream>>>(hist, lut, g);
        OFB_CHECK_LAUNCH();
        k_clahe_interp<T, T><<<dim3((g.Ny + 255) / 256, g.Nx, nf), 256, 0, stream>>>(src, lut, (T*)out + plane * f0, g, plane);
        OFB_CHECK_LAUNCH();
    }
    return 0;
}

}  // namespace

extern "C" {

int ofb_liu_shen(const void* P, const void* C, void* vx, void* vy, void* speed, void* tmp_vx, void* tmp_vy, double alpha,
                 int iterations, double scale, int B, int Nx, int Ny, int dtype, void* stream) {
    if (B < 1 || Nx < 2 || Ny < 2) return fail(-1, "ofb_liu_shen: bad shape %d x %d x %d (frames must be at least 2x2)", B, Nx, Ny);
    if (dtype != OFB_F32 && dtype != OFB_F64) return fail(-3, "dtype must be OFB_F32 or OFB_F64 (got %d)", dtype);
    if (iterations < 0) return fail(-6, "ofb_liu_shen: iterations must be >= 0 (got %d)", iterations);
    if (!P || !C || !vx || !vy || !speed || !tmp_vx || !tmp_vy) return fail(-6, "ofb_liu_shen: null pointer");
    if (vx == tmp_vx || vy == tmp_vy || vx == vy || tmp_vx == tmp_vy) return fail(-7, "ofb_liu_shen: velocity planes must not alias");
    return dtype == OFB_F32 ? liu_shen_impl<float>(P, C, vx, vy, speed, tmp_vx, tmp_vy, alpha, iterations, scale, B, Nx, Ny, (cudaStream_t)stream)
                            : liu_shen_impl<double>(P, C, vx, vy, speed, tmp_vx, tmp_vy, alpha, iterations, scale, B, Nx, Ny, (cudaStream_t)stream);
}

int ofb_movie_max(const void* in, long long n, int dtype, double* max_out, void* stream) {
    if (n < 1 || !in || !max_out) return fail(-1, "ofb_movie_max: empty input");
    if (dtype != OFB_F32 && dtype != OFB_F64) return fail(-3, "dtype must be OFB_F32 or OFB_F64 (got %d)", dtype);
    return dtype == OFB_F32 ? movie_max_impl<float>(in, n, max_out, (cudaStream_t)stream)
                            : movie_max_impl<double>(in, n, max_out, (cudaStream_t)stream);
}

int ofb_adaptive_threshold(const void* in, const double* movie_max, unsigned char* tmp_u8, unsigned char* out, int window, int idelta,
                           int F, int Nx, int Ny, int dtype, void* stream) {
    if (F < 1 || Nx < 1 || Ny < 1) return fail(-1, "ofb_adaptive_threshold: bad shape %d x %d x %d", F, Nx, Ny);
    if (dtype != OFB_F32 && dtype != OFB_F64) return fail(-3, "dtype must be OFB_F32 or OFB_F64 (got %d)", dtype);
    // cv2.adaptiveThreshold asserts blockSize % 2 == 1 && blockSize > 1; 255 keeps the row sums in 16 bits
    if (window < 3 || window > 255 || window % 2 == 0) return fail(-6, "ofb_adaptive_threshold: window must be odd and in 3..255 (got %d)", window);
    if (!in || !movie_max || !tmp_u8 || !out || tmp_u8 == out) return fail(-7, "ofb_adaptive_threshold: null or aliasing buffers");
    return dtype == OFB_F32 ? adaptive_threshold_impl<float>(in, movie_max, tmp_u8, out, window, idelta, F, Nx, Ny, (cudaStream_t)stream)
                            : adaptive_threshold_impl<double>(in, movie_max, tmp_u8, out, window, idelta, F, Nx, Ny, (cudaStream_t)stream);
}

size_t ofb_clahe_workspace_bytes(int frames_in_flight, int tiles_x, int tiles_y) {
    if (frames_in_flight < 1 || tiles_x < 1 || tiles_y < 1) return 0;
    return (size_t)frames_in_flight * clahe_frame_bytes(tiles_x, tiles_y);
}

int ofb_clahe(const void* in, void* out, double clip_limit, int tiles_x, int tiles_y, void* workspace, size_t workspace_bytes, int F,
              int Nx, int Ny, int dtype, void* stream) {
    if (F < 1 || Nx < 1 || Ny < 1) return fail(-1, "ofb_clahe: bad shape %d x %d x %d", F, Nx, Ny);
    if (dtype != OFB_F32 && dtype != OFB_F64) return fail(-3, "dtype must be OFB_F32 or OFB_F64 (got %d)", dtype);
    if (tiles_x < 1 || tiles_y < 1) return fail(-6, "ofb_clahe: tile grid must be at least 1 x 1 (got %d x %d)", tiles_x, tiles_y);
    if (!in || !out || in == out || !workspace) return fail(-7, "ofb_clahe: null or aliasing buffers");
    ClaheGeom g;
    g.Nx = Nx; g.Ny = Ny; g.tiles_x = tiles_x; g.tiles_y = tiles_y;
    if (Ny % tiles_x == 0 && Nx % tiles_y == 0) {
        g.ext_x = Nx; g.ext_y = Ny;
    } else {        // clahe.cpp: copyMakeBorder(0, tilesY - rows % tilesY, 0, tilesX - cols % tilesX, BORDER_REFLECT_101)
        g.ext_x = Nx + (tiles_y - Nx % tiles_y);
        g.ext_y = Ny + (tiles_x - Ny % tiles_x);
        if (g.ext_x - Nx > Nx - 1 || g.ext_y - Ny > Ny - 1)
            return fail(-6, "ofb_clahe: a %d x %d tile grid needs more reflected border than a %d x %d frame has", tiles_x, tiles_y, Nx, Ny);
    }
    g.tw = g.ext_y / tiles_x;
    g.th = g.ext_x / tiles_y;
    const int area = g.tw * g.th;
    g.lut_scale = (float)(CLAHE_BINS - 1) / area;
    g.clip = 0;
    if (clip_limit > 0.0) g.clip = std::max((int)(clip_limit * area / CLAHE_BINS), 1);
    const size_t per_frame = clahe_frame_bytes(tiles_x, tiles_y);
    const int in_flight = (int)std::min<size_t>(workspace_bytes / per_frame, (size_t)F);
    if (in_flight < 1) return fail(-5, "ofb_clahe: workspace too small: %zu bytes given, %zu per frame needed", workspace_bytes, per_frame);
    return dtype == OFB_F32 ? clahe_impl<float>(in, out, g, workspace, in_flight, F, (cudaStream_t)stream)
                            : clahe_impl<double>(in, out, g, workspace, in_flight, F, (cudaStream_t)stream);
}

int ofb_widen(const void* in_f32, void* out_f64, long long n, void* stream) {
    if (n < 1 || !in_f32 || !out_f64) return fail(-1, "ofb_widen: empty input");
    if (!ptr_aligned16(in_f32) || !ptr_aligned16(out_f64)) return fail(-7, "ofb_widen: buffers must be 16-byte aligned");
    k_widen<<<stream_grid(n, 256 * 8), 256, 0, (cudaStream_t)stream>>>((const float*)in_f32, (double*)out_f64, n);
    OFB_CHECK_LAUNCH();
    return 0;
}

int ofb_subtract_background(const void* movie, const void* blurred, void* out, double background, long long n, int dtype, void* stream) {
    if (n < 1 || !movie || !blurred || !out) return fail(-1, "ofb_subtract_background: empty input");
    if (dtype != OFB_F32 && dtype != OFB_F64) return fail(-3, "dtype must be OFB_F32 or OFB_F64 (got %d)", dtype);
    if (dtype == OFB_F32)
        k_subtract_background<float><<<stream_grid(n, 256 * 4), 256, 0, (cudaStream_t)stream>>>((const float*)movie, (const float*)blurred, (float*)out, n, (float)background);
    else
        k_subtract_background<double><<<stream_grid(n, 256 * 4), 256, 0, (cudaStream_t)stream>>>((const double*)movie, (const double*)blurred, (double*)out, n, background);
    OFB_CHECK_LAUNCH();
    return 0;
}

}  // extern "C"

// =====================================================================================================
// Allocation-free variant (no counterpart in the reference)
// =====================================================================================================
// Callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc) get the
// two-pass Gaussian blur with its intermediate plane in a file-scope device array instead of a caller buffer.
// Every GPU that loads the library holds OFB_BLUR_NOALLOC_MAX_BYTES until the process exits.  All calls share the
// array: it keeps the last call's contents (pass 1 overwrites every element pass 2 reads, so nothing needs clearing)
// and calls that overlap on one GPU race on it.
namespace {

static __device__ __align__(256) unsigned char g_blur_scratch[OFB_BLUR_NOALLOC_MAX_BYTES];

// In host code `g_blur_scratch` is a host-side stand-in, not a device address: resolve the address once per device,
// with that device current, and hand the kernels the resolved pointer.
int blur_scratch_address(void** out) {
    static std::atomic<void*> resolved[MAX_DEVICES];          // zero-initialised
    int dev = 0;
    OFB_CHECK_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MAX_DEVICES) return fail(-11, "device ordinal %d out of range", dev);
    void* p = resolved[dev].load(std::memory_order_acquire);
    if (p == nullptr) {
        OFB_CHECK_CUDA(cudaGetSymbolAddress(&p, g_blur_scratch));
        if (p == nullptr) return fail(-11, "cudaGetSymbolAddress returned a null scratch address");
        resolved[dev].store(p, std::memory_order_release);
    }
    *out = p;
    return 0;
}

template <typename T>
int blur_noalloc_impl(const void* in, void* out, void* scratch, const double* weights, int radius, int F, int Nx, int Ny,
                      cudaStream_t stream) {
    const long long fs = (long long)Nx * Ny;
    dim3 block(128), grid((Ny + 127) / 128, Nx, F);
    k_blur_axis<T, 0><<<grid, block, 0, stream>>>((const T*)in, (T*)scratch, Nx, Ny, fs, weights, radius);
    OFB_CHECK_LAUNCH();
    k_blur_axis<T, 1><<<grid, block, 0, stream>>>((const T*)scratch, (T*)out, Nx, Ny, fs, weights, radius);
    OFB_CHECK_LAUNCH();
    return 0;
}

}  // namespace

extern "C" int ofb_blur_noalloc(const void* in, void* out, const double* weights, int radius, int F, int Nx, int Ny, int dtype,
                                void* stream) {
    if (F < 1 || Nx < 1 || Ny < 1) return fail(-1, "ofb_blur_noalloc: bad shape %d x %d x %d", F, Nx, Ny);
    if (dtype != OFB_F32 && dtype != OFB_F64) return fail(-3, "dtype must be OFB_F32 or OFB_F64 (got %d)", dtype);
    if (radius < 0 || !weights || !in || !out) return fail(-6, "ofb_blur_noalloc: null pointer or negative radius");
    if (in == out) return fail(-7, "ofb_blur_noalloc: in/out must not alias");
    if (Ny > 65535 * 128 || Nx > 65535 || F > 65535) return fail(-1, "ofb_blur_noalloc: shape %d x %d x %d exceeds the launch grid", F, Nx, Ny);
    const unsigned long long need = (unsigned long long)F * Nx * Ny * (dtype == OFB_F32 ? 4ull : 8ull);
    if (need > (unsigned long long)OFB_BLUR_NOALLOC_MAX_BYTES)
        return fail(-12, "ofb_blur_noalloc: %d x %d x %d frames need %llu bytes of scratch, the built-in array holds %llu "
                         "(OFB_BLUR_NOALLOC_MAX_BYTES); use ofb_blur with a caller-provided tmp", F, Nx, Ny, need,
                    (unsigned long long)OFB_BLUR_NOALLOC_MAX_BYTES);
    void* scratch = nullptr;
    if (int rc = blur_scratch_address(&scratch)) return rc;
    return dtype == OFB_F32 ? blur_noalloc_impl<float>(in, out, scratch, weights, radius, F, Nx, Ny, (cudaStream_t)stream)
                            : blur_noalloc_impl<double>(in, out, scratch, weights, radius, F, Nx, Ny, (cudaStream_t)stream);
}

// =====================================================================================================
// Sparse direct solve (direct_solver.cuh): replaces the SuperLU branch, optical_flow.py:1146-1147
// =====================================================================================================
namespace {

#define OFB_CHECK_CUBLAS(expr)                                                                   \
    do {                                                                                         \
        cublasStatus_t s__ = (expr);                                                             \
        if (s__ != CUBLAS_STATUS_SUCCESS) return fail(-3000 - (int)s__, "%s: cuBLAS status %d", #expr, (int)s__); \
    } while (0)
#define OFB_CHECK_CUSOLVER(expr)                                                                 \
    do {                                                                                         \
        cusolverStatus_t s__ = (expr);                                                           \
        if (s__ != CUSOLVER_STATUS_SUCCESS) return fail(-4000 - (int)s__, "%s: cuSOLVER status %d", #expr, (int)s__); \
    } while (0)

// Which levels take the batched cuBLAS LU instead of one cuSOLVER getrf per node: fronts of up to 128 owned variables, and up to 192 where
// the level has at least 128 of them.  Measured (tools/direct_threshold.py; the fronts in question have 189 unknowns): a 1024^2 pair - levels
// of 128 and 256 such nodes - 0.224 -> 0.205 s with them batched; a 512^2 pair - 32 and 64 nodes - 0.055 -> 0.066 s, so those stay with
// cuSOLVER; batching the 381-unknown fronts as well: 0.229 s, the 765s: 0.390 s.  OFB_ND_BATCHED_MAX = n makes it "up to n" for A/B runs.
bool nd_use_batched(int no3, int n_nodes) {
    static const int forced = [] { const char* e = getenv("OFB_ND_BATCHED_MAX"); return e ? atoi(e) : 0; }();
    if (forced > 0) return no3 <= forced;
    return no3 <= 128 || (no3 <= 192 && n_nodes >= 128);
}

constexpr int ND_STREAMS = 8;             // side streams for the per-node LU of the mid-size fronts (one getrf cannot fill the GPU)
struct DenseHandles {
    cublasHandle_t blas = nullptr;
    cusolverDnHandle_t solver = nullptr;                   // on the caller's stream
    cusolverDnHandle_t side_solver[ND_STREAMS] = {};       // one per side stream
    cudaStream_t side[ND_STREAMS] = {};
    cudaEvent_t fork = nullptr, join[ND_STREAMS] = {};
    cudaStream_t graph_stream = nullptr;                   // the captured graph of a frame shape runs here (the caller's stream may be the
    cudaEvent_t caller_ready = nullptr;                    // legacy default stream, which cannot be captured)
    std::mutex busy;                                       // the side streams and events serve one direct solve per device at a time
};
int dense_handles(DenseHandles** out) {
    static DenseHandles handles[MAX_DEVICES];
    static std::mutex mu;
    int dev = 0;
    OFB_CHECK_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MAX_DEVICES) return fail(-11, "device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lock(mu);
    DenseHandles& h = handles[dev];
    if (h.blas == nullptr) OFB_CHECK_CUBLAS(cublasCreate(&h.blas));
    if (h.solver == nullptr) OFB_CHECK_CUSOLVER(cusolverDnCreate(&h.solver));
    if (h.fork == nullptr) {
        OFB_CHECK_CUDA(cudaStreamCreateWithFlags(&h.graph_stream, cudaStreamNonBlocking));
        OFB_CHECK_CUDA(cudaEventCreateWithFlags(&h.caller_ready, cudaEventDisableTiming));
        OFB_CHECK_CUDA(cudaEventCreateWithFlags(&h.fork, cudaEventDisableTiming));
        for (int k = 0; k < ND_STREAMS; ++k) {
            OFB_CHECK_CUDA(cudaStreamCreateWithFlags(&h.side[k], cudaStreamNonBlocking));
            OFB_CHECK_CUDA(cudaEventCreateWithFlags(&h.join[k], cudaEventDisableTiming));
            OFB_CHECK_CUSOLVER(cusolverDnCreate(&h.side_solver[k]));
            OFB_CHECK_CUSOLVER(cusolverDnSetStream(h.side_solver[k], h.side[k]));
        }
    }
    *out = &h;
    return 0;
}
// side streams wait for what the caller's stream has queued so far / the caller's stream waits for the side streams
int nd_fork(DenseHandles* H, cudaStream_t stream, int n_used) {
    OFB_CHECK_CUDA(cudaEventRecord(H->fork, stream));
    for (int k = 0; k < n_used; ++k) OFB_CHECK_CUDA(cudaStreamWaitEvent(H->side[k], H->fork, 0));
    return 0;
}
int nd_join(DenseHandles* H, cudaStream_t stream, int n_used) {
    for (int k = 0; k < n_used; ++k) {
        OFB_CHECK_CUDA(cudaEventRecord(H->join[k], H->side[k]));
        OFB_CHECK_CUDA(cudaStreamWaitEvent(stream, H->join[k], 0));
    }
    return 0;
}

struct NdWorkspace {
    NdLevel lv[64];
    double *w, *z, *u, *getrf_work;
    double **Fptr, **F12ptr, **zptr;
    PairState* st;
    int* active;            // per-pair flags k_state_init wants (unused here)
    int* scratch_info;      // devInfo of the triangular solves (getrf's verdict stays in NdLevel::info)
    int* flag;              // 0, or 1 + the level of a front whose LU hit a zero pivot
    void *img_stage, *b_stage, *x_stage;    // fixed slots the captured graph of a frame shape works on
    double* ab_stage;       // the pair's alpha, beta
    int getrf_lwork;
    size_t bytes;
};

// `levels` describe the plan (device index arrays); carve the fronts and the scratch vectors out of `base`.
int nd_carve(const ofb_nd_level* levels, int n_levels, int B, int Nx, int Ny, void* base, NdWorkspace* W, cusolverDnHandle_t solver,
             size_t elem = 8) {
    if (n_levels < 1 || n_levels > 64) return fail(-50, "direct solve: bad number of levels %d", n_levels);
    size_t off = 0;
    auto take = [&](size_t nbytes) { void* p = base ? (char*)base + off : nullptr; off = align_up(off + nbytes, 256); return p; };
    size_t max_nodes = 0, max_z = 0, max_u = 0;
    int lwork = 0;
    for (int l = 0; l < n_levels; ++l) {
        NdLevel& L = W->lv[l];
        L.n_nodes = levels[l].n_nodes; L.n_own = levels[l].n_own; L.n_border = levels[l].n_border;
        if (L.n_nodes < 1 || L.n_own < 1 || L.n_border < 1) return fail(-50, "direct solve: bad level %d", l);
        L.own = levels[l].own; L.border = levels[l].border; L.parent_slot = levels[l].parent_slot;
        const size_t nf = (size_t)L.nf();
        L.F = (double*)take((size_t)L.n_nodes * nf * nf * sizeof(double));
        L.piv = (int*)take((size_t)L.n_nodes * L.no3() * sizeof(int));
        L.info = (int*)take((size_t)L.n_nodes * sizeof(int));
        max_nodes = std::max(max_nodes, (size_t)L.n_nodes);
        max_z = std::max(max_z, (size_t)L.n_nodes * L.no3());
        max_u = std::max(max_u, (size_t)L.n_nodes * L.nb3());
        if (!nd_use_batched(L.no3(), L.n_nodes) && solver != nullptr) {
            int lw = 0;
            OFB_CHECK_CUSOLVER(cusolverDnDgetrf_bufferSize(solver, L.no3(), L.no3(), nullptr, (int)nf, &lw));
            lwork = std::max(lwork, lw);
        }
    }
    const size_t M = (size_t)(Nx - 2) * (Ny - 2);
    W->w = (double*)take(3 * M * sizeof(double));
    W->z = (double*)take(max_z * sizeof(double));
    W->u = (double*)take(max_u * sizeof(double));
    W->Fptr = (double**)take(max_nodes * sizeof(double*));
    W->F12ptr = (double**)take(max_nodes * sizeof(double*));
    W->zptr = (double**)take(max_nodes * sizeof(double*));
    W->getrf_lwork = lwork;
    W->getrf_work = (double*)take((size_t)std::max(lwork, 1) * ND_STREAMS * sizeof(double));       // one slice per side stream
    W->scratch_info = (int*)take(sizeof(int) * ND_STREAMS);
    W->flag = (int*)take(sizeof(int));
    W->img_stage = take((size_t)Nx * Ny * elem);
    W->b_stage = take((size_t)3 * Nx * Ny * elem);
    W->x_stage = take((size_t)3 * Nx * Ny * elem);
    W->ab_stage = (double*)take(2 * sizeof(double));
    // everything above is independent of B (a captured graph holds those addresses); the per-pair arrays come last
    W->st = (PairState*)take(sizeof(PairState) * (size_t)B);
    W->active = (int*)take(sizeof(int) * (size_t)B);
    W->bytes = off;
    return 0;
}

// Everything one pair takes, queued on `stream` (and the side streams): numeric factorisation leaves to root, the two triangular
// sweeps, the solution into x with its ghost ring.  No host synchronisation inside: it is what the CUDA graph of a frame shape captures.
template <typename T>
int nd_enqueue_pair(DenseHandles* H, NdWorkspace& W, int n_levels, const Geom& g, int Nx, int Ny, const T* img, const T* b, T* x,
                    const double* alpha, const double* beta, int pair, cudaStream_t stream) {
    const long long M = (long long)(Nx - 2) * (Ny - 2);
    const double one = 1.0, minus_one = -1.0;
    OFB_CHECK_CUDA(cudaMemsetAsync(W.flag, 0, sizeof(int), stream));
    // ---- numeric factorisation, leaves to root
    for (int l = n_levels - 1; l >= 0; --l) {
        const NdLevel& L = W.lv[l];
        const int nf = L.nf(), no3 = L.no3(), nb3 = L.nb3();
        const long long stride = (long long)nf * nf;
        OFB_CHECK_CUDA(cudaMemsetAsync(L.F, 0, sizeof(double) * stride * L.n_nodes, stream));
        k_nd_assemble<T><<<dim3(L.n_nodes, no3, (nf + 127) / 128), 128, 0, stream>>>(L, g, img, alpha, beta, pair);
        OFB_CHECK_LAUNCH();
        if (l + 1 < n_levels) {
            const NdLevel& C = W.lv[l + 1];
            for (int parity = 0; parity < 2; ++parity) {
                k_nd_extend_add<<<dim3(C.n_nodes / 2, C.nb3(), (C.nb3() + 127) / 128), 128, 0, stream>>>(C, L, parity);
                OFB_CHECK_LAUNCH();
            }
        }
        if (nd_use_batched(no3, L.n_nodes)) {
            k_nd_pointers<<<(L.n_nodes + 127) / 128, 128, 0, stream>>>(L, W.z, W.Fptr, W.F12ptr, W.zptr);
            OFB_CHECK_LAUNCH();
            OFB_CHECK_CUBLAS(cublasDgetrfBatched(H->blas, no3, W.Fptr, nf, L.piv, L.info, L.n_nodes));
            int info = 0;
            OFB_CHECK_CUBLAS(cublasDgetrsBatched(H->blas, CUBLAS_OP_N, no3, nb3, (const double* const*)W.Fptr, nf, L.piv, W.F12ptr, nf,
                                                 &info, L.n_nodes));
            if (info != 0) return fail(-51, "direct solve: cublasDgetrsBatched argument %d invalid", -info);
        } else {
            // the few hundred mid-size fronts: one getrf + getrs per node, the nodes of the level dealt over side streams
            const int used = std::min(ND_STREAMS, L.n_nodes);
            if (int rc = nd_fork(H, stream, used)) return rc;
            for (int n = 0; n < L.n_nodes; ++n) {
                double* F = L.F + stride * n;
                const int k = n % used;
                OFB_CHECK_CUSOLVER(cusolverDnDgetrf(H->side_solver[k], no3, no3, F, nf, W.getrf_work + (size_t)k * W.getrf_lwork,
                                                    L.piv + (long long)n * no3, L.info + n));
                OFB_CHECK_CUSOLVER(cusolverDnDgetrs(H->side_solver[k], CUBLAS_OP_N, no3, nb3, F, nf, L.piv + (long long)n * no3,
                                                    F + (long long)no3 * nf, nf, W.scratch_info + k));
            }
            if (int rc = nd_join(H, stream, used)) return rc;
        }
        k_nd_check_info<<<(L.n_nodes + 255) / 256, 256, 0, stream>>>(L.info, L.n_nodes, l, W.flag);        // a zero pivot: remember where
        OFB_CHECK_LAUNCH();
        // S = F22 - F21 X, in place in the F22 block of every front of the level
        OFB_CHECK_CUBLAS(cublasDgemmStridedBatched(H->blas, CUBLAS_OP_N, CUBLAS_OP_N, nb3, nb3, no3, &minus_one, L.F + no3, nf, stride,
                                                   L.F + (long long)no3 * nf, nf, stride, &one, L.F + (long long)no3 * nf + no3, nf,
                                                   stride, L.n_nodes));
    }
    // ---- forward sweep, leaves to root: z = F11^-1 w[own]; w[own] = z; w[border] -= F21 z
    k_nd_load_rhs<T><<<dim3((Ny - 2 + 127) / 128, Nx - 2, 3), 128, 0, stream>>>(g, b, W.w);
    OFB_CHECK_LAUNCH();
    for (int l = n_levels - 1; l >= 0; --l) {
        const NdLevel& L = W.lv[l];
        const int nf = L.nf(), no3 = L.no3(), nb3 = L.nb3();
        const long long stride = (long long)nf * nf;
        k_nd_gather<<<dim3(L.n_nodes, (no3 + 127) / 128), 128, 0, stream>>>(L, W.w, M, W.z, 0);
        OFB_CHECK_LAUNCH();
        if (nd_use_batched(no3, L.n_nodes)) {
            k_nd_pointers<<<(L.n_nodes + 127) / 128, 128, 0, stream>>>(L, W.z, W.Fptr, W.F12ptr, W.zptr);
            OFB_CHECK_LAUNCH();
            int info = 0;
            OFB_CHECK_CUBLAS(cublasDgetrsBatched(H->blas, CUBLAS_OP_N, no3, 1, (const double* const*)W.Fptr, nf, L.piv, W.zptr, no3, &info,
                                                 L.n_nodes));
        } else {
            const int used = std::min(ND_STREAMS, L.n_nodes);
            if (int rc = nd_fork(H, stream, used)) return rc;
            for (int n = 0; n < L.n_nodes; ++n)
                OFB_CHECK_CUSOLVER(cusolverDnDgetrs(H->side_solver[n % used], CUBLAS_OP_N, no3, 1, L.F + stride * n, nf,
                                                    L.piv + (long long)n * no3, W.z + (long long)n * no3, no3, W.scratch_info + n % used));
            if (int rc = nd_join(H, stream, used)) return rc;
        }
        k_nd_store_own<<<dim3(L.n_nodes, (no3 + 127) / 128), 128, 0, stream>>>(L, W.w, M, W.z);
        OFB_CHECK_LAUNCH();
        const double zero = 0.0;
        OFB_CHECK_CUBLAS(cublasDgemmStridedBatched(H->blas, CUBLAS_OP_N, CUBLAS_OP_N, nb3, 1, no3, &one, L.F + no3, nf, stride, W.z, no3, no3,
                                                   &zero, W.u, nb3, nb3, L.n_nodes));
        k_nd_scatter_border<<<dim3(L.n_nodes, (nb3 + 127) / 128), 128, 0, stream>>>(L, W.w, M, W.u);
        OFB_CHECK_LAUNCH();
    }
    // ---- backward sweep, root to leaves: x[own] = z - X x[border]
    for (int l = 0; l < n_levels; ++l) {
        const NdLevel& L = W.lv[l];
        const int nf = L.nf(), no3 = L.no3(), nb3 = L.nb3();
        const long long stride = (long long)nf * nf;
        k_nd_gather<<<dim3(L.n_nodes, (no3 + 127) / 128), 128, 0, stream>>>(L, W.w, M, W.z, 0);
        OFB_CHECK_LAUNCH();
        k_nd_gather<<<dim3(L.n_nodes, (nb3 + 127) / 128), 128, 0, stream>>>(L, W.w, M, W.u, 1);
        OFB_CHECK_LAUNCH();
        OFB_CHECK_CUBLAS(cublasDgemmStridedBatched(H->blas, CUBLAS_OP_N, CUBLAS_OP_N, no3, 1, nb3, &minus_one, L.F + (long long)no3 * nf, nf,
                                                   stride, W.u, nb3, nb3, &one, W.z, no3, no3, L.n_nodes));
        k_nd_store_own<<<dim3(L.n_nodes, (no3 + 127) / 128), 128, 0, stream>>>(L, W.w, M, W.z);
        OFB_CHECK_LAUNCH();
    }
    k_nd_store_solution<T><<<dim3((Ny + 127) / 128, Nx, 3), 128, 0, stream>>>(g, W.w, x);
    OFB_CHECK_LAUNCH();
    const int ring = 2 * Nx + 2 * Ny;
    k_fill_ghosts<T><<<dim3((ring + 127) / 128, 1), 128, 0, stream>>>(g, x);
    OFB_CHECK_LAUNCH();
    return 0;
}

// One graph per device: the launch sequence of a pair depends on the frame shape (the plan), the workspace and the element type only, so
// it is captured once over staging slots in the workspace and replayed per pair - a solve is ~16 400 launches, most of them from inside
// cuSOLVER / cuBLAS calls, and the host could not issue them faster than 13 us each (profiles/r02_direct_solve_launches.txt).
struct NdGraph {
    cudaGraphExec_t exec = nullptr;
    void* workspace = nullptr;
    const void* plan = nullptr;
    int Nx = 0, Ny = 0, n_levels = 0, elem = 0;
    bool unsupported = false;            // a library call refused to be captured: stay on the eager path
};

std::atomic<int> g_direct_graph_state{0};       // 0: no direct solve yet, 1: the last one replayed a captured graph, -1: capture refused, eager

template <typename T>
int direct_impl(const void* P_, const void* b3, void* x3, const double* alpha, const double* beta, const ofb_nd_level* levels,
                int n_levels, void* workspace, size_t workspace_bytes, double rtol, double* relres_out, int* converged_out, int B, int Nx,
                int Ny, cudaStream_t stream) {
    DenseHandles* H = nullptr;
    if (int rc = dense_handles(&H)) return rc;
    std::lock_guard<std::mutex> one_at_a_time(H->busy);
    OFB_CHECK_CUBLAS(cublasSetStream(H->blas, stream));
    OFB_CHECK_CUSOLVER(cusolverDnSetStream(H->solver, stream));
    NdWorkspace W;
    if (int rc = nd_carve(levels, n_levels, B, Nx, Ny, workspace, &W, H->solver, sizeof(T))) return rc;
    if (workspace == nullptr || workspace_bytes < W.bytes)
        return fail(-4, "direct solve: workspace too small: need %zu bytes, got %zu", W.bytes, workspace_bytes);
    LevelPlan lp{Nx, Ny, Ny, (long long)Nx * Ny, 2, 2, 0.0, 0.0};
    const Geom g = make_geom<T>(lp, false);
    static NdGraph graphs[MAX_DEVICES];
    int dev = 0;
    OFB_CHECK_CUDA(cudaGetDevice(&dev));
    NdGraph& G = graphs[std::min(std::max(dev, 0), MAX_DEVICES - 1)];
    // OFB_DIRECT_GRAPH=1: replay a captured graph per pair instead of issuing the ~16 400 launches.  Measured: 0.21 s per 1024^2 pair
    // either way - the solve is bound by the chains of dependent LU panel kernels of the large fronts on the device, not by the host -
    // so it only frees the host thread; off by default.
    const char* graph_env = getenv("OFB_DIRECT_GRAPH");
    const bool want_graph = !G.unsupported && graph_env != nullptr && graph_env[0] == '1';
    T* img_stage = (T*)W.img_stage;
    T* b_stage = (T*)W.b_stage;
    T* x_stage = (T*)W.x_stage;
    for (int pair = 0; pair < B; ++pair) {
        const T* img = (const T*)P_ + (long long)pair * g.img_pair;
        const T* b = (const T*)b3 + (long long)pair * g.pair;
        T* x = (T*)x3 + (long long)pair * g.pair;
        bool done = false;
        cudaStream_t flag_stream = stream;
        if (want_graph && !G.unsupported) {
            cudaStream_t gs = H->graph_stream;
            const bool cached = G.exec != nullptr && G.workspace == workspace && G.plan == (const void*)levels[0].own && G.Nx == Nx && G.Ny == Ny &&
                                G.n_levels == n_levels && G.elem == (int)sizeof(T);
            if (!cached) {
                if (G.exec != nullptr) { cudaGraphExecDestroy(G.exec); G.exec = nullptr; }
                cudaGraph_t graph = nullptr;
                OFB_CHECK_CUDA(cudaStreamSynchronize(gs));
                OFB_CHECK_CUBLAS(cublasSetStream(H->blas, gs));
                OFB_CHECK_CUSOLVER(cusolverDnSetStream(H->solver, gs));
                int rc = 0;
                cudaError_t ce = cudaStreamBeginCapture(gs, cudaStreamCaptureModeThreadLocal);
                if (ce == cudaSuccess) {
                    rc = nd_enqueue_pair<T>(H, W, n_levels, g, Nx, Ny, img_stage, b_stage, x_stage, W.ab_stage, W.ab_stage + 1, 0, gs);
                    ce = cudaStreamEndCapture(gs, &graph);
                }
                if (rc != 0 || ce != cudaSuccess || graph == nullptr || cudaGraphInstantiate(&G.exec, graph, 0) != cudaSuccess) {
                    cudaGetLastError();
                    G.exec = nullptr;
                    G.unsupported = true;           // a call refused to be captured: eager from now on
                } else {
                    G.workspace = workspace; G.plan = (const void*)levels[0].own; G.Nx = Nx; G.Ny = Ny; G.n_levels = n_levels; G.elem = (int)sizeof(T);
                }
                if (graph != nullptr) cudaGraphDestroy(graph);
                OFB_CHECK_CUBLAS(cublasSetStream(H->blas, stream));
                OFB_CHECK_CUSOLVER(cusolverDnSetStream(H->solver, stream));
            }
            if (G.exec != nullptr) {
                OFB_CHECK_CUDA(cudaEventRecord(H->caller_ready, stream));           // inputs queued on the caller's stream are ready
                OFB_CHECK_CUDA(cudaStreamWaitEvent(gs, H->caller_ready, 0));
                OFB_CHECK_CUDA(cudaMemcpyAsync(img_stage, img, sizeof(T) * g.plane, cudaMemcpyDeviceToDevice, gs));
                OFB_CHECK_CUDA(cudaMemcpyAsync(b_stage, b, sizeof(T) * g.pair, cudaMemcpyDeviceToDevice, gs));
                OFB_CHECK_CUDA(cudaMemcpyAsync(W.ab_stage, alpha + pair, sizeof(double), cudaMemcpyDeviceToDevice, gs));
                OFB_CHECK_CUDA(cudaMemcpyAsync(W.ab_stage + 1, beta + pair, sizeof(double), cudaMemcpyDeviceToDevice, gs));
                OFB_CHECK_CUDA(cudaGraphLaunch(G.exec, gs));
                g_launches.fetch_add(1);
                OFB_CHECK_CUDA(cudaMemcpyAsync(x, x_stage, sizeof(T) * g.pair, cudaMemcpyDeviceToDevice, gs));
                flag_stream = gs;                   // the host waits for this stream below: what the caller queues next comes after it
                done = true;
            }
        }
        g_direct_graph_state.store(done ? 1 : (G.unsupported ? -1 : 0));
        if (!done) {
            if (int rc = nd_enqueue_pair<T>(H, W, n_levels, g, Nx, Ny, img, b, x, alpha, beta, pair, stream)) return rc;
        }
        // a zero pivot anywhere makes the factorisation useless: report it instead of returning garbage
        int flag = 0;
        OFB_CHECK_CUDA(cudaMemcpyAsync(&flag, W.flag, sizeof(int), cudaMemcpyDeviceToHost, flag_stream));
        OFB_CHECK_CUDA(cudaStreamSynchronize(flag_stream));
        if (flag != 0) return fail(-52, "direct solve: singular front (pair %d, level %d)", pair, flag - 1);
    }
    // the reference's independent residual (:1150-1151) of what is returned, in fp64
    const long long nvec = g.pair;
    const int sb = (B + 127) / 128;
    const dim3 blas_grid((unsigned)std::min<long long>((nvec + 1023) / 1024, 148 * 8), B);
    k_state_init<<<sb, 128, 0, stream>>>(W.st, W.active, B);
    OFB_CHECK_LAUNCH();
    k_norm2<T><<<blas_grid, 256, 0, stream>>>(nvec, nvec, (const T*)b3, W.st);
    OFB_CHECK_LAUNCH();
    k_state_bnorm<<<sb, 128, 0, stream>>>(W.st, B);
    OFB_CHECK_LAUNCH();
    k_true_residual<T><<<dim3((Ny + 127) / 128, std::min(std::max(Nx - 2, 1), 64), B), 128, 0, stream>>>(g, (const T*)P_, (const T*)x3,
                                                                                                     (const T*)b3, alpha, beta, W.st);
    OFB_CHECK_LAUNCH();
    k_nd_relres<<<sb, 128, 0, stream>>>(W.st, relres_out, converged_out, B, rtol,
                                        sizeof(T) == 4 ? 5.9604644775390625e-08 : 1.1102230246251565e-16);
    OFB_CHECK_LAUNCH();
    return 0;
}

}  // namespace

extern "C" {

int ofb_direct_graph_state(void) { return g_direct_graph_state.load(); }

size_t ofb_direct_workspace_bytes(const ofb_nd_level* levels, int n_levels, int B, int Nx, int Ny) {
    if (levels == nullptr || B < 1 || Nx < 5 || Ny < 5) return 0;
    DenseHandles* H = nullptr;
    if (dense_handles(&H)) return 0;
    NdWorkspace W;
    if (nd_carve(levels, n_levels, B, Nx, Ny, nullptr, &W, H->solver)) return 0;
    return W.bytes;
}

int ofb_direct_solve(const void* P, const void* b3, void* x3, const double* alpha, const double* beta, const ofb_nd_level* levels,
                     int n_levels, void* workspace, size_t workspace_bytes, double rtol, double* relres_out, int* converged_out, int B,
                     int Nx, int Ny, int dtype, void* stream) {
    if (int rc = check_shape(B, Nx, Ny, dtype)) return rc;
    if (!P || !b3 || !x3 || !alpha || !beta || !levels) return fail(-6, "ofb_direct_solve: null pointer");
    return dtype == OFB_F32 ? direct_impl<float>(P, b3, x3, alpha, beta, levels, n_levels, workspace, workspace_bytes, rtol, relres_out,
                                                 converged_out, B, Nx, Ny, (cudaStream_t)stream)
                            : direct_impl<double>(P, b3, x3, alpha, beta, levels, n_levels, workspace, workspace_bytes, rtol, relres_out,
                                                  converged_out, B, Nx, Ny, (cudaStream_t)stream);
}

}  // extern "C"
