// C ABI of the B200 cost-propagation core (declared in include/tasi_gpu.h).
//
// Host side of the solver: owns the device arrays of one planner (padded cost grid, padded value field, tile
// states, work queue), mirrors the flag logic of ReplannerBase::step (ProjectToolkit/include/ReplannerBase.h:43-75)
// and launches the kernels of tasi_solver.cuh. No CPU fallback: every entry point needs the CUDA device.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include <cuda.h>  // CUtensorMap (types only: the driver entry point is looked up through the runtime, no -lcuda)

#include "../../include/tasi_gpu.h"
#include "tasi_solver.cuh"
#include "tasi_extract.cuh"
#include "tasi_parents.cuh"

using namespace tasi;

#ifndef TASI_COMPACT_DEFAULT
#define TASI_COMPACT_DEFAULT 0
#endif

struct tasi_shard;
static thread_local std::string g_create_err = "null planner";

struct tasi_planner {
    tasi_shard *shard = nullptr;  // set on the handle of a sharded planner (tasi_shard.cuh): every entry point forwards
    int band_rt = 0;              // > 0: this grid is a stack of row bands of band_rt tile rows (a member of a shard)
    int shard_world = 1;          // ... of a grid spread over this many GPUs
    bool fresh_plan = false;      // the relax launch in flight is an initial plan (tasi_step), not a repair
    const FusedTopo *ft = nullptr;  // fused multi-GPU mode (device pointer; tasi_shard.cuh), nullptr = off
    int *inbox = nullptr;
    int inbox_cap = 0;
    int kind = TASI_FD, level = 0, device = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, evs0 = nullptr, evs1 = nullptr;
    int num_sms = 0;

    // the reference's flags (ReplannerBase.h:144-152)
    bool initialize_graph = true, initialize_search = true, goal_set = false, new_goal = false, new_start = false;
    int thr = 254;  // Graph.h:33
    float heuristic_multiplier = 1.0f;
    float start_x = 0, start_y = 0, goal_x = 0, goal_y = 0;
    bool have_goal_elem = false;
    int64_t goal_ex = -1, goal_ey = -1;  // goal element, global coordinates
    std::vector<int2> xgoals;            // further goals (tasi_set_goals), band-local element coordinates
    int2 *xgoals_dev = nullptr;
    size_t xgoals_cap = 0;
    uint64_t num_nodes_updated = 0, num_nodes_expanded = 0;

    // geometry
    int W = 0, rows = 0;  // stored cell columns / rows (band: includes the extra rows of open sides)
    int EX = 0, EY = 0, TX = 0, TY = 0;
    int64_t global_row0 = 0;
    bool open_top = false, open_bottom = false;

    // device memory
    float *g_base = nullptr;
    uint8_t *c_base = nullptr;
    int64_t gpitch = 0, cpitch = 0, g_rows = 0;
    int *state = nullptr, *queue = nullptr;
    uint8_t *seed = nullptr;
    int *key = nullptr, *far = nullptr;
    int far_cap = 0;
    int *rkey = nullptr; // upwind deferral (TASI_DEFER): claim keys of the running tiles
    float defer_tiles = -1.0f;  // margin of the deferral in tile widths of mean cost (< 0: off)
    int *bq = nullptr;   // rings of the bucketed near queue (TASI_BUCKETS), nb x bqsize slots
    int nb = 30;
    unsigned bqsize = 0;
    unsigned long long *cost_sum = nullptr;  // device: {sum of free-cell costs, number of free cells}
    bool cost_stats_valid = false;
    float mean_cost = 1.0f;
    float delta_factor = 2.0f;  // width of the value band per threshold advance, in tiles of mean cost
    bool delta_auto = true;     // ... widened on small grids (band_width()), unless TASI_DELTA_FACTOR fixes it
    DevCtl *ctl = nullptr;
    TmaMaps tm{};            // the two CUtensorMap of the TMA-staged solver instances (TASI_TMA=1), else zero
    float *stage = nullptr;  // 2 x EY receive rows (band mode)
    uint8_t *patch_buf = nullptr, *chg_buf = nullptr;
    size_t patch_cap = 0;
    uint8_t *walk_buf = nullptr, *walk_host = nullptr;  // queries and results of tasi_extract_paths (device, pinned)
    size_t walk_cap = 0;
    unsigned qsize = 0;
    int n_tiles_alloc = 0;

    // pending replan seeds
    bool patch_pending = false;
    uint64_t last_seed_count = 0;

    int solver_blocks = 0, max_blocks = 0;
    int strip = 4;  // elements per thread of a tile row in the solver instance this grid uses (alloc_grid)
    // focused planning (tasi_set_focus): plans stop once the start is settled, the rest stays parked
    bool focus = false;
    float focus_slack = 4.0f * 254.0f;
    bool pile_open = false;  // a focused plan left parked tiles: the field is final only below settled_below
    // band mode: the parked tiles and the threshold survive from one launch to the next
    bool band_started = false;
    int sv_far_sel = 0, sv_far_count[2] = {0, 0}, sv_thr_bits = 0, sv_far_min_bits = INF_BITS;
    double timeout_s = 60.0;
    tasi_stats stats{};
    std::string err;
};

namespace {

#define CU_TRY(p, call)                                                                              \
    do {                                                                                             \
        cudaError_t e__ = (call);                                                                    \
        if (e__ != cudaSuccess) {                                                                    \
            (p)->err = std::string(#call) + ": " + cudaGetErrorString(e__);                          \
            return TASI_ERR_CUDA;                                                                    \
        }                                                                                            \
    } while (0)

inline int fail(tasi_planner *p, int code, const char *what) {
    p->err = what;
    return code;
}

inline bool node_based(const tasi_planner *p) { return p->kind != TASI_DFM; }

// Value width of one threshold advance. A wider band releases more tiles at once (more CTAs busy) and re-runs
// more of them; measured (scripts/sweep_delta.sh, all three planners): the best factor is 2 from 16384^2 up, 3 at
// 8192^2 / 4096^2, 6-8 at 2048^2 / 1024^2, within a 15 % spread — small grids have too short a front to feed the
// CTAs otherwise. Row bands keep the plain factor (their threshold is shared across ranks), and so do focused plans
// (the band is the granularity at which they can stop).
// rings of the bucketed near queue a relax launch of this planner uses (0 = the FIFO ring): unsharded grids and the fused
// launches of a sharded grid; the iteration-based exchange and row bands keep the FIFO ring
inline int active_rings(const tasi_planner *p) {
    static const bool fused_buckets = [] { const char *e = getenv("TASI_FUSED_BUCKETS"); return !e || atoi(e) != 0; }();
    return (p->bq && !p->open_top && !p->open_bottom && p->delta_factor > 0.0f &&
            (p->band_rt == 0 || (p->ft && fused_buckets))) ? p->nb : 0;
}

inline float band_width(const tasi_planner *p) {
    float f = p->delta_factor;
    if (p->delta_auto && f > 0.0f && !p->focus && !p->open_top && !p->open_bottom) {
        // a member of a sharded grid holds 1/world of the tiles and sees 1/world of the front of the WHOLE grid (its
        // length goes with the square root of world x local tiles): the single-GPU grid with the same front per GPU has
        // local tiles / world tiles (TASI_SHARD_DELTA_RULE=0: the local tile count, as before)
        static const bool shard_rule = [] { const char *e = getenv("TASI_SHARD_DELTA_RULE"); return !e || atoi(e) != 0; }();
        double nt = std::max(1.0, (double) p->TX * (double) p->TY);
        if (shard_rule && p->shard_world > 1) nt = std::max(1.0, nt / (double) p->shard_world);
        f = (float) std::min(6.0, std::max(2.0, 2.0 + 0.5 * std::log2(262144.0 / nt)));
        // bucketed near queue: the order inside the band comes from the rings, so the band can be wide and the phase
        // barriers few (measured, 30 rings: best factor 32 up to 4096^2, 24 at 8192^2, 8-16 at 16384^2, 8-12 beyond)
        // (initial plans; a repair works on a small active region: factor 4 balances the three patch kinds of cfg3,
        // profiles/r02_cfg3_repair_factor.txt)
        if (active_rings(p) > 0 && (p->fresh_plan || p->band_rt > 0)) f = (float) std::min(32.0, std::max(8.0, 12.0 + 5.0 * std::log2(262144.0 / nt)));
        static const float repair_f = [] { const char *e = getenv("TASI_REPAIR_FACTOR"); return e ? (float) atof(e) : 0.0f; }();
        if (active_rings(p) > 0 && !p->fresh_plan && p->band_rt == 0) f = repair_f > 0.0f ? repair_f : std::max(f, 4.0f);  // cfg3 sweep: 3 / 4 / 6 / 8 / 12
    }
    return f * (float) TS * p->mean_cost;
}

inline float *g_origin(const tasi_planner *p) { return p->g_base + p->gpitch + GMARGIN; }
inline uint8_t *c_origin(const tasi_planner *p) { return p->c_base + p->cpitch + GMARGIN; }

GridDesc make_desc(const tasi_planner *p) {
    GridDesc d;
    d.g = g_origin(p);
    d.gpitch = p->gpitch;
    d.cost = c_origin(p);
    d.cpitch = p->cpitch;
    d.EX = p->EX; d.EY = p->EY; d.TX = p->TX; d.TY = p->TY;
    d.L = p->rows; d.W = p->W;
    d.thr = p->thr;
    int64_t px = p->have_goal_elem ? p->goal_ex - p->global_row0 : -1;
    int64_t py = p->have_goal_elem ? p->goal_ey : -1;
    if (px < 0 || px >= p->EX || py < 0 || py >= p->EY) { px = -1000000; py = -1000000; }
    d.pin_x = (int) px; d.pin_y = (int) py;
    d.xgoals = p->xgoals_dev; d.n_xgoals = (int) p->xgoals.size();
    d.state = p->state; d.queue = p->queue; d.qmask = p->qsize - 1;
    d.seed = p->seed; d.ctl = p->ctl;
    d.key = p->key; d.far = p->far; d.far_cap = p->far_cap;
    d.delta = band_width(p);
    d.inval_tol = (p->kind == TASI_DFM) ? 1e-6f : 0.0f;
    d.n_focus = 0;
    d.focus_slack = p->focus_slack;
    d.band_rt = p->band_rt;
    d.ft = p->ft;
    d.inbox = p->inbox;
    d.inbox_cap = p->inbox_cap;
    d.fmode = 0;
    static const int rel_passes = [] { const char *e = getenv("TASI_REL_PASSES"); return e ? atoi(e) : 8; }();
    d.rel_passes = rel_passes;
    // bucketed near queue: relax launches of an unsharded grid (run_solver switches it off for the invalidate launch)
    d.nb = active_rings(p);
    d.bq = p->bq;
    d.rkey = p->rkey;
    static const int chain = [] { const char *e = getenv("TASI_CHAIN"); return e ? atoi(e) : 0; }();
    d.chain = chain;
    // sliding key window: plain relax launches of an unsharded grid (run_solver switches it off for focused plans,
    // launches with a limit and resumed ones, which need the quiescent points of the threshold phases)
    static const int slide = [] { const char *e = getenv("TASI_SLIDE"); return e ? atoi(e) : 0; }();
    d.slide = (slide && d.nb > 0 && d.nb <= 31 && p->band_rt == 0 && !p->ft && !(p->rkey && p->defer_tiles >= 0.0f)) ? 1 : 0;
    d.bw = d.nb > 0 ? d.delta / (float) d.nb : 0.0f;
    d.defer_margin = (d.nb > 0 && p->rkey && p->defer_tiles >= 0.0f)
                         ? p->defer_tiles * (float) TS * p->mean_cost : -1.0f;
    d.bqmask = p->bqsize ? p->bqsize - 1 : 0;
    d.inv_bw = d.nb > 0 && d.delta > 0.0f ? (float) d.nb / d.delta : 0.0f;
    if (p->focus && !p->open_top && !p->open_bottom && p->band_rt == 0) {
        // Graph::set_start: start_cell_ = Cell(start) rounds (Cell.cpp:20-21); FD*/MFD* wait for the cell's four
        // corner nodes (FieldDPlanner_impl.h:12, Cell.cpp:58-60), MS-DFM for the cell itself
        const int cx = (int) roundf(p->start_x), cy = (int) roundf(p->start_y);
        const int n = p->kind == TASI_DFM ? 1 : 4;
        for (int i = 0; i < n; ++i) {
            const int x = cx + (i >> 1), y = cy + (i & 1);
            if (x >= 0 && y >= 0 && x < p->EX && y < p->EY) d.focus[d.n_focus++] = make_int2(x, y);
        }
    }
    return d;
}

void free_grid(tasi_planner *p) {
    cudaFree(p->g_base); cudaFree(p->c_base); cudaFree(p->state); cudaFree(p->queue); cudaFree(p->seed);
    cudaFree(p->stage); cudaFree(p->key); cudaFree(p->far); cudaFree(p->bq); cudaFree(p->rkey);
    p->bq = nullptr; p->rkey = nullptr;
    p->key = nullptr; p->far = nullptr;
    p->g_base = nullptr; p->c_base = nullptr; p->state = nullptr; p->queue = nullptr; p->seed = nullptr;
    p->stage = nullptr;
}

typedef void (*solver_fn)(GridDesc, TmaMaps);
template <int SW, int CMP>
solver_fn pick_solver_sw(int kind, int mode) {
    switch (kind * 2 + mode) {
        case TASI_FD * 2 + 0: return solver_kernel<TASI_KIND_FD, 0, SW, CMP>;
        case TASI_FD * 2 + 1: return solver_kernel<TASI_KIND_FD, 1, SW, 0>;
        case TASI_DFM * 2 + 0: return solver_kernel<TASI_KIND_DFM, 0, SW, CMP>;
        case TASI_DFM * 2 + 1: return solver_kernel<TASI_KIND_DFM, 1, SW, 0>;
        case TASI_MFD * 2 + 0: return solver_kernel<TASI_KIND_MFD, 0, SW, CMP>;
        default: return solver_kernel<TASI_KIND_MFD, 1, SW, 0>;
    }
}
// How the relax visits evaluate: 0 strips (relax_tile), 1 compacted over the CTA, dirty elements found by scanning the
// windows (relax_tile_compact), 2 compacted inside each warp (relax_tile_wcompact), 3 compacted over the CTA, dirty marks
// pushed by the evaluations (relax_tile_push). TASI_COMPACT=0|1|2|3 overrides the per-planner default.
inline int compact_mode(int kind) {
    static const int env = [] {
        const char *e = getenv("TASI_COMPACT");
        return e ? atoi(e) : -1;
    }();
    if (env >= 0 && env <= 3) return env;
    return kind == TASI_FD ? 3 : 0;
}
// ... per launch: MFD* gains from the CTA list only where the solve is throughput-bound — the initial plan of a large grid
// (16384^2: 134.8 -> 130.2 ms; 4096^2 loses, 17.7 -> 19.9 ms, and so does the small active region of a replan)
inline int compact_for(const tasi_planner *p) {
    const int c = compact_mode(p->kind);
    static const bool fixed = getenv("TASI_COMPACT") != nullptr;
    if (!fixed && p->kind == TASI_MFD && p->fresh_plan && (int64_t) p->TX * p->TY >= 50000 && p->band_rt == 0) return 3;
    return c;
}
// TASI_TMA=1: the relax instances stage their tiles with cp.async.bulk.tensor.2d instead of vector loads (strips only;
// measured alternative, profiles/r02_tma_staging.txt)
inline bool tma_mode() {
    static const bool on = [] {
        const char *e = getenv("TASI_TMA");
        return e && atoi(e) == 1;
    }();
    return on;
}
template <int SW>
solver_fn pick_solver_tma(int kind) {
    switch (kind) {
        case TASI_FD: return solver_kernel<TASI_KIND_FD, 0, SW, 0, 1>;
        case TASI_DFM: return solver_kernel<TASI_KIND_DFM, 0, SW, 0, 1>;
        default: return solver_kernel<TASI_KIND_MFD, 0, SW, 0, 1>;
    }
}
// strip = elements per thread of a tile row: 4 (256-thread CTAs, two per SM) or 2 (512-thread CTAs, one per SM)
solver_fn pick_solver(int kind, int mode, int strip, int cmp = -1) {
    if (tma_mode() && mode == 0) return strip == 2 ? pick_solver_tma<2>(kind) : pick_solver_tma<4>(kind);
    switch (cmp >= 0 ? cmp : compact_mode(kind)) {
        case 3: return strip == 2 ? pick_solver_sw<2, 3>(kind, mode) : pick_solver_sw<4, 3>(kind, mode);
        case 2: return strip == 2 ? pick_solver_sw<2, 2>(kind, mode) : pick_solver_sw<4, 2>(kind, mode);
        case 1: return strip == 2 ? pick_solver_sw<2, 1>(kind, mode) : pick_solver_sw<4, 1>(kind, mode);
        default: return strip == 2 ? pick_solver_sw<2, 0>(kind, mode) : pick_solver_sw<4, 0>(kind, mode);
    }
}

// The two tensor maps of the TMA-staged solver instances: the padded value array as a 2-D float32 tensor with a 40 x 34
// box (a tile's 34 x 34 window from the 16-byte aligned column left of it), the padded cost array as a uint8 tensor with a
// 48 x 33 box. cuTensorMapEncodeTiled is a driver entry point; it is taken from the runtime (cudaGetDriverEntryPoint).
int make_tensor_maps(tasi_planner *p) {
    typedef CUresult (*encode_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void *fn = nullptr;
    cudaDriverEntryPointQueryResult qres;
    CU_TRY(p, cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
    if (!fn || qres != cudaDriverEntryPointSuccess) return fail(p, TASI_ERR_CUDA, "cuTensorMapEncodeTiled not available");
    alignas(64) CUtensorMap m[2];
    const cuuint32_t estr[2] = {1, 1};
    {
        const cuuint64_t dims[2] = {(cuuint64_t) p->gpitch, (cuuint64_t) p->g_rows};
        const cuuint64_t strides[1] = {(cuuint64_t) p->gpitch * sizeof(float)};
        const cuuint32_t box[2] = {(cuuint32_t) TMA_GW, (cuuint32_t) (TS + 2)};
        const CUresult r = ((encode_fn) fn)(&m[0], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, p->g_base, dims, strides, box, estr,
                                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                            CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return fail(p, TASI_ERR_CUDA, "cuTensorMapEncodeTiled (values) failed");
    }
    {
        const cuuint64_t dims[2] = {(cuuint64_t) p->cpitch, (cuuint64_t) p->g_rows};
        const cuuint64_t strides[1] = {(cuuint64_t) p->cpitch};
        const cuuint32_t box[2] = {(cuuint32_t) TMA_CW, (cuuint32_t) (TS + 1)};
        const CUresult r = ((encode_fn) fn)(&m[1], CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, p->c_base, dims, strides, box, estr,
                                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                            CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return fail(p, TASI_ERR_CUDA, "cuTensorMapEncodeTiled (costs) failed");
    }
    static_assert(sizeof(CUtensorMap) == sizeof(p->tm.g), "CUtensorMap is 128 bytes");
    memcpy(p->tm.g, &m[0], sizeof(CUtensorMap));
    memcpy(p->tm.c, &m[1], sizeof(CUtensorMap));
    return TASI_OK;
}

// (Re)allocate the device arrays for a map of `rows` x `W` cells. Margins: cost bytes 255, values INF.
// The geometry is committed only once every allocation has succeeded: a failed cudaMalloc leaves the planner
// without a grid (W = rows = 0, all pointers null), so a retry of set_map with the same shape allocates again
// instead of taking the "same shape" short cut over null arrays.
int alloc_grid_arrays(tasi_planner *p, int EY, int nt, unsigned q) {
    CU_TRY(p, cudaMalloc(&p->g_base, sizeof(float) * p->gpitch * p->g_rows));
    CU_TRY(p, cudaMalloc(&p->c_base, (size_t) p->cpitch * p->g_rows));
    CU_TRY(p, cudaMalloc(&p->state, sizeof(int) * nt));
    CU_TRY(p, cudaMalloc(&p->queue, sizeof(int) * q));
    CU_TRY(p, cudaMalloc(&p->seed, nt));
    CU_TRY(p, cudaMalloc(&p->key, sizeof(int) * (size_t) ((nt + 3) / 4 * 4)));
    CU_TRY(p, cudaMalloc(&p->far, sizeof(int) * 2 * (size_t) p->far_cap));
    CU_TRY(p, cudaMalloc(&p->stage, sizeof(float) * 2 * (size_t) EY));
    CU_TRY(p, cudaMemsetAsync(p->c_base, 255, (size_t) p->cpitch * p->g_rows, p->stream));
    CU_TRY(p, cudaMemsetAsync(p->state, 0, sizeof(int) * nt, p->stream));
    CU_TRY(p, cudaMemsetAsync(p->queue, 0, sizeof(int) * q, p->stream));
    CU_TRY(p, cudaMemsetAsync(p->seed, 0, nt, p->stream));
    p->bqsize = 0;
    if (p->nb > 0 && !p->open_top && !p->open_bottom) {
        unsigned bs = 1;
        while (bs < (unsigned) nt + 1u) bs <<= 1;
        CU_TRY(p, cudaMalloc(&p->bq, sizeof(int) * (size_t) bs * p->nb));
        CU_TRY(p, cudaMemsetAsync(p->bq, 0, sizeof(int) * (size_t) bs * p->nb, p->stream));
        p->bqsize = bs;
        if (p->defer_tiles >= 0.0f) CU_TRY(p, cudaMalloc(&p->rkey, sizeof(int) * (size_t) nt));
    }
    if (tma_mode()) return make_tensor_maps(p);
    return TASI_OK;
}

int alloc_grid_dims(tasi_planner *p, int W, int rows, int EX, int EY);

int alloc_grid(tasi_planner *p, int W, int rows) {
    int EX, EY;
    if (node_based(p)) { EX = rows + 1 - (p->open_top ? 1 : 0) - (p->open_bottom ? 1 : 0); EY = W + 1; }
    else { EX = rows; EY = W; }
    return alloc_grid_dims(p, W, rows, EX, EY);
}

// `rows` cell rows stored, EX x EY elements (a band stack of a sharded planner passes its own extents)
int alloc_grid_dims(tasi_planner *p, int W, int rows, int EX, int EY) {
    if (EX <= 0 || EY <= 0) { p->err = "map too small for this band configuration"; return TASI_ERR_ARG; }
    if (p->g_base && p->W == W && p->rows == rows && p->EX == EX && p->EY == EY) return TASI_OK;
    free_grid(p);
    p->W = p->rows = p->EX = p->EY = p->TX = p->TY = 0;
    p->n_tiles_alloc = 0;
    const int TX = (EX + TS - 1) / TS, TY = (EY + TS - 1) / TS;
    const int nt = TX * TY;
    int dev_blocks = 0;
    // A relaxation round is issue- and latency-bound inside its SM, and a mid-size grid's front never offers more
    // active tiles than SMs x a few. There ONE CTA per SM with 512 threads and 2-element strips halves the serial
    // work of a thread per round and beats two 256-thread CTAs per SM; from some grid size on the second CTA's
    // extra tile in flight wins. Measured cross-over (tiles), per operator weight (scripts/quick_bench.py with
    // TASI_STRIP=2|4; FD* 4096^2 34.1 -> 30.0 ms, 8192^2 97 -> 86 ms; MFD* 4096^2 21.5 -> 18.0 ms; MS-DFM 2048^2 8.4 -> 8.1;
    // re-measured after the in-strip Gauss-Seidel order: MFD* 8192^2 is 52.9 ms with 2-element strips, 46.4 ms with 4):
    // (re-measured under the bucketed near queue, profiles/r02_buckets_strip_x_compaction.txt: FD* 8192^2 = 66 k tiles is
    // 55.6 ms with 2-element strips, 48.4 ms with 4)
    const int strip2_upto = p->kind == TASI_DFM ? 4200 : 33000;
    int strip = nt <= strip2_upto ? 2 : 4;
    if (const char *e = getenv("TASI_STRIP")) {
        const int v = atoi(e);
        if (v == 2 || v == 4) strip = v;
    }
    const int nthreads = TS * TS / strip;
    CU_TRY(p, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&dev_blocks, pick_solver(p->kind, 0, strip), nthreads, 0));
    int dev_blocks1 = 0;
    CU_TRY(p, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&dev_blocks1, pick_solver(p->kind, 1, strip), nthreads, 0));
    if (dev_blocks1 < dev_blocks) dev_blocks = dev_blocks1;
    if (p->kind == TASI_MFD) {  // compact_for() may pick the CTA-list instance for an initial plan
        CU_TRY(p, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&dev_blocks1, pick_solver(p->kind, 0, strip, 3), nthreads, 0));
        if (dev_blocks1 < dev_blocks) dev_blocks = dev_blocks1;
    }
    if (dev_blocks < 1) { p->err = "solver kernel does not fit on an SM"; return TASI_ERR_CUDA; }
    int blocks = dev_blocks * p->num_sms;
    if (const char *e = getenv("TASI_BLOCKS_PER_SM")) {
        int v = atoi(e);
        if (v >= 1 && v <= dev_blocks) blocks = v * p->num_sms;
    }
    if (blocks > nt) blocks = nt;
    // ring of the work queue: a tile is queued at most once and q_push publishes without waiting for its slot, so the
    // ring holds 2 x max(tiles, CTAs) + 1 slots rounded up to a power of two
    unsigned q = 1;
    while (q < 2u * (unsigned) std::max(nt, blocks) + 1u) q <<= 1;
    p->gpitch = (int64_t) TY * TS + 2 * GMARGIN;
    p->cpitch = (int64_t) TY * TS + 2 * GMARGIN;
    p->g_rows = (int64_t) TX * TS + 2;
    p->far_cap = 2 * nt + 64;
    const int rc = alloc_grid_arrays(p, EY, nt, q);
    if (rc != TASI_OK) {
        free_grid(p);
        return rc;
    }
    p->W = W; p->rows = rows; p->EX = EX; p->EY = EY; p->TX = TX; p->TY = TY;
    p->strip = strip;
    p->solver_blocks = blocks;
    p->qsize = q;
    p->n_tiles_alloc = nt;
    p->initialize_search = true;  // a new geometry invalidates any previous field
    p->patch_pending = false;
    return TASI_OK;
}

inline double now_ms() {
    return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

int read_ctl(tasi_planner *p, DevCtl *h) {
    CU_TRY(p, cudaMemcpyAsync(h, p->ctl, sizeof(DevCtl), cudaMemcpyDeviceToHost, p->stream));
    CU_TRY(p, cudaStreamSynchronize(p->stream));
    return TASI_OK;
}

// Mean free-cell cost of the current map (lazily, once per set_map): the band width of the schedule scales with it.
int ensure_cost_stats(tasi_planner *p) {
    if (p->cost_stats_valid) return TASI_OK;
    if (!p->cost_sum) CU_TRY(p, cudaMalloc(&p->cost_sum, 2 * sizeof(unsigned long long)));
    CU_TRY(p, cudaMemsetAsync(p->cost_sum, 0, 2 * sizeof(unsigned long long), p->stream));
    // sample at most ~16M cells: whole rows from the top of the band
    const int rows = (int) std::min<int64_t>(p->rows, std::max<int64_t>(1, (int64_t) (1 << 24) / p->W));
    cost_stats_kernel<<<p->num_sms * 4, 256, 0, p->stream>>>(c_origin(p), p->cpitch, rows, p->W, p->thr, p->cost_sum);
    CU_TRY(p, cudaGetLastError());
    unsigned long long h[2];
    CU_TRY(p, cudaMemcpyAsync(h, p->cost_sum, sizeof h, cudaMemcpyDeviceToHost, p->stream));
    CU_TRY(p, cudaStreamSynchronize(p->stream));
    p->mean_cost = h[1] ? (float) ((double) h[0] / (double) h[1]) : 1.0f;
    if (p->mean_cost < 1.0f) p->mean_cost = 1.0f;
    p->cost_stats_valid = true;
    return TASI_OK;
}

// The pushed-mark list instances (80 registers) fit three CTAs per SM. The third one pays on large grids, where the
// solve is bound by what an SM issues (MFD* 16384^2 101 -> 92 ms, FD* 122 -> 116 ms), and costs on smaller ones, where
// more workers only inflate the visits (8192^2: 35.2 -> 36.2 / 43.7 -> 47.3 ms): three from 150 k tiles. Every CTA of
// the launch must be resident, so the count comes from the occupancy of THE instance that is launched.
// Returns the CTA count of such a launch, 0 = keep p->solver_blocks, < 0 = error.
int relax_blocks(tasi_planner *p, int mode, int nt) {
    static const int third_from = [] { const char *e = getenv("TASI_THIRD_CTA_FROM"); return e ? atoi(e) : 150000; }();
    static const bool own_occ = [] { const char *e = getenv("TASI_OWN_OCC"); return e && atoi(e) != 0; }();
    if (p->max_blocks > 0 || mode != 0 || p->strip != 4) return 0;
    if (!(own_occ || (compact_for(p) == 3 && nt >= third_from))) return 0;
    int occ = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, pick_solver(p->kind, mode, p->strip, compact_for(p)),
                                                      TS * TS / p->strip, 0) != cudaSuccess) return 0;
    return std::min(std::max(occ, 1) * p->num_sms, nt);
}

// Queue the flagged tiles and run the persistent solver in `mode` until the queue drains.
// Adds the launch's counters to p->stats; *ms = CUDA-event time of the three launches.
int run_solver(tasi_planner *p, int mode, float *ms, bool resume = false, float limit = INFINITY) {
    int rcs = ensure_cost_stats(p);
    if (rcs != TASI_OK) return rcs;
    GridDesc d = make_desc(p);
    if (mode == 1) { d.n_focus = 0; d.nb = 0; d.defer_margin = -1.0f; }
    if (mode == 1 || d.n_focus > 0 || resume || limit < INFINITY || d.nb <= 0) d.slide = 0;
    if (d.slide) CU_TRY(p, cudaMemsetAsync(p->far, 0, sizeof(int) * (size_t) p->far_cap, p->stream));  // 0 = slot not written
    static const bool repair_buckets = [] { const char *e = getenv("TASI_REPAIR_BUCKETS"); return !e || atoi(e) != 0; }();
    if (!p->fresh_plan && !repair_buckets) d.nb = 0;
    DevCtl z;
    memset(&z, 0, sizeof z);
    z.timeout_ns = (unsigned long long) (p->timeout_s * 1e9);
    z.thr_bits = (mode == 1 || p->delta_factor <= 0.0f) ? INF_BITS : 0;  // invalidation needs no order
    memcpy(&z.limit_bits, &limit, sizeof(int));
    z.far_min_bits = INF_BITS;
    if (resume) {  // the far pile (device arrays + tile states) is still as the previous launch left it
        z.far_sel = p->sv_far_sel;
        z.far_count[0] = p->sv_far_count[0];
        z.far_count[1] = p->sv_far_count[1];
        z.thr_bits = (mode == 1) ? INF_BITS : p->sv_thr_bits;
    }
    // reset queue tickets + per-launch counters, keep the patch bookkeeping
    CU_TRY(p, cudaMemcpyAsync(p->ctl, &z, offsetof(DevCtl, n_changed), cudaMemcpyHostToDevice, p->stream));
    CU_TRY(p, cudaEventRecord(p->ev0, p->stream));
    const int nt = p->TX * p->TY;
    const int ordered = (mode == 0 && p->delta_factor > 0.0f) ? 1 : 0;
    if (ordered) seed_keys_kernel<<<std::min((nt + 7) / 8, p->num_sms * 8), 256, 0, p->stream>>>(d);
    enqueue_seeds_kernel<<<std::min((nt + 255) / 256, 1024), 256, 0, p->stream>>>(d, ordered);
    arm_solver_kernel<<<1, 1, 0, p->stream>>>(d);
    int nblocks = (p->max_blocks > 0 && p->max_blocks < p->solver_blocks) ? p->max_blocks : p->solver_blocks;
    {
        const int rb = relax_blocks(p, mode, nt);
        if (rb > 0) { nblocks = rb; p->stats.solver_blocks = nblocks; }
    }
    pick_solver(p->kind, mode, p->strip, compact_for(p))<<<nblocks, TS * TS / p->strip, 0, p->stream>>>(d, p->tm);
    CU_TRY(p, cudaGetLastError());
    CU_TRY(p, cudaEventRecord(p->ev1, p->stream));
    DevCtl h;
    int rc = read_ctl(p, &h);
    if (rc != TASI_OK) return rc;
    CU_TRY(p, cudaEventElapsedTime(ms, p->ev0, p->ev1));
#ifdef TASI_DEBUG_COMPACT
    if (h.kick >= 100000) { p->err = "compact debug flag " + std::to_string(h.kick); return TASI_ERR_STATE; }
#endif
    if (d.slide && getenv("TASI_SLIDE_DEBUG"))
        fprintf(stderr, "slide: abase %d scans %llu far_count %d far_scanned %d jumps %d visits %llu bw %.1f\n", h.abase,
                h.slide_scans, h.far_count[0], h.far_scanned, h.phases, h.visits, d.bw);
    p->sv_far_sel = h.far_sel;
    p->sv_far_count[0] = std::min(h.far_count[0], p->far_cap);
    p->sv_far_count[1] = std::min(h.far_count[1], p->far_cap);
    p->sv_thr_bits = h.thr_bits;
    p->sv_far_min_bits = (p->sv_far_count[h.far_sel] > 0) ? h.far_min_bits : INF_BITS;
    if (mode == 0) {
        p->pile_open = p->sv_far_count[h.far_sel] > 0 && !p->open_top && !p->open_bottom;
        p->stats.focus_hit = h.focus_hit;
        memcpy(&p->stats.settled_below, &p->sv_far_min_bits, sizeof(float));
    }
    p->stats.solver_launches += 3 + ordered;
    p->stats.queue_pushes += h.pushes;
    p->stats.phases += (uint64_t) h.phases;
    p->stats.far_parked += h.far_parked;
    if (mode == 0) {
        p->stats.tile_visits += h.visits;
        p->stats.tile_rounds += h.rounds;
        p->stats.elem_lowered += h.lowered;
        p->stats.elem_evals += h.evals;
        p->stats.cyc_pop += h.cyc_pop; p->stats.cyc_stage += h.cyc_stage;
        p->stats.cyc_relax += h.cyc_relax; p->stats.cyc_post += h.cyc_post;
        p->stats.cyc_advance += h.cyc_advance;
    } else {
        p->stats.inval_visits += h.visits;
        p->stats.elem_invalidated += h.invalidated;
    }
    if (h.error) {
        p->err = "solver watchdog fired (work queue stalled)";
        // leave the queue/state arrays clean for the next attempt
        cudaMemsetAsync(p->state, 0, sizeof(int) * nt, p->stream);
        cudaMemsetAsync(p->queue, 0, sizeof(int) * p->qsize, p->stream);
        if (p->bq) cudaMemsetAsync(p->bq, 0, sizeof(int) * (size_t) p->bqsize * p->nb, p->stream);
        p->initialize_search = true;
        return TASI_ERR_TIMEOUT;
    }
    return TASI_OK;
}

// init(): whole field INF, goal 0, goal tile flagged (FieldDPlanner_impl.h:15-21 and siblings).
int init_field(tasi_planner *p) {
    GridDesc d = make_desc(p);
    const int nt = p->TX * p->TY;
    CU_TRY(p, cudaMemsetAsync(p->seed, 0, nt, p->stream));
    CU_TRY(p, cudaMemsetAsync(p->state, 0, sizeof(int) * nt, p->stream));
    const int64_t n4 = p->gpitch * p->g_rows / 4;
    fill_inf_kernel<<<p->num_sms * 8, 256, 0, p->stream>>>((float4 *) p->g_base, n4);
    fill_inf_kernel<<<64, 256, 0, p->stream>>>((float4 *) p->key, (nt + 3) / 4);
    seed_goal_kernel<<<1, 1, 0, p->stream>>>(d);
    CU_TRY(p, cudaGetLastError());
    // patch bookkeeping restarts: the first plan reads the patched grid (ReplannerBase.h:52)
    CU_TRY(p, cudaMemsetAsync(&p->ctl->n_changed, 0, sizeof(DevCtl) - offsetof(DevCtl, n_changed), p->stream));
    p->patch_pending = false;
    p->last_seed_count = 0;
    p->stats.solver_launches += 2;
    return TASI_OK;
}

}  // namespace

#include "tasi_shard.cuh"

extern "C" {

int tasi_create(int kind, int level, int device, tasi_planner **out) {
    if (!out) return TASI_ERR_ARG;
    *out = nullptr;
    if (kind < TASI_FD || kind > TASI_MFD || level < 0 || level > (kind == TASI_MFD ? 2 : 1)) return TASI_ERR_ARG;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0 || device < 0 || device >= ndev) return TASI_ERR_CUDA;
    tasi_planner *p = new (std::nothrow) tasi_planner();
    if (!p) return TASI_ERR_ARG;
    p->kind = kind; p->level = level; p->device = device;
    if (cudaSetDevice(device) != cudaSuccess) { delete p; return TASI_ERR_CUDA; }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) { delete p; return TASI_ERR_CUDA; }
    p->num_sms = prop.multiProcessorCount;
    if (cudaStreamCreateWithFlags(&p->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaEventCreate(&p->ev0) != cudaSuccess || cudaEventCreate(&p->ev1) != cudaSuccess ||
        cudaEventCreate(&p->evs0) != cudaSuccess || cudaEventCreate(&p->evs1) != cudaSuccess ||
        cudaMalloc(&p->ctl, sizeof(DevCtl)) != cudaSuccess ||
        cudaMemset(p->ctl, 0, sizeof(DevCtl)) != cudaSuccess) {
        tasi_destroy(p);
        return TASI_ERR_CUDA;
    }
    if (const char *e = getenv("TASI_TIMEOUT_S")) {
        double v = atof(e);
        if (v > 0) p->timeout_s = v;
    }
    if (const char *e = getenv("TASI_DEFER")) p->defer_tiles = (float) atof(e);  // margin in tile widths; < 0 = off
    if (const char *e = getenv("TASI_BUCKETS")) {  // key-ordered rings of the near queue (0 = one FIFO ring)
        const int v = atoi(e);
        if (v >= 0 && v <= 62) p->nb = v;
    }
    if (const char *e = getenv("TASI_DELTA_FACTOR")) {  // fixed factor; <= 0: no ordering
        p->delta_factor = (float) atof(e);
        p->delta_auto = false;
    }
    *out = p;
    return TASI_OK;
}

int tasi_destroy(tasi_planner *p) {
    if (!p) return TASI_OK;
    if (p->shard) {
        shard_destroy(p->shard);
        delete p;
        return TASI_OK;
    }
    cudaSetDevice(p->device);
    if (p->stream) cudaStreamSynchronize(p->stream);
    free_grid(p);
    cudaFree(p->ctl); cudaFree(p->cost_sum); cudaFree(p->patch_buf); cudaFree(p->chg_buf); cudaFree(p->xgoals_dev); cudaFree(p->walk_buf); cudaFreeHost(p->walk_host);
    if (p->ev0) cudaEventDestroy(p->ev0);
    if (p->ev1) cudaEventDestroy(p->ev1);
    if (p->evs0) cudaEventDestroy(p->evs0);
    if (p->evs1) cudaEventDestroy(p->evs1);
    if (p->stream) cudaStreamDestroy(p->stream);
    delete p;
    return TASI_OK;
}

int tasi_reset(tasi_planner *p) {
    if (!p) return TASI_ERR_ARG;
    if (p->shard) p->shard->initialize_search = true;
    p->initialize_search = true;
    return TASI_OK;
}

int tasi_set_occupancy_threshold(tasi_planner *p, float t) {
    if (!p) return TASI_ERR_ARG;
    p->thr = (int) (unsigned char) (t * 255.0f);  // Graph.cpp:18-20 (member is an unsigned char)
    if (p->shard) {
        p->shard->thr = p->thr;
        for (ShardMember &m : p->shard->local) m.p->thr = p->thr;
    }
    return TASI_OK;
}

int tasi_set_heuristic_multiplier(tasi_planner *p, float mult) {
    if (!p) return TASI_ERR_ARG;
    if (p->shard) p->shard->heuristic_multiplier = mult;
    p->heuristic_multiplier = mult;
    return TASI_OK;
}

static int set_map_common(tasi_planner *p, const uint8_t *src, int width, int length, int64_t pitch,
                          cudaMemcpyKind kind) {
    if (!p || !src || width <= 0 || length <= 0 || pitch < width) return TASI_ERR_ARG;
    CU_TRY(p, cudaSetDevice(p->device));
    int rc = alloc_grid(p, width, length);
    if (rc != TASI_OK) return rc;
    // band with an open top: the first stored row is the cell row above the first owned node row
    const int dst_row0 = (node_based(p) && p->open_top) ? -1 : 0;
    CU_TRY(p, cudaMemcpy2DAsync(c_origin(p) + (int64_t) dst_row0 * p->cpitch, p->cpitch, src, pitch, width, length,
                                kind, p->stream));
    CU_TRY(p, cudaStreamSynchronize(p->stream));  // the host pointer is not retained
    p->initialize_graph = false;
    p->cost_stats_valid = false;
    return TASI_OK;
}

int tasi_set_map(tasi_planner *p, const uint8_t *host_map, int width, int length) {
    if (p && p->shard) return shard_set_map(p->shard, host_map, width, length);
    return set_map_common(p, host_map, width, length, width, cudaMemcpyHostToDevice);
}

int tasi_set_map_device(tasi_planner *p, const uint8_t *dev_map, int width, int length, int64_t pitch) {
    if (p && p->shard) return sh_fail(p->shard, TASI_ERR_STATE, "set_map_device: not available on a sharded planner");
    return set_map_common(p, dev_map, width, length, pitch, cudaMemcpyDeviceToDevice);
}

int tasi_set_map_stack(tasi_planner *p, const uint8_t *const *maps, int n_maps, const int *index, int n_blocks,
                       int width, int length) {
    if (!p || !maps || !index || n_maps <= 0 || n_blocks <= 0 || width <= 0 || length <= 0) return TASI_ERR_ARG;
    if (p->shard) return sh_fail(p->shard, TASI_ERR_STATE, "set_map_stack: not available on a sharded planner");
    if (p->open_top || p->open_bottom) return fail(p, TASI_ERR_STATE, "set_map_stack: not available on a row band");
    for (int b = 0; b < n_blocks; ++b)
        if (index[b] < 0 || index[b] >= n_maps || !maps[index[b]]) return fail(p, TASI_ERR_ARG, "set_map_stack: bad index");
    CU_TRY(p, cudaSetDevice(p->device));
    const int64_t rows = (int64_t) n_blocks * (length + 1) - 1;
    if (rows > INT32_MAX) return fail(p, TASI_ERR_ARG, "set_map_stack: too many rows");
    int rc = alloc_grid(p, width, (int) rows);
    if (rc != TASI_OK) return rc;
    // every distinct map crosses the bus once (into the first block that shows it); its other blocks are device copies
    std::vector<int> first(n_maps, -1);
    for (int b = 0; b < n_blocks; ++b) {
        uint8_t *dst = c_origin(p) + (int64_t) b * (length + 1) * p->cpitch;
        const int m = index[b];
        if (first[m] < 0) {
            first[m] = b;
            CU_TRY(p, cudaMemcpy2DAsync(dst, p->cpitch, maps[m], width, width, length, cudaMemcpyHostToDevice, p->stream));
        } else {
            const uint8_t *src = c_origin(p) + (int64_t) first[m] * (length + 1) * p->cpitch;
            CU_TRY(p, cudaMemcpy2DAsync(dst, p->cpitch, src, p->cpitch, width, length, cudaMemcpyDeviceToDevice, p->stream));
        }
        if (b + 1 < n_blocks)  // the obstacle row that keeps the blocks apart
            CU_TRY(p, cudaMemset2DAsync(dst + (int64_t) length * p->cpitch, p->cpitch, 255, width, 1, p->stream));
    }
    CU_TRY(p, cudaStreamSynchronize(p->stream));  // host pointers are not retained
    p->initialize_graph = false;
    p->cost_stats_valid = false;
    return TASI_OK;
}

int tasi_synth_map(tasi_planner *p, uint64_t seed, int width, int length, int64_t row0, int p_obs, int sx, int sy,
                   int gx, int gy) {
    if (!p || width <= 0 || length <= 0) return TASI_ERR_ARG;
    if (p->shard) {
        if (row0 != 0) return sh_fail(p->shard, TASI_ERR_ARG, "synth_map: a sharded planner generates the whole map (row0 = 0)");
        return shard_synth_map(p->shard, seed, width, length, p_obs, sx, sy, gx, gy);
    }
    CU_TRY(p, cudaSetDevice(p->device));
    int rc = alloc_grid(p, width, length);
    if (rc != TASI_OK) return rc;
    const int dst_row0 = (node_based(p) && p->open_top) ? -1 : 0;
    synth_map_kernel<<<p->num_sms * 8, 256, 0, p->stream>>>(c_origin(p), p->cpitch, length, width, row0, dst_row0, seed,
                                                           p_obs, sx, sy, gx, gy);
    CU_TRY(p, cudaGetLastError());
    CU_TRY(p, cudaStreamSynchronize(p->stream));
    p->initialize_graph = false;
    p->cost_stats_valid = false;
    return TASI_OK;
}

int tasi_read_map(tasi_planner *p, uint8_t *host_out) {
    if (p && p->shard) return shard_read_map(p->shard, host_out);
    if (!p || !host_out || p->initialize_graph) return TASI_ERR_ARG;
    CU_TRY(p, cudaSetDevice(p->device));
    const int dst_row0 = (node_based(p) && p->open_top) ? -1 : 0;
    CU_TRY(p, cudaMemcpy2DAsync(host_out, p->W, c_origin(p) + (int64_t) dst_row0 * p->cpitch, p->cpitch, p->W, p->rows,
                                cudaMemcpyDeviceToHost, p->stream));
    CU_TRY(p, cudaStreamSynchronize(p->stream));
    return TASI_OK;
}

int tasi_patch_map(tasi_planner *p, const uint8_t *host_patch, int x, int y, int w, int h, uint64_t *n_changed) {
    if (p && p->shard) return shard_patch_map(p->shard, host_patch, x, y, w, h, n_changed);
    if (!p || !host_patch || p->initialize_graph) return TASI_ERR_ARG;
    if (x < 0 || y < 0 || w <= 0 || h <= 0 || (int64_t) x + h > p->rows || (int64_t) y + w > p->W) {
        p->err = "patch outside the map";  // the reference asserts (Graph.cpp:38-41)
        return TASI_ERR_ARG;
    }
    if (p->open_top || p->open_bottom) { p->err = "patch_map is not supported on a band"; return TASI_ERR_STATE; }
    CU_TRY(p, cudaSetDevice(p->device));
    const size_t n = (size_t) w * h;
    if (n > p->patch_cap) {
        cudaFree(p->patch_buf); cudaFree(p->chg_buf);
        p->patch_buf = p->chg_buf = nullptr; p->patch_cap = 0;
        CU_TRY(p, cudaMalloc(&p->patch_buf, n));
        CU_TRY(p, cudaMalloc(&p->chg_buf, n));
        p->patch_cap = n;
    }
    CU_TRY(p, cudaMemcpyAsync(p->patch_buf, host_patch, n, cudaMemcpyHostToDevice, p->stream));
    CU_TRY(p, cudaMemsetAsync(&p->ctl->n_changed, 0, sizeof(unsigned long long), p->stream));
    GridDesc d = make_desc(p);
    const int blocks = (int) std::min<size_t>((n + 255) / 256, (size_t) p->num_sms * 16);
    patch_apply_kernel<<<blocks, 256, 0, p->stream>>>(d, p->patch_buf, p->chg_buf, x, y, w, h, 0, p->rows);
    switch (p->kind) {
        case TASI_FD: patch_seed_kernel<TASI_KIND_FD><<<blocks, 256, 0, p->stream>>>(d, p->chg_buf, x, y, w, h, 0, p->EX); break;
        case TASI_DFM: patch_seed_kernel<TASI_KIND_DFM><<<blocks, 256, 0, p->stream>>>(d, p->chg_buf, x, y, w, h, 0, p->EX); break;
        default: patch_seed_kernel<TASI_KIND_MFD><<<blocks, 256, 0, p->stream>>>(d, p->chg_buf, x, y, w, h, 0, p->EX); break;
    }
    CU_TRY(p, cudaGetLastError());
    DevCtl hc;
    int rc = read_ctl(p, &hc);
    if (rc != TASI_OK) return rc;
    if (n_changed) *n_changed = hc.n_changed;
    p->patch_pending = true;
    return TASI_OK;
}

// ---- tasi_patch_map_static: the same patch, no allocation ------------------------------------------------------------
// tasi_patch_map keeps two scratch buffers per planner (the device copy of the patch and the per-cell change flags that
// patch_apply_kernel hands to patch_seed_kernel) and grows them with cudaFree + cudaMalloc when a larger patch arrives.
// Callers that must not allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc)
// get this variant: the scratch is a file-scope array in device memory, 2 x TASI_PATCH_STATIC_MAX_CELLS bytes, that every
// GPU holds from the moment it loads the library until the process exits. In HOST code `g_patch_scratch` is only a
// stand-in for the symbol, not a device address: the address the kernels get comes from cudaGetSymbolAddress, taken once
// per device with that device current (static_patch_scratch below) — no allocation, the array is part of the module image.
// Every call on a device shares the array; it keeps the last call's contents (both halves are overwritten for the cells
// of the new patch before they are read: the H2D copy fills the patch half, the flag half is cleared and then written by
// patch_apply_kernel), and two calls that overlap on one GPU — two planners on their own streams, two host threads —
// race on it: serialise them.
static __device__ __align__(256) uint8_t g_patch_scratch[2 * TASI_PATCH_STATIC_MAX_CELLS];

static constexpr int STATIC_SCRATCH_MAX_DEVICES = 64;
static uint8_t *g_patch_scratch_addr[STATIC_SCRATCH_MAX_DEVICES] = {nullptr};
static std::mutex g_patch_scratch_mutex;

// device address of g_patch_scratch on the planner's device (which must be current), resolved once per device
static int static_patch_scratch(tasi_planner *p, uint8_t **out) {
    if (p->device < 0 || p->device >= STATIC_SCRATCH_MAX_DEVICES) return fail(p, TASI_ERR_ARG, "device index out of range");
    std::lock_guard<std::mutex> lock(g_patch_scratch_mutex);
    if (!g_patch_scratch_addr[p->device]) {
        void *ptr = nullptr;
        CU_TRY(p, cudaGetSymbolAddress(&ptr, g_patch_scratch));
        if (!ptr) return fail(p, TASI_ERR_CUDA, "cudaGetSymbolAddress(g_patch_scratch) returned a null address");
        g_patch_scratch_addr[p->device] = (uint8_t *) ptr;
    }
    *out = g_patch_scratch_addr[p->device];
    return TASI_OK;
}

int tasi_patch_map_static(tasi_planner *p, const uint8_t *host_patch, int x, int y, int w, int h, uint64_t *n_changed) {
    if (!p) return TASI_ERR_ARG;
    if (p->shard) return sh_fail(p->shard, TASI_ERR_STATE, "tasi_patch_map_static: not available on a sharded planner");
    if (!host_patch || p->initialize_graph) return TASI_ERR_ARG;
    if (x < 0 || y < 0 || w <= 0 || h <= 0 || (int64_t) x + h > p->rows || (int64_t) y + w > p->W) {
        p->err = "patch outside the map";  // the reference asserts (Graph.cpp:38-41)
        return TASI_ERR_ARG;
    }
    if ((int64_t) w * h > (int64_t) TASI_PATCH_STATIC_MAX_CELLS) {
        p->err = "tasi_patch_map_static: patch of " + std::to_string((int64_t) w * h) + " cells exceeds the maximum of " +
                 std::to_string((int64_t) TASI_PATCH_STATIC_MAX_CELLS) + " cells of the static scratch (use tasi_patch_map)";
        return TASI_ERR_ARG;
    }
    if (p->open_top || p->open_bottom) { p->err = "patch_map is not supported on a band"; return TASI_ERR_STATE; }
    CU_TRY(p, cudaSetDevice(p->device));
    uint8_t *scratch = nullptr;
    int rc = static_patch_scratch(p, &scratch);
    if (rc != TASI_OK) return rc;
    uint8_t *patch_buf = scratch, *chg_buf = scratch + TASI_PATCH_STATIC_MAX_CELLS;
    const size_t n = (size_t) w * h;
    CU_TRY(p, cudaMemcpyAsync(patch_buf, host_patch, n, cudaMemcpyHostToDevice, p->stream));
    CU_TRY(p, cudaMemsetAsync(chg_buf, 0, n, p->stream));  // the last call's flags
    CU_TRY(p, cudaMemsetAsync(&p->ctl->n_changed, 0, sizeof(unsigned long long), p->stream));
    GridDesc d = make_desc(p);
    const int blocks = (int) std::min<size_t>((n + 255) / 256, (size_t) p->num_sms * 16);
    patch_apply_kernel<<<blocks, 256, 0, p->stream>>>(d, patch_buf, chg_buf, x, y, w, h, 0, p->rows);
    switch (p->kind) {
        case TASI_FD: patch_seed_kernel<TASI_KIND_FD><<<blocks, 256, 0, p->stream>>>(d, chg_buf, x, y, w, h, 0, p->EX); break;
        case TASI_DFM: patch_seed_kernel<TASI_KIND_DFM><<<blocks, 256, 0, p->stream>>>(d, chg_buf, x, y, w, h, 0, p->EX); break;
        default: patch_seed_kernel<TASI_KIND_MFD><<<blocks, 256, 0, p->stream>>>(d, chg_buf, x, y, w, h, 0, p->EX); break;
    }
    CU_TRY(p, cudaGetLastError());
    DevCtl hc;
    rc = read_ctl(p, &hc);
    if (rc != TASI_OK) return rc;
    if (n_changed) *n_changed = hc.n_changed;
    p->patch_pending = true;
    return TASI_OK;
}

static void round_elem(float x, float y, int64_t *ex, int64_t *ey) {
    *ex = (int64_t) roundf(x);  // Node(Position) / Cell(Position): Node.cpp:14-17, Cell.cpp:20-21
    *ey = (int64_t) roundf(y);
}

int tasi_set_start(tasi_planner *p, float x, float y) {
    if (!p) return TASI_ERR_ARG;
    if (p->shard) {
        p->shard->start_x = x; p->shard->start_y = y;
        p->shard->new_start = true;
    }
    p->start_x = x; p->start_y = y;
    p->new_start = true;
    return TASI_OK;
}

int tasi_set_goal(tasi_planner *p, float x, float y) {
    if (!p) return TASI_ERR_ARG;
    int64_t ex, ey;
    round_elem(x, y, &ex, &ey);
    p->new_goal = !p->have_goal_elem || ex != p->goal_ex || ey != p->goal_ey;  // ReplannerBase.h:99-108
    if (!p->xgoals.empty()) { p->xgoals.clear(); p->new_goal = true; }
    p->goal_x = x; p->goal_y = y; p->goal_ex = ex; p->goal_ey = ey;
    p->have_goal_elem = true;
    p->goal_set = true;
    if (tasi_shard *s = p->shard) {
        s->new_goal = p->new_goal;
        s->goal_x = x; s->goal_y = y; s->goal_ex = ex; s->goal_ey = ey;
        s->have_goal_elem = true;
        s->goal_set = true;
    }
    return TASI_OK;
}

int tasi_set_goals(tasi_planner *p, const float *xy, int n) {
    if (!p || !xy || n < 1) return TASI_ERR_ARG;
    if (p->shard) return sh_fail(p->shard, TASI_ERR_STATE, "set_goals: not available on a sharded planner");
    CU_TRY(p, cudaSetDevice(p->device));
    int rc = tasi_set_goal(p, xy[0], xy[1]);
    if (rc != TASI_OK) return rc;
    p->xgoals.clear();
    for (int i = 1; i < n; ++i) {
        int64_t ex, ey;
        round_elem(xy[2 * i], xy[2 * i + 1], &ex, &ey);
        ex -= p->global_row0;
        if (ex < INT32_MIN / 2 || ex > INT32_MAX / 2) continue;
        p->xgoals.push_back(make_int2((int) ex, (int) ey));
    }
    if (p->xgoals.size() > p->xgoals_cap) {
        cudaFree(p->xgoals_dev);
        p->xgoals_dev = nullptr; p->xgoals_cap = 0;
        CU_TRY(p, cudaMalloc(&p->xgoals_dev, sizeof(int2) * p->xgoals.size()));
        p->xgoals_cap = p->xgoals.size();
    }
    if (!p->xgoals.empty()) {
        CU_TRY(p, cudaMemcpyAsync(p->xgoals_dev, p->xgoals.data(), sizeof(int2) * p->xgoals.size(),
                                  cudaMemcpyHostToDevice, p->stream));
        CU_TRY(p, cudaStreamSynchronize(p->stream));
    }
    p->new_goal = true;  // a goal set is always planned from scratch
    return TASI_OK;
}

int tasi_step(tasi_planner *p, float *u_ms, float *p_ms, uint64_t *n_updated, uint64_t *n_expanded) {
    if (!p) return TASI_ERR_ARG;
    if (p->shard) {
        const int rc = shard_step(p->shard, u_ms, p_ms, n_updated, n_expanded);
        p->new_goal = false;
        return rc;
    }
    if (p->initialize_graph) return TASI_LOOP_FAILURE_NO_GRAPH;  // ReplannerBase.h:44
    if (!p->goal_set) return TASI_LOOP_FAILURE_NO_GOAL;         // ReplannerBase.h:45
    CU_TRY(p, cudaSetDevice(p->device));
    const uint64_t nt = (uint64_t) p->TX * p->TY;
    memset(&p->stats, 0, sizeof p->stats);
    p->stats.tiles = nt;
    p->stats.solver_blocks = p->solver_blocks;
    int rc;
    float ms = 0;

    CU_TRY(p, cudaEventRecord(p->evs0, p->stream));
    double t0 = now_ms();
    if (p->initialize_search || p->new_goal) {  // ReplannerBase.h:48-55
        p->num_nodes_updated = 0;
        p->num_nodes_expanded = 0;
        CU_TRY(p, cudaEventRecord(p->ev0, p->stream));
        rc = init_field(p);
        if (rc != TASI_OK) return rc;
        CU_TRY(p, cudaEventRecord(p->ev1, p->stream));
        CU_TRY(p, cudaStreamSynchronize(p->stream));
        CU_TRY(p, cudaEventElapsedTime(&ms, p->ev0, p->ev1));
        p->stats.kernel_ms_init = ms;
    } else if (p->new_start) {  // ReplannerBase.h:56-59 -> update()
        p->new_start = false;
        if (p->patch_pending) {
            DevCtl hc;
            rc = read_ctl(p, &hc);
            if (rc != TASI_OK) return rc;
            p->last_seed_count = hc.n_seed_elems;
            // phase 1 of the repair: raised costs may have left values without support
            if (hc.n_raised > 0) {
                // tiles parked by a focused plan keep their place: the invalidate launch never releases them
                rc = run_solver(p, 1, &ms, p->pile_open, p->pile_open ? 0.0f : INFINITY);
                if (rc != TASI_OK) return rc;
                p->stats.kernel_ms_inval = ms;
            }
            CU_TRY(p, cudaMemsetAsync(&p->ctl->n_changed, 0, sizeof(DevCtl) - offsetof(DevCtl, n_changed), p->stream));
            p->patch_pending = false;
        }
        // the reference keeps Graph::updated_cells_ until the next patch, so update() re-seeds (and
        // re-counts) the same elements on every new start
        p->num_nodes_updated = p->last_seed_count;
        p->stats.seed_elems = p->last_seed_count;
        CU_TRY(p, cudaStreamSynchronize(p->stream));
    }
    double t1 = now_ms();

    const bool fresh = p->initialize_search || p->new_goal;
    if (fresh) p->pile_open = false;
    // ReplannerBase.h:64-69 -> plan(); also whenever a focused plan left work parked (the new start may lie beyond
    // what is settled)
    if (fresh || p->num_nodes_updated > 0 || p->pile_open) {
        p->fresh_plan = fresh;
        rc = run_solver(p, 0, &ms, p->pile_open);
        p->fresh_plan = false;
        if (rc != TASI_OK) return rc;
        p->stats.kernel_ms_solve = ms;
        p->num_nodes_expanded = p->stats.elem_lowered;
    } else {
        p->num_nodes_expanded = 0;
    }
    p->new_goal = p->initialize_search = false;
    if (!p->pile_open) p->stats.settled_below = INFINITY;
    CU_TRY(p, cudaEventRecord(p->evs1, p->stream));
    CU_TRY(p, cudaEventSynchronize(p->evs1));
    CU_TRY(p, cudaEventElapsedTime(&p->stats.step_ms_device, p->evs0, p->evs1));
    double t2 = now_ms();

    if (u_ms) *u_ms = (float) (t1 - t0);
    if (p_ms) *p_ms = (float) (t2 - t1);
    if (n_updated) *n_updated = p->num_nodes_updated;
    if (n_expanded) *n_expanded = p->num_nodes_expanded;
    return TASI_OK;
}

int tasi_field_dims(tasi_planner *p, int *ex, int *ey) {
    if (p && p->shard) {
        if (p->shard->initialize_graph) return TASI_ERR_ARG;
        if (ex) *ex = (int) p->shard->n_elem_rows;
        if (ey) *ey = p->shard->EY;
        return TASI_OK;
    }
    if (!p || p->initialize_graph) return TASI_ERR_ARG;
    if (ex) *ex = p->EX;
    if (ey) *ey = p->EY;
    return TASI_OK;
}

int tasi_read_window(tasi_planner *p, int x0, int y0, int nx, int ny, float *host_out) {
    if (p && p->shard) return shard_read_window(p->shard, x0, y0, nx, ny, host_out);
    if (!p || !host_out || nx <= 0 || ny <= 0 || p->initialize_graph) return TASI_ERR_ARG;
    CU_TRY(p, cudaSetDevice(p->device));
    const int xa = std::max(x0, 0), xb = std::min(x0 + nx, p->EX);
    const int ya = std::max(y0, 0), yb = std::min(y0 + ny, p->EY);
    const bool clipped = xa != x0 || ya != y0 || xb != x0 + nx || yb != y0 + ny;
    if (clipped)
        for (int64_t i = 0; i < (int64_t) nx * ny; ++i) host_out[i] = INFINITY;
    if (xa < xb && ya < yb) {
        CU_TRY(p, cudaMemcpy2DAsync(host_out + (int64_t) (xa - x0) * ny + (ya - y0), sizeof(float) * ny,
                                    g_origin(p) + (int64_t) xa * p->gpitch + ya, sizeof(float) * p->gpitch,
                                    sizeof(float) * (yb - ya), xb - xa, cudaMemcpyDeviceToHost, p->stream));
        CU_TRY(p, cudaStreamSynchronize(p->stream));
    }
    return TASI_OK;
}

int tasi_read_field(tasi_planner *p, float *host_out) {
    if (p && p->shard) {
        if (p->shard->initialize_graph) return TASI_ERR_ARG;
        return shard_read_window(p->shard, 0, 0, (int) p->shard->n_elem_rows, p->shard->EY, host_out);
    }
    if (!p || p->initialize_graph) return TASI_ERR_ARG;
    return tasi_read_window(p, 0, 0, p->EX, p->EY, host_out);
}

int tasi_field_device(tasi_planner *p, float **dev_ptr, int64_t *pitch_elems) {
    if (p && p->shard) return sh_fail(p->shard, TASI_ERR_STATE, "field_device: a sharded field has no single device view");
    if (!p || !p->g_base) return TASI_ERR_ARG;
    if (dev_ptr) *dev_ptr = g_origin(p);
    if (pitch_elems) *pitch_elems = p->gpitch;
    return TASI_OK;
}

int tasi_cost_device(tasi_planner *p, uint8_t **dev_ptr, int64_t *pitch_bytes) {
    if (p && p->shard) return sh_fail(p->shard, TASI_ERR_STATE, "cost_device: a sharded grid has no single device view");
    if (!p || !p->c_base) return TASI_ERR_ARG;
    if (dev_ptr) *dev_ptr = c_origin(p);
    if (pitch_bytes) *pitch_bytes = p->cpitch;
    return TASI_OK;
}

int tasi_set_solver_blocks(tasi_planner *p, int max_blocks) {
    if (!p || max_blocks < 0) return TASI_ERR_ARG;
    if (p->shard) p->shard->max_blocks = max_blocks;
    p->max_blocks = max_blocks;
    return TASI_OK;
}

int tasi_get_stats(tasi_planner *p, tasi_stats *out) {
    if (!p || !out) return TASI_ERR_ARG;
    if (p->shard) {
        *out = p->shard->stats;
        return TASI_OK;
    }
    *out = p->stats;
    return TASI_OK;
}

const char *tasi_last_error(tasi_planner *p) {
    if (!p) return g_create_err.c_str();  // no handle: why the last tasi_create* of this thread failed
    return p->shard ? p->shard->err.c_str() : p->err.c_str();
}

// ---- row bands -------------------------------------------------------------------------------------------------
int tasi_band_config(tasi_planner *p, int64_t global_row0, int open_top, int open_bottom) {
    if (!p) return TASI_ERR_ARG;
    if (p && p->shard) return sh_fail(p->shard, TASI_ERR_STATE, "tasi_band_config: a sharded planner runs its bands itself");
    if (!p->initialize_graph) { p->err = "band_config must precede set_map"; return TASI_ERR_STATE; }
    p->global_row0 = global_row0;
    p->open_top = open_top != 0;
    p->open_bottom = open_bottom != 0;
    return TASI_OK;
}

int tasi_band_rows(tasi_planner *p, int side, float **send_row, float **recv_stage, int64_t *count) {
    if (p && p->shard) return sh_fail(p->shard, TASI_ERR_STATE, "tasi_band_rows: a sharded planner runs its bands itself");
    if (!p || !p->g_base || side < 0 || side > 1) return TASI_ERR_ARG;
    if (send_row) *send_row = g_origin(p) + (side == 0 ? 0 : (int64_t) (p->EX - 1) * p->gpitch);
    if (recv_stage) *recv_stage = p->stage + (size_t) side * p->EY;
    if (count) *count = p->EY;
    return TASI_OK;
}

int tasi_band_import(tasi_planner *p, int side, uint64_t *n_lowered) {
    if (p && p->shard) return sh_fail(p->shard, TASI_ERR_STATE, "tasi_band_import: a sharded planner runs its bands itself");
    if (!p || !p->g_base || side < 0 || side > 1) return TASI_ERR_ARG;
    CU_TRY(p, cudaSetDevice(p->device));
    CU_TRY(p, cudaMemsetAsync(&p->ctl->halo_lowered, 0, sizeof(unsigned long long), p->stream));
    GridDesc d = make_desc(p);
    band_import_kernel<<<(p->EY + 255) / 256, 256, 0, p->stream>>>(d, p->stage + (size_t) side * p->EY,
                                                                 side == 0 ? -1 : p->EX, side == 0 ? 0 : p->TX - 1);
    CU_TRY(p, cudaGetLastError());
    DevCtl hc;
    int rc = read_ctl(p, &hc);
    if (rc != TASI_OK) return rc;
    if (n_lowered) *n_lowered = hc.halo_lowered;
    return TASI_OK;
}

int tasi_band_begin(tasi_planner *p) {
    if (!p) return TASI_ERR_ARG;
    if (p && p->shard) return sh_fail(p->shard, TASI_ERR_STATE, "tasi_band_begin: a sharded planner runs its bands itself");
    if (p->initialize_graph) return TASI_LOOP_FAILURE_NO_GRAPH;
    if (!p->goal_set) return TASI_LOOP_FAILURE_NO_GOAL;
    CU_TRY(p, cudaSetDevice(p->device));
    memset(&p->stats, 0, sizeof p->stats);
    p->stats.tiles = (uint64_t) p->TX * p->TY;
    p->stats.solver_blocks = p->solver_blocks;
    int rc = init_field(p);
    if (rc != TASI_OK) return rc;
    CU_TRY(p, cudaStreamSynchronize(p->stream));
    p->new_goal = p->initialize_search = false;
    p->band_started = false;
    return TASI_OK;
}

int tasi_band_solve_until(tasi_planner *p, float limit, uint64_t *n_tile_visits, float *next_key) {
    if (p && p->shard) return sh_fail(p->shard, TASI_ERR_STATE, "tasi_band_solve_until: a sharded planner runs its bands itself");
    if (!p || !p->g_base) return TASI_ERR_ARG;
    CU_TRY(p, cudaSetDevice(p->device));
    const uint64_t before = p->stats.tile_visits;
    float ms = 0;
    int rc = run_solver(p, 0, &ms, p->band_started, limit);
    if (rc != TASI_OK) return rc;
    p->band_started = true;
    p->stats.kernel_ms_solve += ms;
    if (n_tile_visits) *n_tile_visits = p->stats.tile_visits - before;
    if (next_key) memcpy(next_key, &p->sv_far_min_bits, sizeof(float));
    return TASI_OK;
}

int tasi_band_solve(tasi_planner *p, uint64_t *n_tile_visits) {
    return tasi_band_solve_until(p, INFINITY, n_tile_visits, nullptr);
}

int tasi_band_delta(tasi_planner *p, float *delta) {
    if (p && p->shard) return sh_fail(p->shard, TASI_ERR_STATE, "tasi_band_delta: a sharded planner runs its bands itself");
    if (!p || !p->g_base || !delta) return TASI_ERR_ARG;
    CU_TRY(p, cudaSetDevice(p->device));
    int rc = ensure_cost_stats(p);
    if (rc != TASI_OK) return rc;
    *delta = band_width(p);
    return TASI_OK;
}

// The walks of tasi_extract_paths over the field view `v` (L x W cells), on planner p's stream with p's staging block.
static int run_walks(tasi_planner *p, const FieldView &v, float def_sx, float def_sy, float def_gx, float def_gy,
                     const float *starts_xy, const float *goals_xy, const int *row_ranges, int n, int max_steps,
                     int lookahead, int allow_indirect, float *xy, int cap_points, int *n_points, float *step_costs,
                     int cap_costs, int *n_costs, float *totals, int *status) {
    std::vector<PathQuery> q((size_t) n);
    for (int i = 0; i < n; ++i) {
        q[i].sx = starts_xy ? starts_xy[2 * i] : def_sx;
        q[i].sy = starts_xy ? starts_xy[2 * i + 1] : def_sy;
        q[i].gx = goals_xy ? goals_xy[2 * i] : def_gx;
        q[i].gy = goals_xy ? goals_xy[2 * i + 1] : def_gy;
        q[i].row0 = row_ranges ? row_ranges[2 * i] : 0;
        q[i].rows = row_ranges ? row_ranges[2 * i + 1] : 0;
        if (row_ranges && (q[i].row0 < 0 || q[i].rows < 1 || q[i].row0 + q[i].rows > v.L))
            return fail(p, TASI_ERR_ARG, "extract_paths: row range outside the map");
        const int rows = row_ranges ? q[i].rows : v.L;
        // a walk starts inside the map (Graph::is_valid(Position), Graph.cpp:57-59)
        if (!(q[i].sx >= 0.0f && q[i].sy >= 0.0f && q[i].sx <= (float) rows && q[i].sy <= (float) v.W))
            return fail(p, TASI_ERR_ARG, "extract_paths: start outside the map");
    }
    // one device block: queries | n_points, n_costs, status | totals | xy | piece costs; everything after the
    // queries comes back in ONE copy into a pinned staging block and is scattered to the caller's arrays here
    const size_t capc = step_costs ? (size_t) cap_costs : 0;
    const size_t o_int = sizeof(PathQuery) * n, o_tot = o_int + sizeof(int) * 3 * n, o_xy = o_tot + sizeof(float) * 2 * n,
                 o_sc = o_xy + sizeof(float) * 2 * cap_points * n, bytes = o_sc + sizeof(float) * capc * n;
    if (bytes > p->walk_cap) {
        cudaFree(p->walk_buf);
        cudaFreeHost(p->walk_host);
        p->walk_buf = p->walk_host = nullptr;
        p->walk_cap = 0;
        CU_TRY(p, cudaMalloc(&p->walk_buf, bytes));
        CU_TRY(p, cudaHostAlloc(&p->walk_host, bytes, cudaHostAllocDefault));
        p->walk_cap = bytes;
    }
    uint8_t *b = p->walk_buf, *hb = p->walk_host;
    if (n > 1) {
        memcpy(hb, q.data(), sizeof(PathQuery) * n);
        CU_TRY(p, cudaMemcpyAsync(b, hb, sizeof(PathQuery) * n, cudaMemcpyHostToDevice, p->stream));
    }
    int *ints = (int *) (b + o_int);
    const int warps_per_block = 4;
    extract_paths_kernel<<<(n + warps_per_block - 1) / warps_per_block, 32 * warps_per_block, 0, p->stream>>>(
        v, n > 1 ? (const PathQuery *) b : nullptr, q[0], n, max_steps, lookahead, allow_indirect, (float *) (b + o_xy),
        cap_points, step_costs ? (float *) (b + o_sc) : nullptr, (int) capc, ints, ints + n, (float *) (b + o_tot),
        ints + 2 * n);
    CU_TRY(p, cudaGetLastError());
    CU_TRY(p, cudaMemcpyAsync(hb + o_int, b + o_int, bytes - o_int, cudaMemcpyDeviceToHost, p->stream));
    CU_TRY(p, cudaStreamSynchronize(p->stream));
    const int *hi = (const int *) (hb + o_int);
    memcpy(n_points, hi, sizeof(int) * n);
    if (step_costs) memcpy(n_costs, hi + n, sizeof(int) * n);
    memcpy(status, hi + 2 * n, sizeof(int) * n);
    memcpy(totals, hb + o_tot, sizeof(float) * 2 * n);
    memcpy(xy, hb + o_xy, sizeof(float) * 2 * cap_points * n);
    if (step_costs) memcpy(step_costs, hb + o_sc, sizeof(float) * capc * n);
    return TASI_OK;
}

static bool walk_args_ok(int n, const float *starts_xy, const float *goals_xy, int max_steps, float *xy, int cap_points,
                         int *n_points, float *step_costs, int cap_costs, int *n_costs, float *totals, int *status) {
    return !(n <= 0 || !xy || cap_points < 1 || !n_points || !totals || !status || max_steps < 1 ||
             (step_costs && (cap_costs < 1 || !n_costs)) || ((!starts_xy || !goals_xy) && n != 1));
}

int tasi_extract_paths(tasi_planner *p, const float *starts_xy, const float *goals_xy, const int *row_ranges, int n,
                       int max_steps, int lookahead, int allow_indirect, float *xy, int cap_points, int *n_points,
                       float *step_costs, int cap_costs, int *n_costs, float *totals, int *status, float *e_ms) {
    if (!p || !walk_args_ok(n, starts_xy, goals_xy, max_steps, xy, cap_points, n_points, step_costs, cap_costs, n_costs,
                            totals, status))
        return TASI_ERR_ARG;
    if (p->shard)
        return shard_extract_paths(p->shard, starts_xy, goals_xy, row_ranges, n, max_steps, lookahead, allow_indirect, xy,
                                   cap_points, n_points, step_costs, cap_costs, n_costs, totals, status, e_ms);
    if (p->initialize_graph || p->initialize_search) return fail(p, TASI_ERR_STATE, "extract_paths: no plan yet");
    if (p->open_top || p->open_bottom) return fail(p, TASI_ERR_STATE, "extract_paths: not available on a row band");
    CU_TRY(p, cudaSetDevice(p->device));
    const double t0 = now_ms();
    FieldView v = {g_origin(p), p->gpitch, c_origin(p), p->cpitch, p->rows, p->W, p->thr, p->kind == TASI_DFM};
    const int rc = run_walks(p, v, p->start_x, p->start_y, p->goal_x, p->goal_y, starts_xy, goals_xy, row_ranges, n,
                             max_steps, lookahead, allow_indirect, xy, cap_points, n_points, step_costs, cap_costs, n_costs,
                             totals, status);
    if (e_ms) *e_ms = (float) (now_ms() - t0);
    return rc;
}

int tasi_set_focus(tasi_planner *p, int enabled, float slack) {
    if (!p) return TASI_ERR_ARG;
    if (p->shard && enabled) return sh_fail(p->shard, TASI_ERR_STATE, "set_focus: not available on a sharded planner");
    p->focus = enabled != 0;
    p->focus_slack = slack >= 0.0f ? slack : 4.0f * 254.0f;
    return TASI_OK;
}

int tasi_read_parents(tasi_planner *p, int x0, int y0, int nx, int ny, signed char *out) {
    if (p && p->shard) return sh_fail(p->shard, TASI_ERR_STATE, "read_parents: not available on a sharded planner");
    if (!p || !out || nx <= 0 || ny <= 0 || p->initialize_graph) return TASI_ERR_ARG;
    if (x0 < 0 || y0 < 0 || x0 + nx > p->EX || y0 + ny > p->EY) return fail(p, TASI_ERR_ARG, "read_parents: window outside the field");
    if (p->initialize_search) return fail(p, TASI_ERR_STATE, "read_parents: no plan yet");
    CU_TRY(p, cudaSetDevice(p->device));
    const size_t bytes = (size_t) nx * ny * sizeof(Parents);
    if (bytes > p->walk_cap) {  // shares the staging block of tasi_extract_paths
        cudaFree(p->walk_buf);
        cudaFreeHost(p->walk_host);
        p->walk_buf = p->walk_host = nullptr;
        p->walk_cap = 0;
        CU_TRY(p, cudaMalloc(&p->walk_buf, bytes));
        CU_TRY(p, cudaHostAlloc(&p->walk_host, bytes, cudaHostAllocDefault));
        p->walk_cap = bytes;
    }
    const int blocks = (int) std::min<int64_t>(((int64_t) nx * ny + 255) / 256, (int64_t) p->num_sms * 16);
    Parents *dev = (Parents *) p->walk_buf;
    if (p->kind == TASI_FD)
        parents_kernel<TASI_KIND_FD><<<blocks, 256, 0, p->stream>>>(g_origin(p), p->gpitch, c_origin(p), p->cpitch, p->thr, x0, y0, nx, ny, dev);
    else if (p->kind == TASI_DFM)
        parents_kernel<TASI_KIND_DFM><<<blocks, 256, 0, p->stream>>>(g_origin(p), p->gpitch, c_origin(p), p->cpitch, p->thr, x0, y0, nx, ny, dev);
    else
        parents_kernel<TASI_KIND_MFD><<<blocks, 256, 0, p->stream>>>(g_origin(p), p->gpitch, c_origin(p), p->cpitch, p->thr, x0, y0, nx, ny, dev);
    CU_TRY(p, cudaGetLastError());
    CU_TRY(p, cudaMemcpyAsync(out, dev, bytes, cudaMemcpyDeviceToHost, p->stream));
    CU_TRY(p, cudaStreamSynchronize(p->stream));
    return TASI_OK;
}


}  // extern "C"

// ---- sharded planner (tasi_shard.cuh) -------------------------------------------------------------------------------
namespace {

int shard_extract_paths(tasi_shard *s, const float *starts_xy, const float *goals_xy, const int *row_ranges, int n,
                        int max_steps, int lookahead, int allow_indirect, float *xy, int cap_points, int *n_points,
                        float *step_costs, int cap_costs, int *n_costs, float *totals, int *status, float *e_ms) {
    if (s->initialize_graph || s->initialize_search) return sh_fail(s, TASI_ERR_STATE, "extract_paths: no plan yet");
    if (row_ranges) return sh_fail(s, TASI_ERR_STATE, "extract_paths: row ranges are for map stacks, not for a sharded grid");
    const double t0 = now_ms();
    ShardMember &m0 = s->local[0];
    tasi_planner *p = m0.p;
    const int reach = 3 * max_steps + 8;  // a step crosses at most a few cell edges
    for (int i = 0; i < n; ++i) {
        const float sx = starts_xy ? starts_xy[2 * i] : s->start_x, sy = starts_xy ? starts_xy[2 * i + 1] : s->start_y;
        const float gx = goals_xy ? goals_xy[2 * i] : s->goal_x, gy = goals_xy ? goals_xy[2 * i + 1] : s->goal_y;
        const int64_t lo = std::max<int64_t>((int64_t) floorf(sx) - reach, 0);
        const int64_t hi = std::min<int64_t>((int64_t) ceilf(sx) + reach + 1, s->n_elem_rows);
        const int64_t clo = std::max<int64_t>(lo - 1, 0), chi = std::min<int64_t>(hi, s->L);
        if (hi <= lo) return sh_fail(s, TASI_ERR_ARG, "extract_paths: start outside the map");
        // window of value rows [lo, hi) and cost rows [clo, chi) on the first local device
        SH_TRY(s, cudaSetDevice(m0.device));
        const int64_t gp = (s->EY + 3) / 4 * 4, cp = (s->W + 15) / 16 * 16;
        float *gw = nullptr;
        uint8_t *cw = nullptr;
        SH_TRY(s, cudaMalloc(&gw, sizeof(float) * gp * (hi - lo)));
        if (cudaMalloc(&cw, (size_t) cp * std::max<int64_t>(chi - clo, 1)) != cudaSuccess) {
            cudaFree(gw);
            return sh_fail(s, TASI_ERR_CUDA, "extract_paths: out of device memory for the walk window");
        }
        int64_t missing = 0;
        int rc = shard_gather_rows(s, lo, hi - lo, 0, s->EY, gw, gp, cudaMemcpyDefault, &missing);
        int64_t have_c = 0;
        if (rc == TASI_OK)
            for (ShardMember &m : s->local) {
                cudaSetDevice(m.device);
                for (int g : m.bands) {
                    const ShardBand &b = s->bands[g];
                    const int64_t a0 = std::max(b.a, clo), a1 = std::min(std::min<int64_t>(b.b, s->L), chi);
                    if (a1 <= a0) continue;
                    if (cudaMemcpy2DAsync(cw + (a0 - clo) * cp, cp, c_origin(m.p) + (int64_t) (b.row0 + (a0 - b.a)) * m.p->cpitch,
                                          m.p->cpitch, s->W, a1 - a0, cudaMemcpyDefault, m.p->stream) != cudaSuccess)
                        rc = TASI_ERR_CUDA;
                    have_c += a1 - a0;
                }
            }
        if (rc == TASI_OK) rc = shard_sync_all(s);
        if (rc == TASI_OK && (missing || have_c != chi - clo))
            rc = sh_fail(s, TASI_ERR_STATE, "extract_paths: the rows around the start are owned by another process");
        if (rc == TASI_OK) {
            cudaSetDevice(m0.device);
            // the view's origin is element (0,0) of the WHOLE grid: only rows [lo, hi) of it are backed by memory
            FieldView v = {gw - lo * gp, gp, cw - clo * cp, cp, s->L, s->W, s->thr, s->kind == TASI_DFM};
            const float st[2] = {sx, sy}, gl[2] = {gx, gy};
            rc = run_walks(p, v, sx, sy, gx, gy, st, gl, nullptr, 1, max_steps, lookahead, allow_indirect,
                           xy + (size_t) i * cap_points * 2, cap_points, n_points + i,
                           step_costs ? step_costs + (size_t) i * cap_costs : nullptr, cap_costs, n_costs ? n_costs + i : nullptr,
                           totals + 2 * i, status + i);
            if (rc != TASI_OK) s->err = p->err;
        }
        cudaSetDevice(m0.device);
        cudaFree(gw);
        cudaFree(cw);
        if (rc != TASI_OK) return rc;
    }
    if (e_ms) *e_ms = (float) (now_ms() - t0);
    return TASI_OK;
}

int shard_finish_create(tasi_shard *s, int kind, int level, tasi_planner **out) {
    tasi_planner *top = new (std::nothrow) tasi_planner();
    if (!top) { shard_destroy(s); return TASI_ERR_ARG; }
    top->kind = kind; top->level = level; top->device = s->local[0].device;
    top->shard = s;
    if (const char *e = getenv("TASI_BANDS_PER_GPU")) {
        const int v = atoi(e);
        if (v >= 1 && v <= 64) s->want_per_rank = v;
    }
    if (const char *e = getenv("TASI_BAND_PHASES")) {
        const int v = atoi(e);
        if (v >= 1 && v <= 64) s->k_phases = v;
    }
    if (const char *e = getenv("TASI_SHARD_FUSED")) s->want_fused = atoi(e) != 0;
    *out = top;
    return TASI_OK;
}

}  // namespace

extern "C" {

int tasi_create_sharded(int kind, int level, const int *device_ids, int n_dev, tasi_planner **out) {
    if (!out) return TASI_ERR_ARG;
    *out = nullptr;
    if (!device_ids || n_dev < 1 || n_dev > SHARD_MAX_RANKS || kind < TASI_FD || kind > TASI_MFD || level < 0 ||
        level > (kind == TASI_MFD ? 2 : 1))
        return TASI_ERR_ARG;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) return TASI_ERR_CUDA;
    for (int i = 0; i < n_dev; ++i)
        if (device_ids[i] < 0 || device_ids[i] >= ndev) return TASI_ERR_CUDA;
    tasi_shard *s = new (std::nothrow) tasi_shard();
    if (!s) return TASI_ERR_ARG;
    s->kind = kind; s->level = level; s->world = n_dev; s->my_rank = -1;
    s->local.resize(n_dev);
    bool distinct = true;
    for (int i = 0; i < n_dev; ++i) {
        s->local[i].rank = i;
        s->local[i].device = device_ids[i];
        for (int j = 0; j < i; ++j) distinct = distinct && device_ids[j] != device_ids[i];
    }
    for (ShardMember &m : s->local)
        if (shard_init_member(s, m) != TASI_OK) { g_create_err = s->err; shard_destroy(s); return TASI_ERR_CUDA; }
    // ranks of one process read each other's buffers directly: map every pair of distinct devices
    bool peers = true;
    for (int i = 0; i < n_dev && peers; ++i)
        for (int j = 0; j < n_dev && peers; ++j) {
            if (device_ids[i] == device_ids[j]) continue;
            int can = 0;
            cudaDeviceCanAccessPeer(&can, device_ids[i], device_ids[j]);
            if (!can) { peers = false; break; }
            cudaSetDevice(device_ids[i]);
            const cudaError_t e = cudaDeviceEnablePeerAccess(device_ids[j], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) peers = false;
            cudaGetLastError();
        }
    const char *want = getenv("TASI_SHARD_NCCL");
    s->use_nccl = n_dev > 1 && distinct && ((want && atoi(want) != 0) || !peers);
    if (n_dev > 1 && !peers && !distinct) {
        g_create_err = "create_sharded: repeated devices without peer access";
        shard_destroy(s);
        return TASI_ERR_CUDA;
    }
    if (s->use_nccl) {
        NcclApi &N = nccl_api();
        std::vector<ncclComm_t> comms(n_dev);
        ncclResult_t r = ncclSuccess;
        if (!N.load() || (r = N.CommInitAll(comms.data(), n_dev, device_ids)) != ncclSuccess) {
            g_create_err = N.handle ? std::string("ncclCommInitAll: ") + N.GetErrorString(r) : N.err;
            shard_destroy(s);
            return TASI_ERR_CUDA;
        }
        for (int i = 0; i < n_dev; ++i) s->local[i].comm = comms[i];
    }
    return shard_finish_create(s, kind, level, out);
}

int tasi_nccl_unique_id(void *out128) {
    if (!out128) return TASI_ERR_ARG;
    NcclApi &N = nccl_api();
    if (!N.load()) { g_create_err = N.err; return TASI_ERR_CUDA; }
    ncclUniqueId id;
    if (N.GetUniqueId(&id) != ncclSuccess) { g_create_err = "ncclGetUniqueId failed"; return TASI_ERR_CUDA; }
    memcpy(out128, &id, sizeof id);
    return TASI_OK;
}

int tasi_create_sharded_rank(int kind, int level, int device, int rank, int n_ranks, const void *nccl_unique_id,
                             tasi_planner **out) {
    if (!out) return TASI_ERR_ARG;
    *out = nullptr;
    if (n_ranks < 1 || n_ranks > SHARD_MAX_RANKS || rank < 0 || rank >= n_ranks || (n_ranks > 1 && !nccl_unique_id) ||
        kind < TASI_FD || kind > TASI_MFD || level < 0 || level > (kind == TASI_MFD ? 2 : 1))
        return TASI_ERR_ARG;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) return TASI_ERR_CUDA;
    tasi_shard *s = new (std::nothrow) tasi_shard();
    if (!s) return TASI_ERR_ARG;
    s->kind = kind; s->level = level; s->world = n_ranks; s->my_rank = rank;
    s->local.resize(1);
    s->local[0].rank = rank;
    s->local[0].device = device;
    if (shard_init_member(s, s->local[0]) != TASI_OK) { g_create_err = s->err; shard_destroy(s); return TASI_ERR_CUDA; }
    s->use_nccl = n_ranks > 1;
    if (s->use_nccl) {
        NcclApi &N = nccl_api();
        ncclUniqueId id;
        memcpy(&id, nccl_unique_id, sizeof id);
        cudaSetDevice(device);
        ncclResult_t r = ncclSuccess;
        if (!N.load() || (r = N.CommInitRank(&s->local[0].comm, n_ranks, id, rank)) != ncclSuccess) {
            g_create_err = N.handle ? std::string("ncclCommInitRank: ") + N.GetErrorString(r) : N.err;
            shard_destroy(s);
            return TASI_ERR_CUDA;
        }
    }
    return shard_finish_create(s, kind, level, out);
}

int tasi_shard_config(tasi_planner *p, int bands_per_device, int phases_per_exchange) {
    if (!p || !p->shard) return TASI_ERR_ARG;
    if (!p->shard->initialize_graph) return sh_fail(p->shard, TASI_ERR_STATE, "shard_config must precede set_map");
    if (bands_per_device >= 1) p->shard->want_per_rank = std::min(bands_per_device, 64);
    if (phases_per_exchange >= 1) p->shard->k_phases = std::min(phases_per_exchange, 64);
    return TASI_OK;
}

int tasi_shard_set_fused(tasi_planner *p, int enabled) {
    if (!p || !p->shard) return TASI_ERR_ARG;
    tasi_shard *s = p->shard;
    s->want_fused = enabled != 0;
    if (!s->want_fused) {
        s->fused = false;
        for (ShardMember &m : s->local) m.p->ft = nullptr;
        return TASI_OK;
    }
    return s->initialize_graph || sh_multi_process(s) ? TASI_OK : shard_connect_local(s);
}

namespace {
struct ShardIpcBlob {
    cudaIpcMemHandle_t g, state, key, queue, ctl, fc, inbox, bq;
    int64_t gpitch;
    int qmask, far_cap, TY, inbox_cap, rank, valid, bqmask, has_bq;
};
}  // namespace

int tasi_shard_ipc_size(void) { return (int) ((sizeof(ShardIpcBlob) + 63) / 64 * 64); }

int tasi_shard_ipc_export(tasi_planner *p, void *blob) {
    if (!p || !p->shard || !blob) return TASI_ERR_ARG;
    tasi_shard *s = p->shard;
    if (s->initialize_graph || s->local.size() != 1) return sh_fail(s, TASI_ERR_STATE, "ipc_export: needs a map, one rank per process");
    ShardMember &m = s->local[0];
    tasi_planner *q = m.p;
    ShardIpcBlob b;
    memset(&b, 0, sizeof b);
    SH_TRY(s, cudaSetDevice(m.device));
    SH_TRY(s, cudaIpcGetMemHandle(&b.g, q->g_base));
    SH_TRY(s, cudaIpcGetMemHandle(&b.state, q->state));
    SH_TRY(s, cudaIpcGetMemHandle(&b.key, q->key));
    SH_TRY(s, cudaIpcGetMemHandle(&b.queue, q->queue));
    SH_TRY(s, cudaIpcGetMemHandle(&b.ctl, q->ctl));
    SH_TRY(s, cudaIpcGetMemHandle(&b.fc, m.fc));
    SH_TRY(s, cudaIpcGetMemHandle(&b.inbox, q->inbox));
    if (q->bq) {
        SH_TRY(s, cudaIpcGetMemHandle(&b.bq, q->bq));
        b.has_bq = 1;
        b.bqmask = (int) (q->bqsize - 1);
    }
    b.gpitch = q->gpitch; b.qmask = (int) (q->qsize - 1); b.far_cap = q->far_cap; b.TY = q->TY;
    b.inbox_cap = q->inbox_cap; b.rank = m.rank; b.valid = 1;
    memset(blob, 0, tasi_shard_ipc_size());
    memcpy(blob, &b, sizeof b);
    return TASI_OK;
}

int tasi_shard_ipc_import(tasi_planner *p, const void *blobs_by_rank) {
    if (!p || !p->shard || !blobs_by_rank) return TASI_ERR_ARG;
    tasi_shard *s = p->shard;
    if (s->initialize_graph || s->local.size() != 1) return sh_fail(s, TASI_ERR_STATE, "ipc_import: needs a map, one rank per process");
    s->fused = false;
    ShardMember &m = s->local[0];
    m.p->ft = nullptr;
    if (!s->want_fused || s->world == 1) return s->world == 1 ? shard_connect_local(s) : TASI_OK;
    SH_TRY(s, cudaSetDevice(m.device));
    shard_close_ipc(m);
    const size_t stride = tasi_shard_ipc_size();
    auto blob_of = [&](int r) {
        ShardIpcBlob b;
        memcpy(&b, (const char *) blobs_by_rank + (size_t) r * stride, sizeof b);
        return b;
    };
    auto open = [&](const cudaIpcMemHandle_t &h, void **out) -> int {
        SH_TRY(s, cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
        m.ipc_open.push_back(*out);
        return TASI_OK;
    };
    FusedCtl *fcs[FUSED_MAX_RANKS] = {};
    for (int r = 0; r < s->world; ++r) {
        if (r == m.rank) { fcs[r] = m.fc; continue; }
        const ShardIpcBlob b = blob_of(r);
        if (!b.valid || b.rank != r) return sh_fail(s, TASI_ERR_ARG, "ipc_import: blob " + std::to_string(r) + " is not rank " + std::to_string(r) + "'s");
        void *q = nullptr;
        int rc = open(b.fc, &q);
        if (rc != TASI_OK) return rc;
        fcs[r] = (FusedCtl *) q;
    }
    GridDesc nb[2];
    const int prev = (m.rank + s->world - 1) % s->world, next = (m.rank + 1) % s->world;
    for (int k = 0; k < 2; ++k) {
        const int r = k == 0 ? prev : next;
        if (k == 1 && next == prev) { nb[1] = nb[0]; break; }
        const ShardIpcBlob b = blob_of(r);
        GridDesc d = make_desc(m.p);  // same geometry conventions on every rank; the pointers are replaced below
        void *g = nullptr, *st = nullptr, *ky = nullptr, *qu = nullptr, *ct = nullptr, *ib = nullptr;
        int rc;
        if ((rc = open(b.g, &g)) != TASI_OK || (rc = open(b.state, &st)) != TASI_OK || (rc = open(b.key, &ky)) != TASI_OK ||
            (rc = open(b.queue, &qu)) != TASI_OK || (rc = open(b.ctl, &ct)) != TASI_OK || (rc = open(b.inbox, &ib)) != TASI_OK)
            return rc;
        d.g = (float *) g + b.gpitch + GMARGIN;
        d.gpitch = b.gpitch;
        d.cost = nullptr;
        d.state = (int *) st; d.key = (int *) ky; d.queue = (int *) qu; d.qmask = (unsigned) b.qmask;
        d.far = nullptr; d.far_cap = b.far_cap;
        d.ctl = (DevCtl *) ct;
        d.inbox = (int *) ib; d.inbox_cap = b.inbox_cap;
        d.bq = nullptr; d.bqmask = 0;
        if (b.has_bq) {
            void *bqp = nullptr;
            if ((rc = open(b.bq, &bqp)) != TASI_OK) return rc;
            d.bq = (int *) bqp; d.bqmask = (unsigned) b.bqmask;
        }
        d.TY = b.TY;
        d.seed = nullptr; d.xgoals = nullptr;
        nb[k] = d;
    }
    int rc = shard_upload_topo(s, m, nb[0], nb[1], fcs);
    if (rc != TASI_OK) return rc;
    s->fused = true;
    return TASI_OK;
}

int tasi_shard_get_stats(tasi_planner *p, tasi_shard_stats *out) {
    if (!p || !p->shard || !out) return TASI_ERR_ARG;
    *out = p->shard->sstats;
    return TASI_OK;
}

int tasi_field_checksum(tasi_planner *p, uint64_t *sum) {
    if (!p || !sum) return TASI_ERR_ARG;
    if (p->shard) return shard_checksum(p->shard, sum);
    if (!p->g_base || p->initialize_graph) return TASI_ERR_ARG;
    CU_TRY(p, cudaSetDevice(p->device));
    if (!p->cost_sum) CU_TRY(p, cudaMalloc(&p->cost_sum, 2 * sizeof(unsigned long long)));
    unsigned long long *acc = p->cost_sum;  // scratch: the cost statistics are consumed on the host as soon as computed
    unsigned long long h = 0;
    CU_TRY(p, cudaMemsetAsync(acc, 0, sizeof h, p->stream));
    field_checksum_kernel<<<p->num_sms * 4, 256, 0, p->stream>>>(g_origin(p), p->gpitch, p->EX, p->EY, p->global_row0, acc);
    CU_TRY(p, cudaGetLastError());
    CU_TRY(p, cudaMemcpyAsync(&h, acc, sizeof h, cudaMemcpyDeviceToHost, p->stream));
    CU_TRY(p, cudaStreamSynchronize(p->stream));
    *sum = h;
    return TASI_OK;
}

}  // extern "C"
