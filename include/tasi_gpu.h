/* tasi_gpu.h — C ABI of the B200 cost-propagation core for the three grid replanners of
 * roncapat/UNIGE-TASI-path-planners (FD*, MS-DFM, MFD*).
 *
 * The reference has no FFI layer: its operator API for this path is the public surface of
 * ReplannerBase<Derived,Elem,Info,Key> (ProjectToolkit/include/ReplannerBase.h:29-161). Each entry point
 * below replaces one member of that class (cited per function); the C++ classes in
 * unige_tasi_path_planners_b200/host/GpuPlanners.h rebuild the class surface on top of it.
 *
 * Conventions (ProjectToolkit/Graph.cpp:22-34): the map is `length` rows (x, north to south) by `width`
 * columns (y) of uint8 traversal costs, map[x*width + y]; a byte >= the occupancy threshold is an obstacle.
 * FD* and MFD* values live on the (length+1) x (width+1) nodes, MS-DFM values on the length x width cells;
 * "element" means node or cell accordingly. Values are float32, +INF = unreachable.
 *
 * All functions return an int status: TASI_OK, the reference's step() codes, or a negative error; no C++
 * exception crosses this boundary. One handle owns one CUDA stream on one device; calls on a handle must be
 * serialised by the caller, different handles are independent. There is no CPU fallback: without a CUDA
 * device tasi_create fails with TASI_ERR_CUDA.
 */
#ifndef TASI_GPU_H
#define TASI_GPU_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TASI_OK 0
#define TASI_LOOP_FAILURE_NO_GRAPH (-1) /* ReplannerBase.h:23 */
#define TASI_LOOP_FAILURE_NO_GOAL (-2)  /* ReplannerBase.h:24 */
#define TASI_ERR_ARG (-3)
#define TASI_ERR_CUDA (-4)
#define TASI_ERR_TIMEOUT (-5) /* the solver's device watchdog fired */
#define TASI_ERR_STATE (-6)

#define TASI_FD 0  /* FieldDPlanner<O>        FieldDStar/FieldDPlanner.h */
#define TASI_DFM 1 /* DFMPlanner<O>           DynamicFastMarching/DynamicFastMarching.h */
#define TASI_MFD 2 /* ShiftedGridPlanner<O>   ShiftedGridFastMarching/ShiftedGridPlanner.h */

typedef struct tasi_planner tasi_planner;

/* Work counters of the last step() (the tile-revisit evidence BASELINE.json asks for). */
typedef struct tasi_stats {
    uint64_t tiles;            /* 32x32 tiles of the element grid */
    uint64_t tile_visits;      /* tiles staged into shared memory (solve phase) */
    uint64_t tile_rounds;      /* in-tile relaxation iterations (one pass over the tile's dirty elements) */
    uint64_t elem_evals;       /* element evaluations (local updates computed) in the solve phase */
    uint64_t elem_lowered;     /* accepted decreases of an element value */
    uint64_t queue_pushes;     /* tile activations that entered the work queue */
    uint64_t inval_visits;     /* tile visits of the invalidate phase of a replan */
    uint64_t elem_invalidated; /* elements reset to INF by the invalidate phase */
    uint64_t seed_elems;       /* elements seeded by the patches pending at this step (== num_nodes_updated). Deviation
                                * from Graph::update (Graph.cpp:36-51), which clears updated_cells_ on every patch and so
                                * seeds the LAST patch only: every patch since the last consumed replan is seeded, and
                                * the counter is the sum over them (an element touched by two patches counts twice) */
    uint64_t phases;           /* advances of the near/far threshold (tile-level Delta-stepping) */
    uint64_t far_parked;       /* activations parked in the far pile */
    float kernel_ms_solve;     /* CUDA-event time of the solver kernel(s), re-propagation phase */
    float kernel_ms_inval;     /* CUDA-event time of the invalidate phase */
    float kernel_ms_init;      /* field fill / seeding kernels */
    float step_ms_device;      /* CUDA-event time of the whole step() on the planner's stream */
    int32_t solver_blocks;     /* persistent CTAs launched */
    int32_t solver_launches;   /* kernels launched by the last step() */
    /* where the persistent CTAs spent their time in the solve phase (SM cycles summed over CTAs): waiting for a
     * tile, staging it, relaxing it, writing back + activating neighbours */
    uint64_t cyc_pop, cyc_stage, cyc_relax, cyc_post;
    uint64_t cyc_advance;      /* cycles one CTA spent advancing the threshold while all others waited */
    float settled_below;       /* every field value below this is final; +INF = the whole field is (no focus) */
    int focus_hit;             /* the last plan ended because the start was settled (tasi_set_focus) */
} tasi_stats;

/* Planner object. kind = TASI_FD|TASI_DFM|TASI_MFD; level = the reference's OPT_LVL template argument
 * (0/1, MFD* also 2): all levels of a planner have the same fixed point, so they share one kernel.
 * Replaces: `FieldDPlanner<OPT_LVL> planner{}` (Tests/Planners/FDSTAR/main.cpp:77). */
int tasi_create(int kind, int level, int device, tasi_planner **out);
int tasi_destroy(tasi_planner *p);

/* ReplannerBase::reset (ReplannerBase.h:39-41): the next step() re-initialises the search. */
int tasi_reset(tasi_planner *p);

/* ReplannerBase::set_occupancy_threshold -> Graph::set_occupancy_threshold (Graph.cpp:18-20):
 * threshold byte = (unsigned char)(t * 255). Default 254 (Graph.h:33). */
int tasi_set_occupancy_threshold(tasi_planner *p, float t);

/* ReplannerBase::set_heuristic_multiplier (ReplannerBase.h:81-83). Stored; the heuristic only prunes the
 * reference's expansion order and never changes a settled value, so the solver ignores it. */
int tasi_set_heuristic_multiplier(tasi_planner *p, float mult);

/* ReplannerBase::set_map -> Graph::init (ReplannerBase.h:85-88, Graph.cpp:22-29). Copies the host map to the
 * device during the call; the host pointer is not retained. */
int tasi_set_map(tasi_planner *p, const uint8_t *host_map, int width, int length);

/* Same, from a buffer already resident in device memory (row pitch in bytes). */
int tasi_set_map_device(tasi_planner *p, const uint8_t *dev_map, int width, int length, int64_t pitch);

/* ReplannerBase::patch_map -> Graph::update (ReplannerBase.h:90-92, Graph.cpp:36-51): rows x..x+h-1, columns
 * y..y+w-1 take the bytes of `patch` (h rows of w); cells whose byte changed seed the next replan.
 * n_changed (optional) receives the number of changed cells. */
int tasi_patch_map(tasi_planner *p, const uint8_t *host_patch, int x, int y, int w, int h, uint64_t *n_changed);

/* tasi_patch_map for callers that must not allocate. NO COUNTERPART IN THE REFERENCE: added
 * for code under CUDA graph capture, tight serving loops and harnesses that forbid cudaMalloc. Same arguments,
 * same effect on the cost grid, the same seeds for the next replan and the same n_changed as tasi_patch_map, but the two
 * scratch buffers of the patch kernels (the device copy of the patch and the per-cell change flags) live in a file-scope
 * static device array of 2 * TASI_PATCH_STATIC_MAX_CELLS bytes (2 MiB) that every GPU holds from the moment it loads
 * the library until the process exits; the call makes no cudaMalloc / cudaFree / allocating library call.
 * A patch of more than TASI_PATCH_STATIC_MAX_CELLS cells (w * h) is rejected with TASI_ERR_ARG and a message naming the
 * limit (tasi_last_error); a sharded handle with TASI_ERR_STATE. All planners on one GPU share the array and it keeps
 * the last call's contents (what must start clean is cleared by the call itself): calls that overlap on one GPU — two
 * planners on their own streams, two host threads — race on it and must be serialised by the caller. */
#define TASI_PATCH_STATIC_MAX_CELLS (1 << 20)
int tasi_patch_map_static(tasi_planner *p, const uint8_t *host_patch, int x, int y, int w, int h, uint64_t *n_changed);

/* ReplannerBase::set_start / set_goal (ReplannerBase.h:94-108; Graph.cpp:6-16). Positions are rounded to
 * the element with roundf (Node.cpp:14-17, Cell.cpp:20-21). A goal on a different element forces re-init. */
int tasi_set_start(tasi_planner *p, float x, float y);
int tasi_set_goal(tasi_planner *p, float x, float y);

/* Several goals at once (n pairs x,y): every goal element gets value 0 and the field is the cost to the nearest
 * goal. No reference counterpart; it is what lets a batch of independent queries run as ONE plan: the maps are
 * stacked into one tall map with an obstacle row between them (nothing propagates through an obstacle row, and
 * the node rows either side of it are exactly the boundary rows of the two maps), one goal per map, and all the
 * wavefronts share one launch of the persistent solver. Always re-plans from scratch. */
int tasi_set_goals(tasi_planner *p, const float *xy, int n);

/* ReplannerBase::step (ReplannerBase.h:43-75): initial plan, or two-phase repair after patches; returns when
 * the field is converged. Outputs (all optional): u_ms/p_ms = host wall time of the update and plan halves
 * (the reference's u_time/p_time), n_updated/n_expanded = num_nodes_updated/num_nodes_expanded. */
int tasi_step(tasi_planner *p, float *u_ms, float *p_ms, uint64_t *n_updated, uint64_t *n_expanded);

/* Element-grid extents: (length+1, width+1) for FD and MFD, (length, width) for MS-DFM. */
int tasi_field_dims(tasi_planner *p, int *ex, int *ey);

/* Hand-off to the path extractor: what ExpandedMap::get_g / get_rhs return after convergence
 * (ExpandedMap_impl.h:65-85; g == rhs on a converged field). read_window copies elements
 * [x0,x0+nx) x [y0,y0+ny) (row-major into out, INF outside the grid); read_field the whole grid. */
int tasi_read_window(tasi_planner *p, int x0, int y0, int nx, int ny, float *host_out);
int tasi_read_field(tasi_planner *p, float *host_out);

/* Device view of the converged field: pointer to element (0,0) and row pitch in elements. */
int tasi_field_device(tasi_planner *p, float **dev_ptr, int64_t *pitch_elems);
/* Device view of the cost grid: pointer to cell (0,0) and row pitch in bytes. */
int tasi_cost_device(tasi_planner *p, uint8_t **dev_ptr, int64_t *pitch_bytes);

/* Deterministic synthetic map generated on the device (same function as synth.py / SURVEY.md 8d):
 * rows [row0,row0+length) of the unbounded map of width `width`; 9x9 blocks of cost 1 are cleared around
 * (sx,sy) and (gx,gy) given in GLOBAL cell coordinates (negative = none). Then behaves as set_map. */
int tasi_synth_map(tasi_planner *p, uint64_t seed, int width, int length, int64_t row0, int p_obs, int sx, int sy,
                   int gx, int gy);
/* Copies the device cost grid back (length*width bytes). */
int tasi_read_map(tasi_planner *p, uint8_t *host_out);

/* Cap the number of persistent solver CTAs of this planner (0 = as many as fit on the device). A batch of
 * independent queries (BASELINE.json cfg5) runs several planners side by side, each on its own stream with a
 * share of the SMs: a 1024^2 plan never has more than a few dozen active tiles, so a full-device launch per
 * query would leave most CTAs waiting. No reference counterpart (the reference is single-threaded). */
int tasi_set_solver_blocks(tasi_planner *p, int max_blocks);

int tasi_get_stats(tasi_planner *p, tasi_stats *out);
const char *tasi_last_error(tasi_planner *p);

/* ---- sharded planner: one grid over several GPUs behind the same entry points ----
 * The reference has no multi-device path; ReplannerBase's surface (ReplannerBase.h:39-123) stays what the caller sees:
 * the handle returned here takes tasi_set_map / tasi_patch_map / tasi_set_start / tasi_set_goal / tasi_step /
 * tasi_read_window / tasi_read_field / tasi_extract_paths / tasi_get_stats exactly as a single-device handle does, and
 * the converged field is bit-identical to the single-device field. The element rows are cut into bands dealt out
 * cyclically over the devices; every device runs ONE persistent solver over the stack of its bands, boundary rows are
 * exchanged and the common value limit raised on the device streams (csrc/tasi_shard.cuh).
 *
 * tasi_create_sharded: all devices driven by this process (one stream per device, peer-mapped exchange;
 * TASI_SHARD_NCCL=1 or missing peer access: NCCL via ncclCommInitAll). device_ids may repeat a device (several ranks
 * on one GPU: the single-GPU test of the protocol).
 * tasi_create_sharded_rank: one rank per process, e.g. one per GPU under torchrun; rank 0 obtains 128 bytes from
 * tasi_nccl_unique_id and the caller hands them to every rank (broadcast); halo rows then travel by ncclSend/ncclRecv
 * and the convergence quantities by ncclAllReduce on the planner's stream. Every rank must make the same calls in the
 * same order (set_map with the WHOLE map, patch_map, step ...); tasi_read_window / tasi_extract_paths only reach rows
 * this rank owns, tasi_field_checksum covers the whole field.
 * Not available on a sharded handle (TASI_ERR_STATE): set_map_device, set_map_stack, set_goals, set_focus,
 * read_parents, field_device / cost_device, the tasi_band_* primitives. */
int tasi_create_sharded(int kind, int level, const int *device_ids, int n_dev, tasi_planner **out);
int tasi_nccl_unique_id(void *out128);
int tasi_create_sharded_rank(int kind, int level, int device, int rank, int n_ranks, const void *nccl_unique_id,
                             tasi_planner **out);
/* Bands per device (default: bands of about 32 tile rows, 2 to 8 per device; TASI_BANDS_PER_GPU) and threshold widths of the tile schedule between two exchanges of
 * the iteration-based mode (default 1, TASI_BAND_PHASES); values < 1 keep the current setting. Before set_map. */
int tasi_shard_config(tasi_planner *p, int bands_per_device, int phases_per_exchange);
/* Fused mode on (default) / off (TASI_SHARD_FUSED=0): off keeps the iteration-based exchange (NCCL or peer reads). */
int tasi_shard_set_fused(tasi_planner *p, int enabled);
/* Fused mode across processes: after every set_map / synth_map each rank exports tasi_shard_ipc_size() bytes (CUDA IPC
 * handles of its value field, tile states, keys, work queue and control blocks), the caller all-gathers them (by rank)
 * and hands the whole array to every rank. Without it a multi-process sharded planner runs the NCCL iterations. */
int tasi_shard_ipc_size(void);
int tasi_shard_ipc_export(tasi_planner *p, void *blob);
int tasi_shard_ipc_import(tasi_planner *p, const void *blobs_by_rank);

typedef struct tasi_shard_stats {
    int32_t n_ranks, bands, bands_per_rank, rows_per_band;
    int32_t nccl;               /* 1: NCCL send/recv + all-reduce per outer iteration; 0: peer-mapped memory */
    int32_t fused;              /* 1: fused mode — one persistent launch per GPU, boundary rows and tile activations go
                                 * straight into the neighbour's memory during the tile visits, thresholds advance globally */
    uint64_t exchanges;         /* outer iterations of the relax loop (0 in fused mode) */
    uint64_t inval_exchanges;   /* outer iterations of the invalidate loop of a replan (0 in fused mode) */
    uint64_t fused_launches;    /* fused launches of the step (1 per phase of a plan / replan) */
    uint64_t remote_activations; /* tile activations sent to other ranks (fused mode), summed over the ranks */
    float wall_ms;              /* host wall clock of the step */
    float solver_ms_max_rank;   /* time inside the persistent solver launches, busiest and idlest rank */
    float solver_ms_min_rank;
    float busy_frac_max_rank;   /* share of the CTA time spent on tiles (staging, relaxing, writing back) rather than */
    float busy_frac_min_rank;   /* waiting for one, busiest and idlest rank */
    float pad;
} tasi_shard_stats;
int tasi_shard_get_stats(tasi_planner *p, tasi_shard_stats *out);

/* Order-independent 64-bit checksum of the value field (sum over the elements of a mix of global index and value bits):
 * equal for a sharded and a single-device planner iff their fields are bit-identical. On a multi-process sharded handle
 * it is collective (every rank calls it and gets the sum over the whole field). */
int tasi_field_checksum(tasi_planner *p, uint64_t *sum);

/* ---- row-band primitives (one band per handle; the building blocks the sharded planner grew out of; halo rows are exchanged by the caller, e.g. NCCL send/recv) ----
 * A band holds element rows [0,ex) of its own plus one halo row above (row -1) and below (row ex) in the
 * same device array. tasi_band_config marks which sides are open (have a neighbour band) and gives the
 * band's first global element row so start/goal can be given in global coordinates. The band's map must
 * then contain one extra cost row per open side (the cells between the halo row and the first owned row). */
int tasi_band_config(tasi_planner *p, int64_t global_row0, int open_top, int open_bottom);
/* Device pointers of the rows to send (first/last owned row) and to receive into (staging rows);
 * each is `count` floats. side 0 = top, 1 = bottom. */
int tasi_band_rows(tasi_planner *p, int side, float **send_row, float **recv_stage, int64_t *count);
/* Merge the received staging row into the halo (min), activate the border tiles whose halo got lower;
 * n_lowered = how many halo elements changed. */
int tasi_band_import(tasi_planner *p, int side, uint64_t *n_lowered);
/* Initialise the field (INF, goal = 0 if the goal lies in this band) without solving. */
int tasi_band_begin(tasi_planner *p);
/* Drain the local work queue to convergence; n_active_out = tiles processed. */
int tasi_band_solve(tasi_planner *p, uint64_t *n_tile_visits);
/* Same, but the value threshold of the tile schedule never passes `limit`: tiles whose lowest pending value is at
 * or beyond it stay parked (they survive to the next call). next_key = the lowest parked value (+INF: nothing
 * parked). Lets all bands of a grid work on the same band of values between two halo exchanges. */
int tasi_band_solve_until(tasi_planner *p, float limit, uint64_t *n_tile_visits, float *next_key);
/* Width of the value band one threshold advance covers on this band's map (delta_factor * 32 * mean free cost). */
int tasi_band_delta(tasi_planner *p, float *delta);

/* A stack of n_blocks independent maps of `length` rows each in one planner, one obstacle row (255) between
 * consecutive blocks: the layout tasi_set_goals (one goal per block) and tasi_extract_paths' row_ranges work on
 * (block b = cell rows b * (length + 1) ... + length - 1). Block b shows maps[index[b]]; every distinct map crosses
 * the bus once and is replicated on the device. Host pointers are not retained. Batches of queries
 * (BASELINE.json cfg5); the reference plans one query at a time (set_map + set_goal + step per query). */
int tasi_set_map_stack(tasi_planner *p, const uint8_t *const *maps, int n_maps, const int *index, int n_blocks,
                       int width, int length);

/* ---- focused planning ----
 * The reference's plan() stops once the start is settled (end_condition: FieldDPlanner_impl.h:225-256 waits for the
 * four corner nodes of the start cell, DynamicFastMarching_impl.h:315-320 for the start cell; the shifted-grid
 * planner as FD*). With enabled != 0, step() does the same: the solver's value-ordered schedule ends the launch as
 * soon as every value up to (highest finite value among those elements) + slack is final; tiles with later work
 * stay parked and the next step() — after a patch, a new start, or focus switched off — resumes them. Values at or
 * above tasi_stats.settled_below may still be upper bounds (as the reference's are beyond its queue's top key).
 * slack < 0 = default (4 * 254: covers every node within two cells of the start cell, what the extractor's first
 * step reads). Off by default: step() then converges the whole field. Not available on row bands. */
int tasi_set_focus(tasi_planner *p, int enabled, float slack);

/* ---- parent (back-pointer) field ----
 * What the reference's level-1 planners keep as INFO next to (g, rhs): for FD* / MFD* the node b whose triangle
 * (b, ccw(b)) gave the value (FieldDPlanner_impl.h:91-95, 196-208), for MS-DFM the (lower, higher) cells of the winning
 * stencil (DynamicFastMarching_impl.h:322-342). Computed on demand from the field for the window (x0, y0, nx, ny), which
 * must lie inside the field: out receives nx * ny * 4 signed bytes (dx1, dy1, dx2, dy2) = offsets of the two parents
 * (FD* / MFD*: b and ccw(b)); (0, 0) = absent (goals, unreached elements, one-sided MS-DFM updates). Among triangles
 * that tie, the last in the order of Graph::neighbors_8 is reported (min_rhs<1>); the reference's own choice among
 * ties depends on its expansion order. */
int tasi_read_parents(tasi_planner *p, int x0, int y0, int nx, int ny, signed char *out);

/* ---- path extraction over the converged field ----
 * Replaces LinearInterpolationPathExtractor::extract_path (PathExtraction/LinearInterpolationPathExtractor_impl.h:
 * 12-58; members path_, cost_, total_cost, total_dist, lookahead, max_steps, allow_indirect_traversals of
 * LinearInterpolationPathExtractor.h:14-22) for n independent walks over the planner's field in ONE kernel launch
 * (one warp per walk). starts_xy / goals_xy: n (x, y) pairs; NULL = the planner's own start / goal (n must be 1).
 * row_ranges (may be NULL): n (first cell row, cell rows) pairs — walk q then treats that block of rows as its
 * whole map and its positions are relative to the block (one map of a stack planned with tasi_set_goals).
 * Per walk q: xy[q * cap_points * 2 ...] receives n_points[q] positions (the start first), step_costs (may be
 * NULL) n_costs[q] piece costs, totals[2q] / totals[2q+1] the path cost and length, status[q] one of TASI_WALK_*.
 * The reference's defaults are max_steps 20, lookahead 1, allow_indirect 1; one step adds at most 3 positions and
 * 2 piece costs. e_ms (may be NULL): wall clock of the call, as e_time. Not available on row bands. */
#define TASI_WALK_GOAL 0      /* the walk ended on the goal position */
#define TASI_WALK_MAX_STEPS 1 /* max_steps steps taken (the reference keeps such a path, so do we) */
#define TASI_WALK_STUCK 2     /* no edge around the last position can be crossed */
#define TASI_WALK_FULL 3      /* cap_points / cap_costs reached */
int tasi_extract_paths(tasi_planner *p, const float *starts_xy, const float *goals_xy, const int *row_ranges, int n,
                       int max_steps,
                       int lookahead, int allow_indirect, float *xy, int cap_points, int *n_points,
                       float *step_costs, int cap_costs, int *n_costs, float *totals, int *status, float *e_ms);

#ifdef __cplusplus
}
#endif
#endif /* TASI_GPU_H */
