// Sharded planner: ONE grid cut into row bands over several GPUs, behind the same entry points as a single-device
// planner (tasi_create_sharded / tasi_create_sharded_rank in include/tasi_gpu.h). Included by tasi_capi.cu.
//
// The reference has no multi-device path; this is the sharded form of the same fixed-point solve and
// its contract is bit-identity with the single-device field.
//
//   * Distribution: the element rows are cut into `parts` = ranks x bands_per_rank bands of R rows (R a multiple of the
//     tile side) dealt out CYCLICALLY, band g to rank g % ranks, so that every GPU holds a piece of a wavefront for the
//     whole sweep (contiguous blocks keep a GPU busy only while the front crosses its block).
//   * Layout per GPU: its bands are stacked in ONE grid (one GridDesc, one tile queue, one persistent launch), separated
//     by one DEAD tile row each: the first row of a dead tile row is the halo below the band above it, its last row the
//     halo above the band below it (values and, for the node planners, the cell costs the first node row reads). Dead
//     tile rows are never scheduled (GridDesc::band_rt), everything else of the solver is unchanged.
//   * Outer iteration, enqueued on the streams with no host synchronisation: persistent solver up to a COMMON value
//     limit (tiles whose lowest pending value lies beyond it stay parked) -> pack the first/last row of every band ->
//     exchange with the ring neighbours (rank r's "down" rows all go to rank r+1, its "up" rows to rank r-1: one message
//     per direction) -> merge into the halo rows (min) and flag the border tiles -> all-reduce (max) of {any halo value
//     lowered, -lowest parked value} -> a one-thread kernel raises the limit and sets the convergence flag, both in
//     device memory. The host only watches the flag K iterations behind (pinned memory), so the GPUs never wait for it.
//   * Exchange: ranks in one process (tasi_create_sharded) read each other's send buffers and reduction slots through
//     peer-mapped pointers, ordered by stream events; ranks in different processes (tasi_create_sharded_rank, one per
//     GPU under torchrun) use NCCL send/recv + all-reduce on the planner's stream (libnccl is dlopen'ed: the already
//     loaded one of the hosting process if there is one). TASI_SHARD_NCCL=1 makes the one-process form use NCCL too.
//   * Replans: patch_map applies the patch to every band that stores the cell rows (owners and halo copies); phase 1
//     (invalidate) runs host-paced with a "replace" import until no band changes a halo, phase 2 is the loop above.
#pragma once
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <memory>

namespace {

constexpr int SHARD_MAX_RANKS = 16;
constexpr int SHARD_FLAG_RING = 4;  // the host looks at the convergence flag this many iterations behind

// ---- NCCL, loaded at run time -----------------------------------------------------------------------------------
struct NcclApi {
    void *handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t *, int, const int *) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
    std::string err;

    bool load() {
        if (handle) return true;
        const char *names[] = {getenv("TASI_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
        // the copy the hosting process already uses (torch's bundled libnccl.so.2), else whatever the loader finds
        for (int pass = 0; pass < 2 && !handle; ++pass)
            for (const char *n : names)
                if (n && *n && (handle = dlopen(n, RTLD_NOW | (pass == 0 ? RTLD_NOLOAD : 0) | RTLD_GLOBAL))) break;
        if (!handle) { err = "libnccl.so.2 not found (set TASI_NCCL_LIB)"; return false; }
#define TASI_NCCL_SYM(field, sym)                                               \
        field = (decltype(field)) dlsym(handle, sym);                           \
        if (!field) { err = std::string("libnccl lacks ") + sym; return false; }
        TASI_NCCL_SYM(GetUniqueId, "ncclGetUniqueId")
        TASI_NCCL_SYM(CommInitRank, "ncclCommInitRank")
        TASI_NCCL_SYM(CommInitAll, "ncclCommInitAll")
        TASI_NCCL_SYM(CommDestroy, "ncclCommDestroy")
        TASI_NCCL_SYM(Send, "ncclSend")
        TASI_NCCL_SYM(Recv, "ncclRecv")
        TASI_NCCL_SYM(AllReduce, "ncclAllReduce")
        TASI_NCCL_SYM(GroupStart, "ncclGroupStart")
        TASI_NCCL_SYM(GroupEnd, "ncclGroupEnd")
        TASI_NCCL_SYM(GetErrorString, "ncclGetErrorString")
#undef TASI_NCCL_SYM
        return true;
    }
};

NcclApi &nccl_api() {
    static NcclApi api;
    return api;
}

// ---- device side of the outer iteration -------------------------------------------------------------------------
struct ShardCtl {
    float limit;       // common value limit of the next solver launch
    float step;        // how far past the lowest parked value the limit is raised
    int converged;     // nothing lowered anywhere and nothing parked anywhere
    int iters;         // outer iterations done
    float red[2][2];   // [parity]{any halo value lowered ? 1 : 0, -(lowest parked value)}: reduced with max
    unsigned long long changed;  // invalidate loop: halo elements replaced by the last import
};

struct ImportJob {
    const float *src;  // EY values (a neighbour's packed boundary row), nullptr = none
    int hrow;          // local halo row they are merged into
    int trow;          // tile row of the tiles that read that halo row
};

struct PackJob {
    int row;           // local element row
    float *dst;        // EY floats
};

// Reset the per-launch part of the queue state on the device (what run_solver uploads from the host): tickets, pending
// count, done flag; the far pile, the threshold and the work counters carry over. mode 1 (invalidate) has no order.
__global__ void shard_rearm_kernel(GridDesc d, const ShardCtl *sc, int mode) {
    DevCtl *c = d.ctl;
    c->head = c->tail = 0;
    c->pending = 0;
    c->done = 0;
    c->kick = 0;
    c->focus_hit = 0;
    c->far_min_bits = INF_BITS;
    c->far_count[0] = min(c->far_count[0], d.far_cap);
    c->far_count[1] = min(c->far_count[1], d.far_cap);
    if (mode == 1) {
        c->thr_bits = INF_BITS;
        c->limit_bits = INF_BITS;
    } else {
        c->limit_bits = __float_as_int(sc->limit);
    }
}

__global__ void shard_pack_kernel(GridDesc d, const PackJob *jobs, int n_jobs) {
    const int j = blockIdx.y;
    if (j >= n_jobs) return;
    const float *src = d.g + (int64_t) jobs[j].row * d.gpitch;
    float *dst = jobs[j].dst;
    for (int y = blockIdx.x * blockDim.x + threadIdx.x; y < d.EY; y += gridDim.x * blockDim.x) dst[y] = __ldcg(src + y);
}

// Merge received boundary rows into the halo rows and flag the tiles that read the elements that changed.
// replace == 0 (relax loop): halo = min(halo, received), keys carry the lowered value; replace == 1 (invalidate
// loop): halo = received wherever it differs (values only rise to INF there).
__global__ void shard_import_kernel(GridDesc d, const ImportJob *jobs, int n_jobs, int replace, ShardCtl *sc) {
    const int j = blockIdx.y;
    if (j >= n_jobs || jobs[j].src == nullptr) return;
    const float *src = jobs[j].src;
    const int hrow = jobs[j].hrow, trow = jobs[j].trow;
    unsigned cnt = 0;
    for (int yy = blockIdx.x * blockDim.x + threadIdx.x; yy < d.EY; yy += gridDim.x * blockDim.x) {
        float *hp = d.g + (int64_t) hrow * d.gpitch + yy;
        const float nv = __ldcg(src + yy), ov = *hp;
        const bool take = replace ? (__float_as_int(nv) != __float_as_int(ov)) : (nv < ov);
        if (take) {
            *hp = nv;
            ++cnt;
            const int ty = yy / TS;
            const int kb = replace ? 0 : __float_as_int(nv);
            for (int t = max(ty - ((yy & (TS - 1)) == 0 ? 1 : 0), 0);
                 t <= min(ty + ((yy & (TS - 1)) == TS - 1 ? 1 : 0), d.TY - 1); ++t) {
                atomicMin(&d.key[trow * d.TY + t], kb);  // a parked tile must be released when the schedule reaches nv
                d.seed[trow * d.TY + t] = 1;
            }
        }
    }
    cnt = __reduce_add_sync(0xffffffffu, cnt);
    if ((threadIdx.x & 31) == 0 && cnt) {
        if (replace) atomicAdd(&sc->changed, (unsigned long long) cnt);
        else atomicAdd(&d.ctl->halo_lowered, (unsigned long long) cnt);
    }
}

// This rank's contribution to the reduction of the iteration, and the counters of the next one.
__global__ void shard_reduce_prep_kernel(GridDesc d, ShardCtl *sc, int parity) {
    DevCtl *c = d.ctl;
    const bool lowered = c->halo_lowered != 0;
    c->halo_lowered = 0;
    const bool parked = c->far_count[c->far_sel] > 0;
    const float fmin = parked ? __int_as_float(c->far_min_bits) : __int_as_float(INF_BITS);
    sc->red[parity][0] = lowered ? 1.0f : 0.0f;
    sc->red[parity][1] = -fmin;
}

struct RedPtrs {
    const float *p[SHARD_MAX_RANKS];  // each rank's red[parity] (peer-mapped), or only this rank's own after an all-reduce
    int n;
};

__global__ void shard_limit_kernel(ShardCtl *sc, RedPtrs r, int *flag_out) {
    float lowered = 0.0f, negmin = -__int_as_float(INF_BITS);
    for (int i = 0; i < r.n; ++i) {
        lowered = fmaxf(lowered, __ldcg(r.p[i]));
        negmin = fmaxf(negmin, __ldcg(r.p[i] + 1));
    }
    const float gmin = -negmin;
    const float INF = __int_as_float(INF_BITS);
    if (!(lowered > 0.0f) && gmin == INF) sc->converged = 1;
    if (gmin < INF) sc->limit = fmaxf(sc->limit, gmin + sc->step);
    sc->iters += 1;
    if (flag_out) *flag_out = sc->converged;
}

// fused mode: queue state and control words before a launch (everything is parked first: enqueue_seeds ordered)
__global__ void fused_rearm_kernel(GridDesc d, FusedCtl *fc, FusedTopo *topo, int mode, int rank) {
    DevCtl *c = d.ctl;
    c->head = c->tail = 0;
    c->pending = 0;
    c->done = 0;
    c->kick = 0;
    c->focus_hit = 0;
    c->far_min_bits = INF_BITS;
    c->limit_bits = INF_BITS;
    c->thr_bits = mode == 1 ? INF_BITS : 0;
    c->far_sel = 0;
    c->far_count[0] = c->far_count[1] = 0;
    for (int b = 0; b < QB_MAX; ++b) { c->bhead[b] = c->btail[b] = 0; c->bcount[b] = 0; }
    fc->inbox_count = 0;
    fc->rx = fc->tx = 0;
    if (rank == 0) fc->active = 0;
    topo->mode = mode;
}

__device__ __forceinline__ uint64_t shard_mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

// Order-independent checksum of `rows` element rows: sum over the elements of mix64(global index, value bits), mod 2^64.
// Two fields have the same sum iff they are bit-identical (up to a 2^-64 collision); sums of disjoint row ranges add.
__global__ void field_checksum_kernel(const float *g, int64_t gpitch, int rows, int EY, int64_t global_row0,
                                      unsigned long long *out) {
    unsigned long long acc = 0;
    const int64_t n = (int64_t) rows * EY;
    for (int64_t i = blockIdx.x * (int64_t) blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        const int r = (int) (i / EY), c = (int) (i - (int64_t) r * EY);
        const uint64_t bits = (uint32_t) __float_as_int(__ldcg(g + (int64_t) r * gpitch + c));
        const uint64_t idx = (uint64_t) (global_row0 + r) * (uint64_t) EY + (uint64_t) c;
        acc += shard_mix64(idx * 0x100000001B3ULL ^ (bits << 32 | bits));
    }
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) atomicAdd(out, acc);
}

// ---- host side ---------------------------------------------------------------------------------------------------
struct ShardBand {
    int g = 0;                 // global index
    int64_t a = 0, b = 0;      // element rows [a, b)
    int owner = 0, local = 0;  // rank and position in its stack
    int row0 = 0;              // first local element row
    int nrows = 0;
};

struct ShardMember {
    int rank = 0, device = 0;
    tasi_planner *p = nullptr;  // the stack of this rank's bands (an ordinary planner object in stack mode)
    std::vector<int> bands;     // global band indices, ascending
    float *xbuf = nullptr;      // [parity 2][up, dn, recv_top, recv_bot][B][EY]
    ShardCtl *sctl = nullptr;
    PackJob *pack_jobs = nullptr;      // [parity][2B]
    ImportJob *import_jobs = nullptr;  // [parity][2B]
    int *flags_host = nullptr;         // pinned ring of convergence flags
    unsigned long long *sum_dev = nullptr;
    cudaEvent_t ev_packed = nullptr, ev_red = nullptr, ev_flag[SHARD_FLAG_RING] = {}, ev_t0 = nullptr, ev_t1 = nullptr;
    ncclComm_t comm = nullptr;
    // fused mode
    FusedCtl *fc = nullptr;
    FusedTopo *topo = nullptr;     // device copy
    std::vector<void *> ipc_open;  // pointers opened with cudaIpcOpenMemHandle (multi-process form)
    float *buf(int parity, int what, int B, int EY) const { return xbuf + ((size_t) (parity * 4 + what) * B) * EY; }
};

}  // namespace

struct tasi_shard {
    int kind = 0, level = 0;
    int world = 1;        // ranks of the whole grid
    int my_rank = -1;     // >= 0: one rank per process (multi-process form)
    bool use_nccl = false;
    int want_per_rank = 0, per_rank = 1, parts = 1;  // want_per_rank 0 = by grid size (shard_layout)
    int k_phases = 1;
    bool want_fused = true;   // use the fused mode whenever the ranks can map each other's memory
    bool fused = false;       // ... and they do (set by shard_connect_local / tasi_shard_ipc_import)
    std::vector<ShardMember> local;
    std::vector<ShardBand> bands;
    // geometry
    int L = 0, W = 0, EY = 0, R = 0, RT = 0;
    int64_t n_elem_rows = 0;
    // the reference's flags (ReplannerBase.h:144-152), as in tasi_planner
    bool initialize_graph = true, initialize_search = true, goal_set = false, new_goal = false, new_start = false;
    int thr = 254;
    float heuristic_multiplier = 1.0f;
    float start_x = 0, start_y = 0, goal_x = 0, goal_y = 0;
    bool have_goal_elem = false;
    int64_t goal_ex = -1, goal_ey = -1;
    bool patch_pending = false;
    uint64_t num_nodes_updated = 0, num_nodes_expanded = 0, last_seed_count = 0;
    int max_blocks = 0;
    // results of the last step
    tasi_stats stats{};
    tasi_shard_stats sstats{};
    std::string err;
};

namespace {

#define SH_TRY(s, call)                                                                              \
    do {                                                                                             \
        cudaError_t e__ = (call);                                                                    \
        if (e__ != cudaSuccess) {                                                                    \
            (s)->err = std::string(#call) + ": " + cudaGetErrorString(e__);                          \
            return TASI_ERR_CUDA;                                                                    \
        }                                                                                            \
    } while (0)

#define SH_NCCL(s, call)                                                                             \
    do {                                                                                             \
        ncclResult_t r__ = (call);                                                                   \
        if (r__ != ncclSuccess) {                                                                    \
            (s)->err = std::string(#call) + ": " + nccl_api().GetErrorString(r__);                   \
            return TASI_ERR_CUDA;                                                                    \
        }                                                                                            \
    } while (0)

inline int sh_fail(tasi_shard *s, int code, const std::string &what) {
    s->err = what;
    return code;
}

inline int sh_sub(tasi_shard *s, ShardMember &m, int rc) {
    if (rc != TASI_OK) s->err = "rank " + std::to_string(m.rank) + ": " + m.p->err;
    return rc;
}

inline bool sh_node(const tasi_shard *s) { return s->kind != TASI_DFM; }
inline bool sh_multi_process(const tasi_shard *s) { return s->my_rank >= 0; }

void shard_close_ipc(ShardMember &m) {
    cudaSetDevice(m.device);
    for (void *q : m.ipc_open) cudaIpcCloseMemHandle(q);
    m.ipc_open.clear();
}

void shard_free_member_buffers(ShardMember &m) {
    cudaSetDevice(m.device);
    shard_close_ipc(m);
    cudaFree(m.xbuf); cudaFree(m.pack_jobs); cudaFree(m.import_jobs);
    m.xbuf = nullptr; m.pack_jobs = nullptr; m.import_jobs = nullptr;
    if (m.p) {
        cudaFree(m.p->inbox);
        m.p->inbox = nullptr; m.p->inbox_cap = 0; m.p->ft = nullptr;
    }
}

int shard_connect_local(tasi_shard *s);
int shard_sync_all(tasi_shard *s);

// Cut the element rows into bands and (re)allocate every local stack.
int shard_layout(tasi_shard *s, int W, int L) {
    const bool node = sh_node(s);
    const int64_t n_rows = node ? (int64_t) L + 1 : L;
    const int EY = node ? W + 1 : W;
    // bands per GPU: unless the caller fixed it, bands of about 32 tile rows, between 2 and 8 per GPU (measured, fused mode:
    // 65536^2 on 8 GPUs 665 ms with 4 bands per GPU, 616 ms with 7-8; 16384^2 on 2 GPUs 130.2 -> 128.4 ms; thinner bands
    // only add border traffic)
    int per_rank = s->want_per_rank > 0 ? s->want_per_rank
                                        : (int) std::min<int64_t>(8, std::max<int64_t>(2, n_rows / ((int64_t) s->world * 32 * TS)));
    int R = 0, parts = 0;
    for (;; --per_rank) {
        parts = s->world * per_rank;
        R = (int) (((n_rows + parts - 1) / parts + TS - 1) / TS) * TS;
        if ((int64_t) (parts - 1) * R < n_rows) break;  // every band owns at least one row
        if (per_rank == 1) return sh_fail(s, TASI_ERR_ARG, "map too small to give every device a band of rows");
    }
    const bool same = !s->initialize_graph && s->W == W && s->L == L && s->per_rank == per_rank;
    s->W = W; s->L = L; s->EY = EY; s->R = R; s->RT = R / TS; s->n_elem_rows = n_rows;
    s->per_rank = per_rank; s->parts = parts;
    s->bands.assign(parts, ShardBand());
    for (int g = 0; g < parts; ++g) {
        ShardBand &b = s->bands[g];
        b.g = g;
        b.a = (int64_t) g * R;
        b.b = std::min<int64_t>(b.a + R, n_rows);
        b.owner = g % s->world;
        b.local = g / s->world;
        b.row0 = b.local * (s->RT + 1) * TS;
        b.nrows = (int) (b.b - b.a);
    }
    for (ShardMember &m : s->local) {
        m.bands.clear();
        for (int j = 0; j < per_rank; ++j) m.bands.push_back(m.rank + j * s->world);
        const ShardBand &last = s->bands[m.bands.back()];
        const int EX = last.row0 + last.nrows;
        tasi_planner *p = m.p;
        SH_TRY(s, cudaSetDevice(m.device));
        p->thr = s->thr;
        p->band_rt = s->RT;
        p->shard_world = s->world;
        int rc = alloc_grid_dims(p, W, EX, EX, EY);
        if (rc != TASI_OK) return sh_sub(s, m, rc);
        if (same && m.xbuf) continue;
        shard_free_member_buffers(m);
        const int B = per_rank;
        SH_TRY(s, cudaMalloc(&m.xbuf, sizeof(float) * 8 * (size_t) B * EY));
        SH_TRY(s, cudaMalloc(&m.pack_jobs, sizeof(PackJob) * 4 * B));
        SH_TRY(s, cudaMalloc(&m.import_jobs, sizeof(ImportJob) * 4 * B));
        p->inbox_cap = 4 * B * p->TY + 64;  // every border tile of the stack, twice over
        SH_TRY(s, cudaMalloc(&p->inbox, sizeof(int) * p->inbox_cap));
    }
    s->fused = false;
    for (ShardMember &m : s->local) m.p->ft = nullptr;
    // job tables (they hold peer pointers, so they are built once every member's buffers exist)
    for (ShardMember &m : s->local) {
        const int B = per_rank;
        std::vector<PackJob> pj(4 * B);
        std::vector<ImportJob> ij(4 * B);
        const bool nccl = s->use_nccl && s->world > 1;
        auto member_of = [&](int rank) -> ShardMember * {
            for (ShardMember &o : s->local)
                if (o.rank == rank) return &o;
            return nullptr;
        };
        for (int par = 0; par < 2; ++par)
            for (int j = 0; j < B; ++j) {
                const ShardBand &b = s->bands[m.bands[j]];
                pj[par * 2 * B + j] = {b.row0, m.buf(par, 0, B, EY) + (size_t) j * EY};                    // up: first row
                pj[par * 2 * B + B + j] = {b.row0 + b.nrows - 1, m.buf(par, 1, B, EY) + (size_t) j * EY};  // dn: last row
                ImportJob top = {nullptr, b.row0 - 1, b.row0 / TS};
                ImportJob bot = {nullptr, b.row0 + b.nrows, (b.row0 + b.nrows - 1) / TS};
                if (b.g > 0) {  // the band above is band g-1: its last row is row j' of its owner's "dn" rows
                    const ShardBand &up = s->bands[b.g - 1];
                    if (nccl) top.src = m.buf(par, 2, B, EY) + (size_t) up.local * EY;
                    else top.src = member_of(up.owner)->buf(par, 1, B, EY) + (size_t) up.local * EY;
                }
                if (b.g + 1 < parts) {
                    const ShardBand &dn = s->bands[b.g + 1];
                    if (nccl) bot.src = m.buf(par, 3, B, EY) + (size_t) dn.local * EY;
                    else bot.src = member_of(dn.owner)->buf(par, 0, B, EY) + (size_t) dn.local * EY;
                }
                ij[par * 2 * B + j] = top;
                ij[par * 2 * B + B + j] = bot;
            }
        SH_TRY(s, cudaSetDevice(m.device));
        SH_TRY(s, cudaMemcpy(m.pack_jobs, pj.data(), sizeof(PackJob) * pj.size(), cudaMemcpyHostToDevice));
        SH_TRY(s, cudaMemcpy(m.import_jobs, ij.data(), sizeof(ImportJob) * ij.size(), cudaMemcpyHostToDevice));
    }
    return shard_connect_local(s);
}

// Fused mode for the ranks of ONE process: every rank gets the GridDesc of its ring neighbours and the control blocks of
// all ranks (plain device pointers; distinct devices are peer-mapped at creation). Not with several ranks on one
// device: their persistent kernels would wait on each other inside one GPU.
int shard_upload_topo(tasi_shard *s, ShardMember &m, const GridDesc &up, const GridDesc &dn, FusedCtl *const *fcs) {
    FusedTopo t;
    memset(&t, 0, sizeof t);
    t.world = s->world; t.rank = m.rank; t.parts = s->parts; t.mode = 0;
    t.nbr[0] = up; t.nbr[1] = dn;
    t.nbr[0].ft = nullptr; t.nbr[1].ft = nullptr;
    for (int r = 0; r < s->world; ++r) t.fc[r] = fcs[r];
    SH_TRY(s, cudaSetDevice(m.device));
    SH_TRY(s, cudaMemcpy(m.topo, &t, sizeof t, cudaMemcpyHostToDevice));
    m.p->ft = m.topo;
    return TASI_OK;
}

int shard_connect_local(tasi_shard *s) {
    s->fused = false;
    if (!s->want_fused || sh_multi_process(s) || s->use_nccl) return TASI_OK;
    for (size_t i = 0; i < s->local.size(); ++i)
        for (size_t j = 0; j < i; ++j)
            if (s->local[i].device == s->local[j].device) return TASI_OK;
    FusedCtl *fcs[FUSED_MAX_RANKS] = {};
    for (ShardMember &m : s->local) fcs[m.rank] = m.fc;
    for (ShardMember &m : s->local) {
        const ShardMember &up = s->local[(m.rank + s->world - 1) % s->world], &dn = s->local[(m.rank + 1) % s->world];
        int rc = shard_upload_topo(s, m, make_desc(up.p), make_desc(dn.p), fcs);
        if (rc != TASI_OK) return rc;
    }
    s->fused = true;
    return TASI_OK;
}

// One fused launch per rank: everything flagged is parked, the solvers of all ranks run until the coordinator ends them.
int shard_fused_run(tasi_shard *s, int mode) {
    for (ShardMember &m : s->local) {
        tasi_planner *p = m.p;
        SH_TRY(s, cudaSetDevice(m.device));
        int rc = ensure_cost_stats(p);
        if (rc != TASI_OK) return sh_sub(s, m, rc);
        GridDesc d = make_desc(p);
        d.n_focus = 0;
        d.fmode = mode;
        if (mode == 1) { d.nb = 0; d.defer_margin = -1.0f; }  // the invalidate launch has no order
        const int nt = p->TX * p->TY;
        fused_rearm_kernel<<<1, 1, 0, p->stream>>>(d, m.fc, m.topo, mode, m.rank);
        if (mode == 0) seed_keys_kernel<<<std::min((nt + 7) / 8, p->num_sms * 8), 256, 0, p->stream>>>(d);
        enqueue_seeds_kernel<<<std::min((nt + 255) / 256, 1024), 256, 0, p->stream>>>(d, 1);
        const int cap = s->max_blocks > 0 ? s->max_blocks : p->max_blocks;
        int nblocks = (cap > 0 && cap < p->solver_blocks) ? cap : p->solver_blocks;
        if (cap <= 0) {
            const int rb = relax_blocks(p, mode, nt);  // a third CTA per SM on large stacks (see relax_blocks)
            if (rb > 0) nblocks = rb;
        }
        pick_solver(p->kind, mode, p->strip, compact_for(p))<<<nblocks, TS * TS / p->strip, 0, p->stream>>>(d, p->tm);
        SH_TRY(s, cudaGetLastError());
    }
    int rc = shard_sync_all(s);
    if (rc != TASI_OK) return rc;
    s->sstats.fused_launches += 1;
    return TASI_OK;
}

// cell rows band `b` stores: its own, plus (node planners) the row above its first node row
inline void shard_cell_rows(const tasi_shard *s, const ShardBand &b, int64_t *c0, int64_t *c1) {
    if (sh_node(s)) { *c0 = std::max<int64_t>(b.a - 1, 0); *c1 = std::min<int64_t>(b.b, s->L); }
    else { *c0 = b.a; *c1 = b.b; }
}

int shard_sync_all(tasi_shard *s) {
    for (ShardMember &m : s->local) {
        SH_TRY(s, cudaSetDevice(m.device));
        SH_TRY(s, cudaStreamSynchronize(m.p->stream));
    }
    return TASI_OK;
}

int shard_after_map(tasi_shard *s) {
    for (ShardMember &m : s->local) {
        m.p->initialize_graph = false;
        m.p->cost_stats_valid = false;
    }
    int rc = shard_sync_all(s);
    if (rc != TASI_OK) return rc;
    s->initialize_graph = false;
    s->initialize_search = true;
    s->patch_pending = false;
    return TASI_OK;
}

int shard_set_map(tasi_shard *s, const uint8_t *host_map, int width, int length) {
    if (!host_map || width <= 0 || length <= 0) return TASI_ERR_ARG;
    int rc = shard_layout(s, width, length);
    if (rc != TASI_OK) return rc;
    for (ShardMember &m : s->local) {
        SH_TRY(s, cudaSetDevice(m.device));
        for (int g : m.bands) {
            const ShardBand &b = s->bands[g];
            int64_t c0, c1;
            shard_cell_rows(s, b, &c0, &c1);
            if (c1 <= c0) continue;
            SH_TRY(s, cudaMemcpy2DAsync(c_origin(m.p) + (int64_t) (b.row0 + (c0 - b.a)) * m.p->cpitch, m.p->cpitch,
                                        host_map + (size_t) c0 * width, width, width, c1 - c0, cudaMemcpyHostToDevice,
                                        m.p->stream));
        }
    }
    return shard_after_map(s);
}

int shard_synth_map(tasi_shard *s, uint64_t seed, int width, int length, int p_obs, int sx, int sy, int gx, int gy) {
    if (width <= 0 || length <= 0) return TASI_ERR_ARG;
    int rc = shard_layout(s, width, length);
    if (rc != TASI_OK) return rc;
    for (ShardMember &m : s->local) {
        SH_TRY(s, cudaSetDevice(m.device));
        for (int g : m.bands) {
            const ShardBand &b = s->bands[g];
            int64_t c0, c1;
            shard_cell_rows(s, b, &c0, &c1);
            if (c1 <= c0) continue;
            synth_map_kernel<<<m.p->num_sms * 4, 256, 0, m.p->stream>>>(c_origin(m.p), m.p->cpitch, (int) (c1 - c0), width,
                                                                       c0, (int) (b.row0 + (c0 - b.a)), seed, p_obs, sx,
                                                                       sy, gx, gy);
        }
        SH_TRY(s, cudaGetLastError());
    }
    return shard_after_map(s);
}

// rows [x0, x0 + nx) of the element grid -> dst (row pitch in elements), from whichever local member owns them.
// Rows owned by another process are left untouched and counted in *missing.
int shard_gather_rows(tasi_shard *s, int64_t x0, int64_t nx, int y0, int ny, float *dst, int64_t dst_pitch,
                      cudaMemcpyKind kind, int64_t *missing) {
    int64_t have = 0;
    for (ShardMember &m : s->local) {
        SH_TRY(s, cudaSetDevice(m.device));
        for (int g : m.bands) {
            const ShardBand &b = s->bands[g];
            const int64_t lo = std::max(b.a, x0), hi = std::min(b.b, x0 + nx);
            if (hi <= lo) continue;
            SH_TRY(s, cudaMemcpy2DAsync(dst + (lo - x0) * dst_pitch, sizeof(float) * dst_pitch,
                                        g_origin(m.p) + (int64_t) (b.row0 + (lo - b.a)) * m.p->gpitch + y0,
                                        sizeof(float) * m.p->gpitch, sizeof(float) * ny, hi - lo, kind, m.p->stream));
            have += hi - lo;
        }
    }
    int rc = shard_sync_all(s);
    if (rc != TASI_OK) return rc;
    if (missing) *missing = nx - have;
    return TASI_OK;
}

int shard_read_window(tasi_shard *s, int x0, int y0, int nx, int ny, float *host_out) {
    if (!host_out || nx <= 0 || ny <= 0 || s->initialize_graph) return TASI_ERR_ARG;
    const int64_t xa = std::max(x0, 0), xb = std::min<int64_t>((int64_t) x0 + nx, s->n_elem_rows);
    const int ya = std::max(y0, 0), yb = std::min(y0 + ny, s->EY);
    const bool clipped = xa != x0 || ya != y0 || xb != (int64_t) x0 + nx || yb != y0 + ny;
    if (clipped)
        for (int64_t i = 0; i < (int64_t) nx * ny; ++i) host_out[i] = INFINITY;
    if (xa < xb && ya < yb) {
        int64_t missing = 0;
        int rc = shard_gather_rows(s, xa, xb - xa, ya, yb - ya, host_out + (xa - x0) * ny + (ya - y0), ny,
                                   cudaMemcpyDeviceToHost, &missing);
        if (rc != TASI_OK) return rc;
        if (missing) return sh_fail(s, TASI_ERR_STATE, "read_window: rows owned by another process (read them on their rank)");
    }
    return TASI_OK;
}

int shard_read_map(tasi_shard *s, uint8_t *host_out) {
    if (!host_out || s->initialize_graph) return TASI_ERR_ARG;
    int64_t have = 0;
    for (ShardMember &m : s->local) {
        SH_TRY(s, cudaSetDevice(m.device));
        for (int g : m.bands) {
            const ShardBand &b = s->bands[g];
            const int64_t c0 = b.a, c1 = std::min<int64_t>(b.b, s->L);  // each cell row from the band that owns it
            if (c1 <= c0) continue;
            SH_TRY(s, cudaMemcpy2DAsync(host_out + (size_t) c0 * s->W, s->W,
                                        c_origin(m.p) + (int64_t) (b.row0 + (c0 - b.a)) * m.p->cpitch, m.p->cpitch,
                                        s->W, c1 - c0, cudaMemcpyDeviceToHost, m.p->stream));
            have += c1 - c0;
        }
    }
    int rc = shard_sync_all(s);
    if (rc != TASI_OK) return rc;
    if (have != s->L) return sh_fail(s, TASI_ERR_STATE, "read_map: rows owned by another process");
    return TASI_OK;
}

// ---- collectives over the ranks of the grid ---------------------------------------------------------------------------
// max / sum over all ranks of one value per local member; every rank gets the result. One-process form: on the host.
int shard_allreduce_host(tasi_shard *s, double *vals /* one per local member */, bool is_max, double *out) {
    double acc = is_max ? -INFINITY : 0.0;
    for (size_t i = 0; i < s->local.size(); ++i) acc = is_max ? std::max(acc, vals[i]) : acc + vals[i];
    if (sh_multi_process(s) && s->world > 1) {
        ShardMember &m = s->local[0];
        SH_TRY(s, cudaSetDevice(m.device));
        double *dv = (double *) m.sum_dev;
        SH_TRY(s, cudaMemcpyAsync(dv, &acc, sizeof(double), cudaMemcpyHostToDevice, m.p->stream));
        SH_NCCL(s, nccl_api().AllReduce(dv, dv, 1, ncclFloat64, is_max ? ncclMax : ncclSum, m.comm, m.p->stream));
        SH_TRY(s, cudaMemcpyAsync(&acc, dv, sizeof(double), cudaMemcpyDeviceToHost, m.p->stream));
        SH_TRY(s, cudaStreamSynchronize(m.p->stream));
    }
    *out = acc;
    return TASI_OK;
}

int shard_checksum(tasi_shard *s, uint64_t *sum) {
    if (s->initialize_graph) return TASI_ERR_ARG;
    uint64_t acc = 0;
    for (ShardMember &m : s->local) {
        SH_TRY(s, cudaSetDevice(m.device));
        SH_TRY(s, cudaMemsetAsync(m.sum_dev, 0, sizeof(unsigned long long), m.p->stream));
        for (int g : m.bands) {
            const ShardBand &b = s->bands[g];
            field_checksum_kernel<<<m.p->num_sms * 4, 256, 0, m.p->stream>>>(
                g_origin(m.p) + (int64_t) b.row0 * m.p->gpitch, m.p->gpitch, b.nrows, s->EY, b.a, m.sum_dev);
        }
        SH_TRY(s, cudaGetLastError());
        unsigned long long h = 0;
        SH_TRY(s, cudaMemcpyAsync(&h, m.sum_dev, sizeof h, cudaMemcpyDeviceToHost, m.p->stream));
        SH_TRY(s, cudaStreamSynchronize(m.p->stream));
        acc += h;
    }
    if (sh_multi_process(s) && s->world > 1) {
        ShardMember &m = s->local[0];
        unsigned long long h = acc;
        SH_TRY(s, cudaMemcpyAsync(m.sum_dev, &h, sizeof h, cudaMemcpyHostToDevice, m.p->stream));
        SH_NCCL(s, nccl_api().AllReduce(m.sum_dev, m.sum_dev, 1, ncclUint64, ncclSum, m.comm, m.p->stream));
        SH_TRY(s, cudaMemcpyAsync(&h, m.sum_dev, sizeof h, cudaMemcpyDeviceToHost, m.p->stream));
        SH_TRY(s, cudaStreamSynchronize(m.p->stream));
        acc = h;
    }
    *sum = acc;
    return TASI_OK;
}

// ---- the exchange of one outer iteration ------------------------------------------------------------------------------
// Called with every member's pack kernel enqueued; returns with every member's import kernel enqueued.
int shard_exchange_and_import(tasi_shard *s, int par, int replace) {
    const int B = s->per_rank, EY = s->EY;
    const bool nccl = s->use_nccl && s->world > 1;
    const size_t cnt = (size_t) B * EY;
    if (nccl) {
        NcclApi &N = nccl_api();
        SH_NCCL(s, N.GroupStart());
        for (ShardMember &m : s->local) {
            const int prev = (m.rank + s->world - 1) % s->world, next = (m.rank + 1) % s->world;
            // order matters only when prev == next (two ranks): my first send ("up" rows) meets the peer's first
            // receive (its "bottom" stage), my second send its second
            SH_NCCL(s, N.Send(m.buf(par, 0, B, EY), cnt, ncclFloat32, prev, m.comm, m.p->stream));
            SH_NCCL(s, N.Send(m.buf(par, 1, B, EY), cnt, ncclFloat32, next, m.comm, m.p->stream));
            SH_NCCL(s, N.Recv(m.buf(par, 3, B, EY), cnt, ncclFloat32, next, m.comm, m.p->stream));
            SH_NCCL(s, N.Recv(m.buf(par, 2, B, EY), cnt, ncclFloat32, prev, m.comm, m.p->stream));
        }
        SH_NCCL(s, N.GroupEnd());
    } else if (s->local.size() > 1) {
        for (ShardMember &m : s->local) {
            SH_TRY(s, cudaSetDevice(m.device));
            SH_TRY(s, cudaEventRecord(m.ev_packed, m.p->stream));
        }
        for (ShardMember &m : s->local) {
            SH_TRY(s, cudaSetDevice(m.device));
            for (ShardMember &o : s->local) {
                const int d = (o.rank - m.rank + s->world) % s->world;
                if (&o != &m && (d == 1 || d == s->world - 1)) SH_TRY(s, cudaStreamWaitEvent(m.p->stream, o.ev_packed, 0));
            }
        }
    }
    for (ShardMember &m : s->local) {
        SH_TRY(s, cudaSetDevice(m.device));
        GridDesc d = make_desc(m.p);
        dim3 grid((EY + 255) / 256, 2 * B);
        shard_import_kernel<<<grid, 256, 0, m.p->stream>>>(d, m.import_jobs + par * 2 * B, 2 * B, replace, m.sctl);
        SH_TRY(s, cudaGetLastError());
    }
    return TASI_OK;
}

// queue the flagged tiles and launch the persistent solver of one member without touching the host
int shard_launch_solver(tasi_shard *s, ShardMember &m, int mode) {
    tasi_planner *p = m.p;
    SH_TRY(s, cudaSetDevice(m.device));
    GridDesc d = make_desc(p);
    d.n_focus = 0;
    const int nt = p->TX * p->TY;
    const int ordered = (mode == 0 && p->delta_factor > 0.0f) ? 1 : 0;
    shard_rearm_kernel<<<1, 1, 0, p->stream>>>(d, m.sctl, mode);
    if (ordered) seed_keys_kernel<<<std::min((nt + 7) / 8, p->num_sms * 8), 256, 0, p->stream>>>(d);
    enqueue_seeds_kernel<<<std::min((nt + 255) / 256, 1024), 256, 0, p->stream>>>(d, ordered);
    arm_solver_kernel<<<1, 1, 0, p->stream>>>(d);
    const int cap = s->max_blocks > 0 ? s->max_blocks : p->max_blocks;
    const int nblocks = (cap > 0 && cap < p->solver_blocks) ? cap : p->solver_blocks;
    pick_solver(p->kind, mode, p->strip, compact_for(p))<<<nblocks, TS * TS / p->strip, 0, p->stream>>>(d, p->tm);
    SH_TRY(s, cudaGetLastError());
    return TASI_OK;
}

int shard_pack(tasi_shard *s, ShardMember &m, int par) {
    const int B = s->per_rank;
    GridDesc d = make_desc(m.p);
    dim3 grid(std::min((s->EY + 255) / 256, 64), 2 * B);
    shard_pack_kernel<<<grid, 256, 0, m.p->stream>>>(d, m.pack_jobs + par * 2 * B, 2 * B);
    SH_TRY(s, cudaGetLastError());
    return TASI_OK;
}

// reset the schedule of every member before a loop (fresh: also the work counters)
int shard_reset_ctl(tasi_shard *s, bool counters, float limit, float step) {
    for (ShardMember &m : s->local) {
        SH_TRY(s, cudaSetDevice(m.device));
        DevCtl z;
        memset(&z, 0, sizeof z);
        z.timeout_ns = (unsigned long long) (m.p->timeout_s * 1e9);
        z.limit_bits = INF_BITS;
        z.far_min_bits = INF_BITS;
        if (counters) {
            SH_TRY(s, cudaMemcpyAsync(m.p->ctl, &z, offsetof(DevCtl, n_changed), cudaMemcpyHostToDevice, m.p->stream));
        } else {  // keep the work counters of the phases before: only the schedule restarts
            SH_TRY(s, cudaMemcpyAsync(m.p->ctl, &z, offsetof(DevCtl, visits), cudaMemcpyHostToDevice, m.p->stream));
            SH_TRY(s, cudaMemcpyAsync((char *) m.p->ctl + offsetof(DevCtl, timeout_ns),
                                      (char *) &z + offsetof(DevCtl, timeout_ns),
                                      offsetof(DevCtl, far_parked) - offsetof(DevCtl, timeout_ns), cudaMemcpyHostToDevice,
                                      m.p->stream));
        }
        SH_TRY(s, cudaMemsetAsync(&m.p->ctl->halo_lowered, 0, sizeof(unsigned long long), m.p->stream));
        ShardCtl sc;
        memset(&sc, 0, sizeof sc);
        sc.limit = limit;
        sc.step = step;
        SH_TRY(s, cudaMemcpyAsync(m.sctl, &sc, sizeof sc, cudaMemcpyHostToDevice, m.p->stream));
        SH_TRY(s, cudaStreamSynchronize(m.p->stream));  // z and sc live on this stack frame
    }
    return TASI_OK;
}

// The relax loop (initial plan, or phase 2 of a replan): see the header of this file.
int shard_relax_loop(tasi_shard *s) {
    if (s->fused) return shard_fused_run(s, 0);
    // common step of the limit: k_phases threshold widths of the widest band
    std::vector<double> deltas;
    for (ShardMember &m : s->local) {
        SH_TRY(s, cudaSetDevice(m.device));
        int rc = ensure_cost_stats(m.p);
        if (rc != TASI_OK) return sh_sub(s, m, rc);
        deltas.push_back(band_width(m.p));
    }
    double dmax = 0;
    int rc = shard_allreduce_host(s, deltas.data(), true, &dmax);
    if (rc != TASI_OK) return rc;
    const float step = (float) (std::max(1, s->k_phases) * dmax);
    rc = shard_reset_ctl(s, false, step, step);
    if (rc != TASI_OK) return rc;

    const bool nccl = s->use_nccl && s->world > 1;
    const int max_outer = 1 << 22;
    int it = 0;
    bool converged = false;
    for (; it < max_outer && !converged; ++it) {
        const int par = it & 1, slot = it % SHARD_FLAG_RING;
        if (it >= SHARD_FLAG_RING) {  // the flag written SHARD_FLAG_RING iterations ago
            ShardMember &m0 = s->local[0];
            SH_TRY(s, cudaSetDevice(m0.device));
            SH_TRY(s, cudaEventSynchronize(m0.ev_flag[slot]));
            if (m0.flags_host[slot]) { converged = true; break; }
        }
        for (ShardMember &m : s->local) {
            rc = shard_launch_solver(s, m, 0);
            if (rc != TASI_OK) return rc;
            rc = shard_pack(s, m, par);
            if (rc != TASI_OK) return rc;
        }
        rc = shard_exchange_and_import(s, par, 0);
        if (rc != TASI_OK) return rc;
        for (ShardMember &m : s->local) {
            SH_TRY(s, cudaSetDevice(m.device));
            GridDesc d = make_desc(m.p);
            shard_reduce_prep_kernel<<<1, 1, 0, m.p->stream>>>(d, m.sctl, par);
            SH_TRY(s, cudaGetLastError());
        }
        if (nccl) {
            NcclApi &N = nccl_api();
            SH_NCCL(s, N.GroupStart());
            for (ShardMember &m : s->local)
                SH_NCCL(s, N.AllReduce(m.sctl->red[par], m.sctl->red[par], 2, ncclFloat32, ncclMax, m.comm, m.p->stream));
            SH_NCCL(s, N.GroupEnd());
        } else if (s->local.size() > 1) {
            for (ShardMember &m : s->local) {
                SH_TRY(s, cudaSetDevice(m.device));
                SH_TRY(s, cudaEventRecord(m.ev_red, m.p->stream));
            }
            for (ShardMember &m : s->local) {
                SH_TRY(s, cudaSetDevice(m.device));
                for (ShardMember &o : s->local)
                    if (&o != &m) SH_TRY(s, cudaStreamWaitEvent(m.p->stream, o.ev_red, 0));
            }
        }
        for (ShardMember &m : s->local) {
            SH_TRY(s, cudaSetDevice(m.device));
            RedPtrs rp;
            rp.n = 0;
            if (nccl) rp.p[rp.n++] = m.sctl->red[par];
            else
                for (ShardMember &o : s->local) rp.p[rp.n++] = o.sctl->red[par];
            const bool first = &m == &s->local[0];
            int *flag_dev = nullptr;
            if (first) SH_TRY(s, cudaHostGetDevicePointer((void **) &flag_dev, m.flags_host + slot, 0));
            shard_limit_kernel<<<1, 1, 0, m.p->stream>>>(m.sctl, rp, flag_dev);
            SH_TRY(s, cudaGetLastError());
            if (first) SH_TRY(s, cudaEventRecord(m.ev_flag[slot], m.p->stream));
        }
    }
    rc = shard_sync_all(s);
    if (rc != TASI_OK) return rc;
    if (!converged) {  // the loop ended on the flag of an earlier iteration or ran out: look at the last flags
        ShardCtl sc;
        ShardMember &m0 = s->local[0];
        SH_TRY(s, cudaSetDevice(m0.device));
        SH_TRY(s, cudaMemcpy(&sc, m0.sctl, sizeof sc, cudaMemcpyDeviceToHost));
        if (!sc.converged) return sh_fail(s, TASI_ERR_TIMEOUT, "sharded solve did not converge");
    }
    ShardCtl sc;
    ShardMember &m0 = s->local[0];
    SH_TRY(s, cudaSetDevice(m0.device));
    SH_TRY(s, cudaMemcpy(&sc, m0.sctl, sizeof sc, cudaMemcpyDeviceToHost));
    s->sstats.exchanges += sc.iters;
    return TASI_OK;
}

// Phase 1 of a replan on the sharded grid: invalidate in every band, swap boundary rows, replace the halos, until no
// halo changes anywhere. Host-paced (a replan's invalidated region is small and the loop runs a handful of times).
int shard_invalidate_loop(tasi_shard *s) {
    if (s->fused) return shard_fused_run(s, 1);
    int rc = shard_reset_ctl(s, false, INFINITY, 0.0f);
    if (rc != TASI_OK) return rc;
    for (int it = 0; it < (1 << 20); ++it) {
        const int par = it & 1;
        for (ShardMember &m : s->local) {
            SH_TRY(s, cudaSetDevice(m.device));
            SH_TRY(s, cudaMemsetAsync(&m.sctl->changed, 0, sizeof(unsigned long long), m.p->stream));
            rc = shard_launch_solver(s, m, 1);
            if (rc != TASI_OK) return rc;
            rc = shard_pack(s, m, par);
            if (rc != TASI_OK) return rc;
        }
        rc = shard_exchange_and_import(s, par, 1);
        if (rc != TASI_OK) return rc;
        std::vector<double> ch;
        for (ShardMember &m : s->local) {
            SH_TRY(s, cudaSetDevice(m.device));
            unsigned long long c = 0;
            SH_TRY(s, cudaMemcpyAsync(&c, &m.sctl->changed, sizeof c, cudaMemcpyDeviceToHost, m.p->stream));
            SH_TRY(s, cudaStreamSynchronize(m.p->stream));
            ch.push_back((double) c);
        }
        double total = 0;
        rc = shard_allreduce_host(s, ch.data(), false, &total);
        if (rc != TASI_OK) return rc;
        s->sstats.inval_exchanges += 1;
        if (total == 0) return TASI_OK;
    }
    return sh_fail(s, TASI_ERR_TIMEOUT, "sharded invalidate did not converge");
}

int shard_collect_stats(tasi_shard *s, bool fresh) {
    double busy_max = 0, busy_min = INFINITY, bf_max = 0, bf_min = INFINITY, rtx = 0;
    for (ShardMember &m : s->local) {
        SH_TRY(s, cudaSetDevice(m.device));
        DevCtl h;
        int rc = read_ctl(m.p, &h);
        if (rc != TASI_OK) return sh_sub(s, m, rc);
        if (h.error) return sh_fail(s, TASI_ERR_TIMEOUT, "solver watchdog fired on rank " + std::to_string(m.rank));
        s->stats.tile_visits += h.visits;
        s->stats.tile_rounds += h.rounds;
        s->stats.elem_evals += h.evals;
        s->stats.elem_lowered += h.lowered;
        s->stats.elem_invalidated += h.invalidated;
        s->stats.queue_pushes += h.pushes;
        s->stats.phases += (uint64_t) h.phases;
        s->stats.far_parked += h.far_parked;
        s->stats.solver_launches += h.launches * 9;  // re-arm, keys, enqueue, arm, solver, pack, import, reduce, limit
        s->stats.cyc_pop += h.cyc_pop; s->stats.cyc_stage += h.cyc_stage;
        s->stats.cyc_relax += h.cyc_relax; s->stats.cyc_post += h.cyc_post; s->stats.cyc_advance += h.cyc_advance;
        s->stats.tiles += (uint64_t) m.p->TX * m.p->TY;
        s->stats.solver_blocks = m.p->solver_blocks;
        const double ms = (double) h.busy_ns / 1e6;
        busy_max = std::max(busy_max, ms);
        busy_min = std::min(busy_min, ms);
        const double work = (double) h.cyc_stage + (double) h.cyc_relax + (double) h.cyc_post, all = work + (double) h.cyc_pop;
        const double bf = all > 0 ? work / all : 0.0;
        bf_max = std::max(bf_max, bf);
        bf_min = std::min(bf_min, bf);
        FusedCtl fh;
        SH_TRY(s, cudaMemcpy(&fh, m.fc, sizeof fh, cudaMemcpyDeviceToHost));
        rtx += (double) fh.tx;
        float dev_ms = 0;
        SH_TRY(s, cudaEventElapsedTime(&dev_ms, m.ev_t0, m.ev_t1));
        s->stats.step_ms_device = std::max(s->stats.step_ms_device, dev_ms);
    }
    std::vector<double> v(s->local.size(), busy_max);
    double out = 0;
    int rc = shard_allreduce_host(s, v.data(), true, &out);
    if (rc != TASI_OK) return rc;
    s->sstats.solver_ms_max_rank = (float) out;
    v.assign(s->local.size(), -busy_min);
    rc = shard_allreduce_host(s, v.data(), true, &out);
    if (rc != TASI_OK) return rc;
    s->sstats.solver_ms_min_rank = (float) -out;
    s->stats.kernel_ms_solve = s->sstats.solver_ms_max_rank;
    v.assign(s->local.size(), bf_max);
    rc = shard_allreduce_host(s, v.data(), true, &out);
    if (rc != TASI_OK) return rc;
    s->sstats.busy_frac_max_rank = (float) out;
    v.assign(s->local.size(), -bf_min);
    rc = shard_allreduce_host(s, v.data(), true, &out);
    if (rc != TASI_OK) return rc;
    s->sstats.busy_frac_min_rank = (float) -out;
    v.assign(s->local.size(), 0.0);
    v[0] = rtx;
    rc = shard_allreduce_host(s, v.data(), false, &out);
    if (rc != TASI_OK) return rc;
    s->sstats.remote_activations = (uint64_t) out;
    (void) fresh;
    return TASI_OK;
}

// local element row of global element row x on member m, or -1
inline int shard_local_row(const tasi_shard *s, const ShardMember &m, int64_t x) {
    if (x < 0 || x >= s->n_elem_rows) return -1;
    const ShardBand &b = s->bands[(int) (x / s->R)];
    return b.owner == m.rank ? (int) (b.row0 + (x - b.a)) : -1;
}

int shard_step(tasi_shard *s, float *u_ms, float *p_ms, uint64_t *n_updated, uint64_t *n_expanded) {
    if (s->initialize_graph) return TASI_LOOP_FAILURE_NO_GRAPH;  // ReplannerBase.h:44
    if (!s->goal_set) return TASI_LOOP_FAILURE_NO_GOAL;         // ReplannerBase.h:45
    memset(&s->stats, 0, sizeof s->stats);
    memset(&s->sstats, 0, sizeof s->sstats);
    s->sstats.n_ranks = s->world;
    s->sstats.bands = s->parts;
    s->sstats.bands_per_rank = s->per_rank;
    s->sstats.rows_per_band = s->R;
    s->sstats.nccl = (s->use_nccl && s->world > 1 && !s->fused) ? 1 : 0;
    s->sstats.fused = s->fused ? 1 : 0;
    int rc;
    const double t0 = now_ms();
    for (ShardMember &m : s->local) {
        SH_TRY(s, cudaSetDevice(m.device));
        SH_TRY(s, cudaEventRecord(m.ev_t0, m.p->stream));
    }
    const bool fresh = s->initialize_search || s->new_goal;
    bool need_plan = fresh;
    if (fresh) {  // ReplannerBase.h:48-55
        s->num_nodes_updated = s->num_nodes_expanded = 0;
        for (ShardMember &m : s->local) {
            SH_TRY(s, cudaSetDevice(m.device));
            tasi_planner *p = m.p;
            const int lr = s->have_goal_elem ? shard_local_row(s, m, s->goal_ex) : -1;
            p->have_goal_elem = lr >= 0 && s->goal_ey >= 0 && s->goal_ey < s->EY;
            p->goal_ex = lr; p->goal_ey = s->goal_ey; p->global_row0 = 0;
            p->goal_set = true;
            rc = init_field(p);
            if (rc != TASI_OK) return sh_sub(s, m, rc);
            p->new_goal = p->initialize_search = false;
        }
        rc = shard_reset_ctl(s, true, 0.0f, 0.0f);  // the counters of the whole step start here
        if (rc != TASI_OK) return rc;
        s->patch_pending = false;
        s->last_seed_count = 0;
    } else if (s->new_start) {  // ReplannerBase.h:56-59 -> update()
        s->new_start = false;
        for (ShardMember &m : s->local) {  // per-step counters
            SH_TRY(s, cudaSetDevice(m.device));
            SH_TRY(s, cudaMemsetAsync(&m.p->ctl->visits, 0, offsetof(DevCtl, timeout_ns) - offsetof(DevCtl, visits),
                                      m.p->stream));
            SH_TRY(s, cudaMemsetAsync(&m.p->ctl->far_parked, 0, offsetof(DevCtl, n_changed) - offsetof(DevCtl, far_parked),
                                      m.p->stream));
        }
        if (s->patch_pending) {
            std::vector<double> seeds, raised;
            for (ShardMember &m : s->local) {
                SH_TRY(s, cudaSetDevice(m.device));
                DevCtl hc;
                rc = read_ctl(m.p, &hc);
                if (rc != TASI_OK) return sh_sub(s, m, rc);
                seeds.push_back((double) hc.n_seed_elems);
                raised.push_back((double) hc.n_raised);
            }
            double n_seed = 0, n_raised = 0;
            rc = shard_allreduce_host(s, seeds.data(), false, &n_seed);
            if (rc != TASI_OK) return rc;
            rc = shard_allreduce_host(s, raised.data(), false, &n_raised);
            if (rc != TASI_OK) return rc;
            s->last_seed_count = (uint64_t) n_seed;
            if (n_raised > 0) {
                rc = shard_invalidate_loop(s);
                if (rc != TASI_OK) return rc;
            }
            for (ShardMember &m : s->local) {
                SH_TRY(s, cudaSetDevice(m.device));
                SH_TRY(s, cudaMemsetAsync(&m.p->ctl->n_changed, 0, sizeof(DevCtl) - offsetof(DevCtl, n_changed), m.p->stream));
            }
            s->patch_pending = false;
        }
        s->num_nodes_updated = s->last_seed_count;
        s->stats.seed_elems = s->last_seed_count;
        need_plan = s->num_nodes_updated > 0;
        rc = shard_sync_all(s);
        if (rc != TASI_OK) return rc;
    }
    const double t1 = now_ms();
    if (need_plan) {
        rc = shard_relax_loop(s);
        if (rc != TASI_OK) return rc;
    }
    for (ShardMember &m : s->local) {
        SH_TRY(s, cudaSetDevice(m.device));
        SH_TRY(s, cudaEventRecord(m.ev_t1, m.p->stream));
    }
    rc = shard_sync_all(s);
    if (rc != TASI_OK) return rc;
    const double t2 = now_ms();
    s->sstats.wall_ms = (float) (t2 - t0);
    rc = shard_collect_stats(s, fresh);
    if (rc != TASI_OK) return rc;
    s->num_nodes_expanded = need_plan ? s->stats.elem_lowered : 0;
    s->new_goal = s->initialize_search = false;
    s->stats.settled_below = INFINITY;
    if (u_ms) *u_ms = (float) (t1 - t0);
    if (p_ms) *p_ms = (float) (t2 - t1);
    if (n_updated) *n_updated = s->num_nodes_updated;
    if (n_expanded) *n_expanded = s->num_nodes_expanded;
    return TASI_OK;
}

// ReplannerBase::patch_map on the sharded grid: every band that stores some of the patch's cell rows (as owner or as the
// halo copy above its first node row) takes them; changes are counted where the row is owned, elements are seeded by
// the band that owns them.
int shard_patch_map(tasi_shard *s, const uint8_t *host_patch, int x, int y, int w, int h, uint64_t *n_changed) {
    if (!host_patch || s->initialize_graph) return TASI_ERR_ARG;
    if (x < 0 || y < 0 || w <= 0 || h <= 0 || (int64_t) x + h > s->L || (int64_t) y + w > s->W)
        return sh_fail(s, TASI_ERR_ARG, "patch outside the map");  // the reference asserts (Graph.cpp:38-41)
    const bool node = sh_node(s);
    std::vector<double> changed;
    for (ShardMember &m : s->local) {
        tasi_planner *p = m.p;
        SH_TRY(s, cudaSetDevice(m.device));
        SH_TRY(s, cudaMemsetAsync(&p->ctl->n_changed, 0, sizeof(unsigned long long), p->stream));
        for (int g : m.bands) {
            const ShardBand &b = s->bands[g];
            int64_t c0, c1;
            shard_cell_rows(s, b, &c0, &c1);
            const int64_t lo = std::max<int64_t>(c0, x), hi = std::min<int64_t>(c1, (int64_t) x + h);
            if (hi <= lo) continue;
            const int hh = (int) (hi - lo);
            const size_t n = (size_t) w * hh;
            if (n > p->patch_cap) {
                SH_TRY(s, cudaStreamSynchronize(p->stream));
                cudaFree(p->patch_buf); cudaFree(p->chg_buf);
                p->patch_buf = p->chg_buf = nullptr; p->patch_cap = 0;
                const size_t cap = std::max(n, (size_t) w * std::min<int64_t>(h, (int64_t) s->R + 1));
                SH_TRY(s, cudaMalloc(&p->patch_buf, cap));
                SH_TRY(s, cudaMalloc(&p->chg_buf, cap));
                p->patch_cap = cap;
            }
            SH_TRY(s, cudaMemcpyAsync(p->patch_buf, host_patch + (size_t) (lo - x) * w, n, cudaMemcpyHostToDevice, p->stream));
            GridDesc d = make_desc(p);
            const int lx = (int) (b.row0 + (lo - b.a));  // local row of the first patched cell row
            // rows counted: the cell rows this band OWNS (cell row c belongs to the band owning element row c)
            const int cnt_lo = (int) (b.row0 + std::max<int64_t>(lo, b.a) - b.a), cnt_hi = (int) (b.row0 + (hi - b.a));
            // elements seeded: those of this band
            const int seed_lo = b.row0, seed_hi = b.row0 + b.nrows;
            const int blocks = (int) std::min<size_t>((n + 255) / 256, (size_t) p->num_sms * 16);
            patch_apply_kernel<<<blocks, 256, 0, p->stream>>>(d, p->patch_buf, p->chg_buf, lx, y, w, hh, cnt_lo, cnt_hi);
            switch (s->kind) {
                case TASI_FD: patch_seed_kernel<TASI_KIND_FD><<<blocks, 256, 0, p->stream>>>(d, p->chg_buf, lx, y, w, hh, seed_lo, seed_hi); break;
                case TASI_DFM: patch_seed_kernel<TASI_KIND_DFM><<<blocks, 256, 0, p->stream>>>(d, p->chg_buf, lx, y, w, hh, seed_lo, seed_hi); break;
                default: patch_seed_kernel<TASI_KIND_MFD><<<blocks, 256, 0, p->stream>>>(d, p->chg_buf, lx, y, w, hh, seed_lo, seed_hi); break;
            }
            SH_TRY(s, cudaGetLastError());
            SH_TRY(s, cudaStreamSynchronize(p->stream));  // patch_buf is reused by the next band
        }
        DevCtl hc;
        int rc = read_ctl(p, &hc);
        if (rc != TASI_OK) return sh_sub(s, m, rc);
        changed.push_back((double) hc.n_changed);
    }
    (void) node;
    double total = 0;
    int rc = shard_allreduce_host(s, changed.data(), false, &total);
    if (rc != TASI_OK) return rc;
    if (n_changed) *n_changed = (uint64_t) total;
    s->patch_pending = true;
    return TASI_OK;
}

// Path walks on the sharded grid: the rows a walk of max_steps steps can reach are gathered onto one device and walked
// there (a step crosses one cell edge, so max_steps + 8 rows either side of the start cover it).
int shard_extract_paths(tasi_shard *s, const float *starts_xy, const float *goals_xy, const int *row_ranges, int n,
                        int max_steps, int lookahead, int allow_indirect, float *xy, int cap_points, int *n_points,
                        float *step_costs, int cap_costs, int *n_costs, float *totals, int *status, float *e_ms);

void shard_destroy(tasi_shard *s) {
    if (!s) return;
    for (ShardMember &m : s->local) {
        cudaSetDevice(m.device);
        if (m.p && m.p->stream) cudaStreamSynchronize(m.p->stream);
        if (m.comm && nccl_api().CommDestroy) nccl_api().CommDestroy(m.comm);
        shard_free_member_buffers(m);
        cudaFree(m.sctl); cudaFree(m.sum_dev); cudaFree(m.fc); cudaFree(m.topo);
        cudaFreeHost(m.flags_host);
        if (m.ev_packed) cudaEventDestroy(m.ev_packed);
        if (m.ev_red) cudaEventDestroy(m.ev_red);
        if (m.ev_t0) cudaEventDestroy(m.ev_t0);
        if (m.ev_t1) cudaEventDestroy(m.ev_t1);
        for (cudaEvent_t e : m.ev_flag)
            if (e) cudaEventDestroy(e);
        if (m.p) tasi_destroy(m.p);
    }
    delete s;
}

int shard_init_member(tasi_shard *s, ShardMember &m) {
    int rc = tasi_create(s->kind, s->level, m.device, &m.p);
    if (rc != TASI_OK) return sh_fail(s, rc, "tasi_create failed on device " + std::to_string(m.device));
    SH_TRY(s, cudaSetDevice(m.device));
    SH_TRY(s, cudaMalloc(&m.sctl, sizeof(ShardCtl)));
    SH_TRY(s, cudaMemset(m.sctl, 0, sizeof(ShardCtl)));
    SH_TRY(s, cudaMalloc(&m.sum_dev, 16));
    SH_TRY(s, cudaMalloc(&m.fc, sizeof(FusedCtl)));
    SH_TRY(s, cudaMemset(m.fc, 0, sizeof(FusedCtl)));
    SH_TRY(s, cudaMalloc(&m.topo, sizeof(FusedTopo)));
    SH_TRY(s, cudaHostAlloc(&m.flags_host, sizeof(int) * SHARD_FLAG_RING, cudaHostAllocMapped));
    memset(m.flags_host, 0, sizeof(int) * SHARD_FLAG_RING);
    SH_TRY(s, cudaEventCreateWithFlags(&m.ev_packed, cudaEventDisableTiming));
    SH_TRY(s, cudaEventCreateWithFlags(&m.ev_red, cudaEventDisableTiming));
    SH_TRY(s, cudaEventCreate(&m.ev_t0));
    SH_TRY(s, cudaEventCreate(&m.ev_t1));
    for (cudaEvent_t &e : m.ev_flag) SH_TRY(s, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    // the band factor follows band_width()'s rule on tiles per GPU / world (every rank holds the same share, so the ranks
    // agree; the fused mode takes rank 0's width, the iteration mode the widest)
    return TASI_OK;
}

}  // namespace
