// Tiled active-list solver for the cost-to-go field of FD*, MS-DFM and MFD* (sm_100a).
//
// Replaces the reference's open-list expansion — plan() in FieldDStar/FieldDPlanner_impl.h:23-116,
// DynamicFastMarching/DynamicFastMarching_impl.h:13-104, ShiftedGridFastMarching/ShiftedGridPlanner_impl.h:17-170
// with its PriorityQueue (ProjectToolkit/include/PriorityQueue.h) and ExpandedMap
// (ProjectToolkit/include/ExpandedMap.h) — by a label-correcting solve to the same fixed point:
//
//   * the element grid is cut into 32x32 tiles; a persistent grid of CTAs pulls active tiles from a device
//     ticket queue, stages the tile (+1 halo ring) and the cell costs around it in shared memory, and relaxes it
//     to tile-local convergence (relax_tile: one 1x4 strip per thread with a register copy of its neighbourhood),
//     accepting only decreases;
//   * a tile whose border changed activates the neighbour tiles that read that border;
//   * the same machinery run in "invalidate" mode is phase 1 of a replan: an element whose stored value is
//     no longer supported by its neighbours under the new costs is reset to INF (what the reference's
//     under-consistent branch does lazily, FieldDPlanner_impl.h:98-112), then phase 2 re-relaxes.
//
// Tile states: 0 idle, DIRTY queued, RUN being processed, RUN|DIRTY halo changed while processing (the
// same CTA reloads and continues). A tile is in the queue at most once, so a ring of 2 x max(tiles, CTAs) + 1
// slots (rounded up to a power of two, alloc_grid) can never overflow and a pusher never has to wait for its
// slot; poppers take tickets and wait for their slot, `pending` (queued + running tiles) reaching 0 ends the
// kernel. A wall-clock watchdog on GLOBAL progress (no push anywhere for timeout_ns) turns any would-be hang
// into an error code.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "tasi_update.cuh"

namespace tasi {

constexpr int TS = 32;         // tile side (elements)
constexpr int GP = 35;         // shared-memory pitch of the value tile (34 used; odd => row and column
                               // sweeps are both bank-conflict free)
constexpr int CP = 35;         // shared-memory pitch of the cost tile (33 used for node planners)
constexpr int NTHREADS = 256;  // one thread per 1x4 strip of the tile
constexpr int ST_DIRTY = 1;
constexpr int ST_RUN = 2;
constexpr int ST_FAR = 4;      // queued, but parked in the far pile (key >= current threshold)
constexpr int ST_DEFER = 8;    // popped, found an upwind neighbour still at work, set aside until that one leaves (queued, in
                               // no ring, still counted as pending)
constexpr int ST_MASK = 15;    // the four state bits; bits ST_SHIFT.. of the same word hold the IN_* direction
constexpr int ST_SHIFT = 8;    // bits (which halo sides changed since the tile was last staged)
constexpr int INF_BITS = 0x7f800000;
constexpr int GMARGIN = 32;    // columns of margin left of element/cell column 0 (keeps rows 128B aligned)

constexpr int QB_MAX = 64;     // size of the ring counter arrays of the bucketed near queue (at most 62 rings: two per lane of
                               // the popping warp, lane 31 watches the plain ring)

struct DevCtl {
    unsigned long long head, tail;  // ticket counters of the work queue
    int pending;                    // tiles queued or running
    int done;                       // set when pending drops to 0 (or on error)
    int error;                      // 1 = watchdog fired
    int thr_bits;                   // float bits of the near/far threshold: tiles keyed below it run now
    unsigned long long visits, rounds, lowered, pushes, invalidated;
    unsigned long long timeout_ns;
    int far_sel;                    // which of the two far lists takes new entries
    int far_count[2];
    int phases;                     // threshold advances
    int limit_bits;                 // band mode: the threshold never passes this value (float bits; INF = none)
    int far_min_bits;               // lowest parked key seen by the last advance (INF = nothing parked)
    int kick;                       // set by arm_solver: nothing queued but tiles are parked, CTA 0 starts by advancing
    unsigned long long far_parked;  // activations that went to the far pile
    unsigned long long evals;       // element evaluations
    unsigned long long cyc_pop, cyc_stage, cyc_relax, cyc_post;  // SM cycles summed over CTAs (thread 0's clock)
    unsigned long long cyc_advance;  // cycles the advancing CTA spent raising the threshold (all others wait)
    int focus_hit;                  // the launch ended because the focus elements were settled (work stays parked)
    int launches;                   // solver launches since the counters were last cleared (sharded loop)
    unsigned long long busy_ns;     // globaltimer span of the solver launches as CTA 0 saw them, summed
    // bucketed near queue (GridDesc::nb > 0): ticket counters of the key-ordered rings, one 128-byte line each
    alignas(128) unsigned long long bhead[QB_MAX];
    alignas(128) unsigned long long btail[QB_MAX];
    alignas(128) int bcount[QB_MAX];  // published entries of each ring not yet claimed by a popper
    unsigned long long deferred;      // tiles set aside by the upwind deferral
    // sliding key window (GridDesc::slide): absolute index of the slice the lowest ring stands for, the prefix of the
    // (append-only) far list already dealt with, and the lock of whoever scans it
    int abase, far_scanned, scan_lock;
    unsigned long long slide_scans;
    alignas(128) int brun[QB_MAX];    // sliding window: tiles being visited that were claimed from each ring
    // patch bookkeeping
    unsigned long long n_changed, n_raised, n_lowered_cells, n_seed_elems;
    unsigned long long halo_lowered;
};

struct GridDesc {
    float *g;              // element (0,0); rows -1..TX*32 and GMARGIN columns either side exist
    int64_t gpitch;        // elements per row
    uint8_t *cost;         // cell (0,0); rows -1..TX*32 and margins exist, margin bytes are 255
    int64_t cpitch;        // bytes per row
    int EX, EY;            // element-grid extents owned by this planner (band)
    int TX, TY;            // tiles
    int L, W;              // cell rows / columns stored (band-local)
    int thr;               // occupancy threshold byte: cost >= thr is an obstacle (Graph.cpp:262-268)
    int pin_x, pin_y;      // goal element in local coordinates (outside the grid = none)
    const int2 *xgoals;    // further goal elements (local coordinates) of a multi-goal plan, e.g. a stack of
    int n_xgoals;          // independent maps solved in one launch (tasi_set_goals)
    int *state;            // TX*TY tile states
    int *queue;            // ring of tile id + 1, 0 = empty
    unsigned qmask;        // ring size - 1
    uint8_t *seed;         // TX*TY: tiles to enqueue before the next solver launch
    int *key;              // TX*TY: float bits of the lowest value that entered the tile since its last visit
    int *far;              // 2 * far_cap tile ids: the far pile, double-buffered
    int far_cap;
    float delta;           // width of the value band processed per phase
    DevCtl *ctl;
    float inval_tol;       // relative slack of the invalidate test (0 for FD*/MFD*)
    // focused plan (the reference's end_condition, FieldDPlanner_impl.h:225-256, DynamicFastMarching_impl.h:315-320):
    // the launch ends once every value up to the highest finite one among these elements (+ slack) is final
    int n_focus;           // 0 = solve to convergence
    int2 focus[4];         // the start cell's corner nodes / the start cell, local coordinates
    float focus_slack;
    // a stack of row bands in one grid (sharded planner, tasi_shard.cuh): every band_rt tile rows are followed by one
    // DEAD tile row that is never scheduled; it holds the halo rows of the bands above and below it. 0 = no stack.
    int band_rt;
    // fused multi-GPU mode (see "fused multi-GPU mode" below): nullptr = off
    const struct FusedTopo *ft;
    int *inbox;            // tiles parked here by OTHER ranks since the last phase boundary (this rank drains it)
    int inbox_cap;
    int fmode;             // fused launch: 0 relax (value-ordered), 1 invalidate (no order: everything parked is released)
    int rel_passes;        // slices of rising key in which a threshold advance releases the parked tiles (TASI_REL_PASSES)
    // bucketed near queue (TASI_BUCKETS = nb > 0, relax launches of an unsharded grid): tiles that are due now go to one
    // of nb rings by key — ring b holds the keys of slice b of the current value band [threshold - delta, threshold) —
    // and a CTA looking for work takes from the lowest non-empty ring (after the plain ring `queue`, which keeps the
    // entries without an order: overflow fall-backs). No barrier between the rings: an ordered-by-integer-metric
    // worklist inside each threshold phase.
    int nb;
    int *bq;               // nb rings of bqmask + 1 slots (a power of two > tiles: a tile is queued at most once)
    unsigned bqmask;
    float inv_bw;          // nb / delta
    // upwind deferral (TASI_DEFER, with the bucketed queue only): see "upwind deferral" below. margin < 0: off
    float defer_margin;
    int *rkey;             // TX*TY: the key a running tile was claimed with
    int chain;             // TASI_CHAIN: a CTA goes on with the lowest-keyed neighbour it has just activated (see solver_kernel)
    // sliding key window (TASI_SLIDE; see "sliding key window" below): ring = (key / bw) mod nb, no threshold phases
    int slide;
    float bw;              // ring width = delta / nb
};

// TMA staging (solver instances with TMA = 1, TASI_TMA=1): the two CUtensorMap (128 opaque bytes each) the kernel takes
// as a __grid_constant__ parameter: g = the padded value array (float32, box 40 x 34), c = the padded cost array (uint8,
// box 48 x 33). All-zero for the instances that stage with vector loads.
struct TmaMaps {
    alignas(64) unsigned long long g[16];
    alignas(64) unsigned long long c[16];
};

// tile row `tx` of a band stack is a dead row (halo storage between two bands)
__device__ __forceinline__ bool dead_tile_row(const GridDesc &d, int tx) {
    return d.band_rt > 0 && (tx % (d.band_rt + 1)) == d.band_rt;
}

// Release/acquire fence at device scope. __threadfence() compiles to the sequentially consistent MEMBAR.SC.GPU;
// the queue protocol only needs "my (and, after a CTA barrier, my CTA's) earlier writes are visible before my next
// atomic", which the acq_rel fence gives (cumulative over the barrier) at a fraction of the latency.
__device__ __forceinline__ void fence_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }

__device__ __forceinline__ unsigned long long globaltimer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// ---- work queue (called by one thread per CTA) ------------------------------------------------------------
__device__ __forceinline__ void q_fail(const GridDesc &d) {
    atomicExch(&d.ctl->error, 1);
    atomicExch(&d.ctl->done, 1);
}

__device__ __forceinline__ int cur_thr_bits(const GridDesc &d);

// ring of the bucketed near queue for key `kbits` (-1 = the plain ring)
__device__ __forceinline__ int bucket_of(const GridDesc &d, int kbits) {
    if (d.nb <= 0) return -1;
    if (d.slide) {  // absolute, cyclic: slice of the key, clamped into the window [base, base + nb)
        const int a = *(volatile int *) &d.ctl->abase;
        int si = (kbits == INF_BITS) ? a : (int) (__int_as_float(kbits) * d.inv_bw);
        si = max(si, a);
        si = min(si, a + d.nb - 1);
        return si % d.nb;
    }
    const int tb = cur_thr_bits(d);
    if (tb == INF_BITS || kbits == INF_BITS) return 0;
    const float x = (__int_as_float(kbits) - (__int_as_float(tb) - d.delta)) * d.inv_bw;
    const int b = x > 0.0f ? (int) x : 0;
    return b < d.nb ? b : d.nb - 1;
}

// push with a key: into the ring of its slice of the band (bucketed near queue), else the plain ring
__device__ inline void q_push_key(const GridDesc &d, int tile, int kbits) {
    const int b = bucket_of(d, kbits);
    if (b < 0) {
        atomicAdd(&d.ctl->pending, 1);
        unsigned long long t = atomicAdd(&d.ctl->tail, 1ULL);
        atomicExch(&d.queue[(unsigned) t & d.qmask], tile + 1);
        return;
    }
    atomicAdd(&d.ctl->pending, 1);
    unsigned long long t = atomicAdd(&d.ctl->btail[b], 1ULL);
    atomicExch(&d.bq[(size_t) b * (d.bqmask + 1u) + ((unsigned) t & d.bqmask)], tile + 1);
    // no fence between the slot and the count: a popper that wins a unit of the count takes the next head ticket and
    // waits for THAT slot to be filled — by this pusher or by one that already holds the ticket — so the order in which
    // the two stores become visible does not matter (one L2 round trip less per push)
    atomicAdd(&d.ctl->bcount[b], 1);
}

// back into its ring a tile that is already counted as pending (a deferred tile that is released)
__device__ inline void q_repush(const GridDesc &d, int tile) {
    const int b = max(bucket_of(d, *(volatile int *) &d.key[tile]), 0);
    unsigned long long t = atomicAdd(&d.ctl->btail[b], 1ULL);
    atomicExch(&d.bq[(size_t) b * (d.bqmask + 1u) + ((unsigned) t & d.bqmask)], tile + 1);
    __threadfence();
    atomicAdd(&d.ctl->bcount[b], 1);
}

__device__ inline void q_push(const GridDesc &d, int tile) {
    if (d.nb > 0) {  // bucketed near queue: by the tile's current key
        q_push_key(d, tile, *(volatile int *) &d.key[tile]);
        return;
    }
    atomicAdd(&d.ctl->pending, 1);
    unsigned long long t = atomicAdd(&d.ctl->tail, 1ULL);
    // The ring has 2 x max(tiles, CTAs) + 1 slots and a tile is queued at most once, so the slot of this ticket was
    // emptied by its previous owner long ago (CTAs are not preempted individually): publish without waiting for a reply.
    atomicExch(&d.queue[(unsigned) t & d.qmask], tile + 1);
}

#ifndef Q_SLEEP_MAX_NS
#define Q_SLEEP_MAX_NS 256
#endif
__device__ inline int q_pop(const GridDesc &d) {
    unsigned long long t = atomicAdd(&d.ctl->head, 1ULL);
    volatile int *slot = &d.queue[(unsigned) t & d.qmask];
    volatile int *done = &d.ctl->done;
    unsigned long long t0 = 0, tail0 = 0;
    unsigned ns = 32, polls = 0;
    for (;;) {
        int v = *slot;
        if (v != 0) {
            *slot = 0;
            return v - 1;
        }
        // the slot is this ticket's own word; `done` shares its line with the queue counters, so it is looked at
        // on every fourth poll only and the back-off can stay short (a waiting CTA sits on the critical path)
        if ((++polls & 3u) == 0 && *done) return -1;
        if (t0 == 0) {
            t0 = globaltimer_ns();
            tail0 = *(volatile unsigned long long *) &d.ctl->tail;
        } else if (globaltimer_ns() - t0 > d.ctl->timeout_ns) {
            // the watchdog fires on lack of GLOBAL progress: if anything was pushed since this wait began (other CTAs
            // are working; or the device is shared and some CTAs of this launch were not resident yet), wait on
            const unsigned long long tl = *(volatile unsigned long long *) &d.ctl->tail;
            if (tl == tail0) { q_fail(d); return -1; }
            tail0 = tl;
            t0 = globaltimer_ns();
        }
        __nanosleep(ns);
        if (ns < Q_SLEEP_MAX_NS) ns <<= 1;
    }
}

// ---- near / far scheduling -------------------------------------------------------------------------------------
// Every activation carries a key: the lowest changed value the tile is about to read. Tiles keyed below the
// current threshold go to the work queue ("near"); the others are parked in the far pile until the queue has
// drained and the threshold advances by `delta` (a tile-level Delta-stepping order: without it a
// label-correcting solve revisits each tile dozens of times, with it a few).
// near / far boundary. Sliding window: the window of nb rings above the base slice.
__device__ __forceinline__ int cur_thr_bits(const GridDesc &d) {
    if (d.slide) return __float_as_int((float) (*(volatile int *) &d.ctl->abase + d.nb) * d.bw);
    return *(volatile int *) &d.ctl->thr_bits;
}

// park a tile whose state already carries ST_FAR; falls back to the queue if the pile is full
__device__ inline void far_append(const GridDesc &d, int tile) {
    if (d.slide) {
        // sliding window: one append-only list of tile + 1 (0 = slot taken, not written yet). The window may pass the
        // key while the tile is being parked, and a scan may already be beyond this slot: look again after the append
        // and take the tile back if it has become due (whoever clears ST_FAR owns the push).
        const int slot = atomicAdd(&d.ctl->far_count[0], 1);
        if (slot < d.far_cap) {
            *(volatile int *) &d.far[slot] = tile + 1;
            __threadfence();
            atomicAdd(&d.ctl->far_parked, 1ULL);
            const int k = *(volatile int *) &d.key[tile];
            if (k < cur_thr_bits(d) && (atomicAnd(&d.state[tile], ~ST_FAR) & ST_FAR)) {
                q_push_key(d, tile, k);
                atomicAdd(&d.ctl->pushes, 1ULL);
            }
        } else {
            atomicAnd(&d.state[tile], ~ST_FAR);
            q_push(d, tile);
            atomicAdd(&d.ctl->pushes, 1ULL);
        }
        return;
    }
    const int sel = *(volatile int *) &d.ctl->far_sel;
    const int slot = atomicAdd(&d.ctl->far_count[sel], 1);
    if (slot < d.far_cap) {
        d.far[sel * d.far_cap + slot] = tile;
        __threadfence();
        atomicAdd(&d.ctl->far_parked, 1ULL);
    } else {
        atomicAnd(&d.state[tile], ~ST_FAR);
        q_push(d, tile);
        atomicAdd(&d.ctl->pushes, 1ULL);
    }
}

// hold: if the tile was idle and is due now, do NOT queue it but return true — the caller queues it or runs it itself
__device__ inline bool activate(const GridDesc &d, int tile, int kbits, int dirbit, bool hold = false) {
    const int oldk = atomicMin(&d.key[tile], kbits);  // may land after the state bit: a stale low key only makes the tile look due
    const bool near = kbits < cur_thr_bits(d);
    // one atomic both flags the tile and records which of its halo sides changed (bits ST_SHIFT..)
    const int old = atomicOr(&d.state[tile], ST_DIRTY | (dirbit << ST_SHIFT));
    const int st = old & ST_MASK;
    if ((st & ST_DEFER) && kbits < oldk) {
        // set aside with a higher key: what it waited for may now wait for IT, so it goes back into the rings
        if (atomicAnd(&d.state[tile], ~ST_DEFER) & ST_DEFER) q_repush(d, tile);
        return false;
    }
    if (st == 0) {  // it was idle: ours to queue
        if (near) {
            if (hold) return true;
            q_push_key(d, tile, kbits);
            atomicAdd(&d.ctl->pushes, 1ULL);
        } else {
            atomicOr(&d.state[tile], ST_FAR);
            far_append(d, tile);
        }
    } else if ((st & ST_FAR) && near) {
        // parked with a higher key: promote (whoever clears ST_FAR owns the push; the pile entry goes stale)
        if (atomicAnd(&d.state[tile], ~ST_FAR) & ST_FAR) {
            q_push_key(d, tile, kbits);
            atomicAdd(&d.ctl->pushes, 1ULL);
        }
    }
    return false;
}

// ---- sliding key window (TASI_SLIDE) -----------------------------------------------------------------------------------
// With the rings the threshold phases are only the trade between a wide band (tiles far up the band run on provisional
// borders) and many phase barriers (CTAs wait at each). Here there are no phases: ring = (key / bw) mod nb, absolute and
// cyclic; the window is the nb slices above the BASE slice, which follows the lowest non-empty ring (a popper that finds
// its tile at offset o > 0 raises the base by o with atomicMax; a push below the base is clamped into the base ring); the
// near / far boundary is base + nb and moves with it, so a tile activated for the first time — its key lies a tile width
// or two above the front — is never parked. What lies beyond the window goes to an append-only far list; the popper
// that moves the base across a multiple of SLIDE_SCAN_EVERY slices scans the list (one scanner at a time) and releases
// what the window has reached. Hand-over of a parked tile (appender that finds the window has passed it, scanner,
// promoting activation): whoever clears ST_FAR owns the push. Quiescence (pending == 0) only happens when everything
// left is parked beyond the window: the base jumps to the lowest parked key.
constexpr int SLIDE_SCAN_EVERY = 4;

// One pass over the unscanned part of the far list by the 32 lanes of a warp. jump: the grid is quiescent — first move the
// base to the slice of the lowest parked key. Returns (to every lane) the number of live entries left parked.
__device__ inline int slide_scan_warp(const GridDesc &d, int lane, bool jump) {
    DevCtl *c = d.ctl;
    const unsigned full = 0xffffffffu;
    // one scanner at a time
    int got = 0;
    if (lane == 0) {
        if (jump) { while (atomicCAS(&c->scan_lock, 0, 1) != 0) __nanosleep(64); got = 1; }
        else got = atomicCAS(&c->scan_lock, 0, 1) == 0;
    }
    got = __shfl_sync(full, got, 0);
    if (!got) return -1;
    __threadfence();
    const int s0 = *(volatile int *) &c->far_scanned;
    const int n = min(*(volatile int *) &c->far_count[0], d.far_cap);
    // first slot that is taken but not written yet: the prefix cannot pass it
    int stop = n;
    for (int i = s0 + lane; i < n; i += 32)
        if (*(volatile int *) &d.far[i] == 0) { stop = i; break; }
    stop = __reduce_min_sync(full, stop);
    if (jump) {
        int m = INF_BITS;
        for (int i = s0 + lane; i < n; i += 32) {
            const int e = *(volatile int *) &d.far[i];
            if (e > 0 && (*(volatile int *) &d.state[e - 1] & ST_FAR)) m = min(m, *(volatile int *) &d.key[e - 1]);
        }
        m = __reduce_min_sync(full, m);
        if (lane == 0) {
            c->far_min_bits = m;
            if (m != INF_BITS) atomicMax(&c->abase, (int) (__int_as_float(m) * d.inv_bw));
            atomicAdd(&c->phases, 1);
            __threadfence();
        }
        __syncwarp();
    }
    const int thr = cur_thr_bits(d);
    const bool relall = jump && *(volatile int *) &c->far_min_bits == INF_BITS;  // only tiles without a finite key are left
    int left = 0;
    for (int i = s0 + lane; i < n; i += 32) {
        const int e = *(volatile int *) &d.far[i];
        if (e <= 0) continue;
        const int t = e - 1;
        if (!(*(volatile int *) &d.state[t] & ST_FAR)) continue;  // stale: promoted or taken back meanwhile
        const int k = *(volatile int *) &d.key[t];
        if (k < thr || relall) {
            if (atomicAnd(&d.state[t], ~ST_FAR) & ST_FAR) {
                q_push_key(d, t, k);
                atomicAdd(&c->pushes, 1ULL);
            }
        } else {
            ++left;
            if (i < stop) {  // keep it: behind the new prefix it would never be looked at again, so it moves to the tail
                const int slot = atomicAdd(&c->far_count[0], 1);
                if (slot < d.far_cap) *(volatile int *) &d.far[slot] = e;
                else if (atomicAnd(&d.state[t], ~ST_FAR) & ST_FAR) {  // list full: run it now
                    q_push_key(d, t, k);
                    atomicAdd(&c->pushes, 1ULL);
                }
            }
        }
    }
    left = __reduce_add_sync(full, left);
    __threadfence();
    if (lane == 0) {
        c->far_scanned = stop;
        atomicAdd(&c->slide_scans, 1ULL);
        __threadfence();
        atomicExch(&c->scan_lock, 0);
    }
    __syncwarp();
    return left;
}

// The queue has drained and no tile is running: raise the threshold to (lowest parked key + delta) and move
// the parked tiles below it into the queue. Runs on ONE CTA while all others wait in q_pop, so the pile is
// not appended to until the first push below. Returns with s_flag = 1 if nothing is left anywhere.
// Returns true (to every thread) if the lowest parked key is at or beyond the launch's limit: nothing is released,
// the pile stays as it is and the launch ends (band mode: the host exchanges halos and raises the limit).
__device__ inline bool advance_threshold(const GridDesc &d, int *s_red, int tid) {
    DevCtl *c = d.ctl;
    const int nthreads = blockDim.x;  // 256 or 512 (the strip width of the kernel instance)
    const int cur = *(volatile int *) &c->far_sel;
    const int n = min(*(volatile int *) &c->far_count[cur], d.far_cap);
    const int *src = d.far + cur * d.far_cap;
    int m = INF_BITS;
    for (int i = tid; i < n; i += nthreads) {
        const int t = __ldcg(src + i);
        if (*(volatile int *) &d.state[t] & ST_FAR) m = min(m, *(volatile int *) &d.key[t]);
    }
    m = __reduce_min_sync(0xffffffffu, m);
    if ((tid & 31) == 0) s_red[tid >> 5] = m;
    __syncthreads();
    m = s_red[0];
    for (int w = 1; w < nthreads / 32; ++w) m = min(m, s_red[w]);
    const int lim = *(volatile int *) &c->limit_bits;
    if (tid == 0) c->far_min_bits = m;
    if (m != INF_BITS && m >= lim) return true;
    if (d.n_focus > 0 && m != INF_BITS) {
        // the queue is empty and every parked key is >= m: all values below m are final. Values only depend on
        // lower ones, so the focus elements are settled once the highest finite one lies below m.
        float top = -1.0f;
        for (int i = 0; i < d.n_focus; ++i) {
            const float v = __ldcg(d.g + (int64_t) d.focus[i].x * d.gpitch + d.focus[i].y);
            if (v < __int_as_float(INF_BITS)) top = fmaxf(top, v);
        }
        if (top >= 0.0f && top + d.focus_slack < __int_as_float(m)) {
            if (tid == 0) c->focus_hit = 1;
            return true;
        }
    }
    int nthr = INF_BITS;
    if (m != INF_BITS) {
        const float t = __int_as_float(m) + d.delta;
        nthr = __float_as_int(t);
        if (nthr <= m) nthr = m + 1;  // delta lost to rounding: still let the lowest tile through
        if (nthr > lim) nthr = lim;
    }
    if (tid == 0) {
        c->thr_bits = nthr;
        c->far_count[cur ^ 1] = 0;
        c->far_sel = cur ^ 1;  // new parkings (from the tiles pushed below) go to the other list
        atomicAdd(&c->phases, 1);
        __threadfence();
    }
    __syncthreads();
    int *dst = d.far + (cur ^ 1) * d.far_cap;
    // The released tiles enter the FIFO queue in REL_PASSES slices of rising key (lowest quarter of the band first), so
    // that the CTAs start on the tiles whose inputs are final and reach the upper part of the band after its upwind
    // neighbours have run once (fewer visits on provisional borders). d.rel_passes <= 1: pile order.
    const float lo = __int_as_float(m), width = (nthr == INF_BITS || m == INF_BITS) ? 0.0f : __int_as_float(nthr) - lo;
    const int passes = (width > 0.0f && d.rel_passes > 1) ? d.rel_passes : 1;
    for (int ps = 0; ps < passes; ++ps) {
        const int cut = (ps == passes - 1) ? nthr : __float_as_int(lo + width * (float) (ps + 1) / (float) passes);
        for (int i = tid; i < n; i += nthreads) {
            const int t = __ldcg(src + i);
            if (!(*(volatile int *) &d.state[t] & ST_FAR)) continue;  // stale entry (tile was promoted), or released by
                                                                     // an earlier pass
            const int k = *(volatile int *) &d.key[t];
            if (k < cut || nthr == INF_BITS) {
                if (atomicAnd(&d.state[t], ~ST_FAR) & ST_FAR) {
                    q_push_key(d, t, k);
                    atomicAdd(&c->pushes, 1ULL);
                }
            } else if (ps == passes - 1) {
                const int slot = atomicAdd(&c->far_count[cur ^ 1], 1);
                if (slot < d.far_cap) dst[slot] = t;
                else if (atomicAnd(&d.state[t], ~ST_FAR) & ST_FAR) q_push(d, t);
            }
        }
        __syncthreads();
    }
    if (tid == 0) c->far_count[cur] = 0;
    return false;
}

// The count of pending tiles reached 0 in this CTA (s_adv set by thread 0): raise the threshold until tiles run
// again, or end the launch (nothing parked, or everything parked lies beyond the limit).
__device__ inline void drain_and_advance(const GridDesc &d, int *s_red, int *s_adv, int tid) {
    __syncthreads();
    if (d.slide) {
        // sliding window: the count reached 0, so whatever is left is parked beyond the window — the base jumps to the
        // lowest parked key (slide_scan_warp with jump), until nothing is left anywhere
        while (*s_adv) {
            __syncthreads();
            if (tid == 0) atomicAdd(&d.ctl->pending, 1);  // hold the count while tiles are being pushed
            int left = 0;
            if (tid < 32) left = slide_scan_warp(d, tid, true);
            __syncthreads();
            if (tid == 0) {
                int more = 0;
                if (atomicSub(&d.ctl->pending, 1) == 1) {  // nothing was pushed (or all of it already finished)
                    if (left > 0) more = 1;
                    else { __threadfence(); atomicExch(&d.ctl->done, 1); }
                }
                if (*(volatile int *) &d.ctl->error) more = 0;
                *s_adv = more;
            }
            __syncthreads();
        }
        return;
    }
    while (*s_adv) {
        __syncthreads();
        const long long ca0 = clock64();
        if (tid == 0) atomicAdd(&d.ctl->pending, 1);  // hold the count while tiles are being pushed
        const bool stalled = advance_threshold(d, s_red, tid);
        __syncthreads();
        if (tid == 0) {
            int more = 0;
            if (atomicSub(&d.ctl->pending, 1) == 1) {
                // nothing was pushed (or all of it already finished)
                const int sel = *(volatile int *) &d.ctl->far_sel;
                if (!stalled && *(volatile int *) &d.ctl->far_count[sel] > 0) more = 1;
                else { __threadfence(); atomicExch(&d.ctl->done, 1); }
            }
            if (*(volatile int *) &d.ctl->error) more = 0;
            *s_adv = more;
            atomicAdd(&d.ctl->cyc_advance, (unsigned long long) (clock64() - ca0));
        }
        __syncthreads();
    }
}

// ---- fused multi-GPU mode ---------------------------------------------------------------------------------------------
// One persistent solver per GPU, all running at the same time on the bands of ONE grid (tasi_shard.cuh), with no outer
// iteration: a visit of a tile on a band border stores its changed boundary row straight into the halo row of the rank
// that owns the band across the edge (peer-mapped memory over NVLink) and activates that rank's tiles with system-scope
// atomics on ITS tile states, keys and work queue — the transfer is part of the tile visit, tile by tile. The value
// order of the schedule is global: thresholds advance in lock-step. Rank 0's CTA 0 is the coordinator; it sees the grid
// quiescent when `active` (the number of ranks with pending tiles, kept exact by +1 BEFORE any 0 -> 1 transition of a
// rank's pending count and -1 AFTER any 1 -> 0) is 0, asks every rank for the lowest key of its pile (MIN), and answers
// with the new threshold (RELEASE) or the end of the launch (DONE). Commands and replies are single words in each
// other's memory; every rank's CTA 0 looks at its command word whenever it waits for a tile.
constexpr int FUSED_MAX_RANKS = 16;
constexpr int FOP_MIN = 1, FOP_RELEASE = 2, FOP_DONE = 3;

struct FusedCtl {
    int cmd;                    // (seq << 2) | op, written by the coordinator
    int cmd_thr_bits;           // RELEASE: the new threshold
    int seen_seq;               // seq of the last command this rank served (persists across launches)
    int seq;                    // coordinator: last sequence number issued (persists across launches)
    int active;                 // coordinator: ranks with pending tiles (+ pushes in flight)
    int inbox_count;            // entries of this rank's inbox
    unsigned long long rx;      // remote activations received (statistics)
    unsigned long long tx;      // remote activations sent (statistics)
    int ack[FUSED_MAX_RANKS];   // coordinator: ack[r] = seq of the last command rank r completed
    int rmin[FUSED_MAX_RANKS];  // coordinator: rank r's reply to MIN (float bits)
};

struct FusedTopo {
    int world, rank, parts, mode;
    GridDesc nbr[2];                // [0] the rank holding the bands above ours, [1] below; pointers valid from this device
    FusedCtl *fc[FUSED_MAX_RANKS];  // every rank's control block (peer-mapped)
};

__device__ __forceinline__ void fence_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }
__device__ __forceinline__ int ld_sys(const int *p) { return *(const volatile int *) p; }

// the pending count of this rank is about to leave / has just reached 0 (see `active` above)
__device__ __forceinline__ void fused_rank_wakes(const GridDesc &d) { atomicAdd_system(&d.ft->fc[0]->active, 1); }
__device__ __forceinline__ void fused_rank_idles(const GridDesc &d) { atomicAdd_system(&d.ft->fc[0]->active, -1); }

// activate() on a tile of another rank `r` (its GridDesc, peer-mapped). `near` is judged by OUR threshold: thresholds
// are global in this mode.
__device__ inline void activate_remote(const GridDesc &d, const GridDesc &r, FusedCtl *rfc, int tile, int kbits, int dirbit) {
    const int oldk = atomicMin_system(&r.key[tile], kbits);
    const bool near = kbits < cur_thr_bits(d);
    const int old = atomicOr_system(&r.state[tile], ST_DIRTY | (dirbit << ST_SHIFT));
    const int st = old & ST_MASK;
    if ((st & ST_DEFER) && kbits < oldk) {
        // set aside over there with a higher key (upwind deferral): back into its rings, it is counted as pending already
        if ((atomicAnd_system(&r.state[tile], ~ST_DEFER) & ST_DEFER) && r.bq) {
            const int b = max(bucket_of(d, kbits), 0);
            const unsigned long long t = atomicAdd_system(&r.ctl->btail[b], 1ULL);
            atomicExch_system(&r.bq[(size_t) b * (r.bqmask + 1u) + ((unsigned) t & r.bqmask)], tile + 1);
            fence_sys();
            atomicAdd_system(&r.ctl->bcount[b], 1);
        }
        atomicAdd(&d.ft->fc[d.ft->rank]->tx, 1ULL);
        return;
    }
    bool push = false;
    if (st == 0) {
        if (near) push = true;
        else {
            atomicOr_system(&r.state[tile], ST_FAR);
            const int slot = atomicAdd_system(&rfc->inbox_count, 1);
            if (slot < r.inbox_cap) { r.inbox[slot] = tile; fence_sys(); }
            else { atomicAnd_system(&r.state[tile], ~ST_FAR); push = true; }  // inbox full: run it now
        }
    } else if ((st & ST_FAR) && near) {
        if (atomicAnd_system(&r.state[tile], ~ST_FAR) & ST_FAR) push = true;
    }
    if (push) {
        fused_rank_wakes(d);  // BEFORE the count of the other rank can leave 0 ...
        if (atomicAdd_system(&r.ctl->pending, 1) != 0) fused_rank_idles(d);  // ... it was awake already
        if (d.nb > 0 && r.bq) {  // bucketed near queue (every rank runs the same configuration): the ring of the key
            const int b = bucket_of(d, kbits);
            const unsigned long long t = atomicAdd_system(&r.ctl->btail[b], 1ULL);
            atomicExch_system(&r.bq[(size_t) b * (r.bqmask + 1u) + ((unsigned) t & r.bqmask)], tile + 1);
            fence_sys();
            atomicAdd_system(&r.ctl->bcount[b], 1);
        } else {
            const unsigned long long t = atomicAdd_system(&r.ctl->tail, 1ULL);
            atomicExch_system(&r.queue[(unsigned) t & r.qmask], tile + 1);
        }
        atomicAdd_system(&rfc->rx, 1ULL);
    }
    atomicAdd(&d.ft->fc[d.ft->rank]->tx, 1ULL);
}

// lowest key of this rank's pile (all threads of the CTA; the inbox is moved into the pile first: the grid is quiescent)
__device__ inline int fused_pile_min(const GridDesc &d, int *s_red, int tid) {
    DevCtl *c = d.ctl;
    FusedCtl *fc = d.ft->fc[d.ft->rank];
    const int nthreads = blockDim.x;
    const int cur = ld_sys(&c->far_sel);
    const int nin = min(ld_sys(&fc->inbox_count), d.inbox_cap);
    for (int i = tid; i < nin; i += nthreads) {
        const int t = __ldcg(d.inbox + i);
        if (ld_sys(&d.state[t]) & ST_FAR) {
            const int slot = atomicAdd(&c->far_count[cur], 1);
            if (slot < d.far_cap) d.far[cur * d.far_cap + slot] = t;
            else if (atomicAnd(&d.state[t], ~ST_FAR) & ST_FAR) {  // pile full: run it with the next release
                if (atomicAdd(&c->pending, 1) == 0) fused_rank_wakes(d);
                const unsigned long long tk = atomicAdd(&c->tail, 1ULL);
                atomicExch(&d.queue[(unsigned) tk & d.qmask], t + 1);
            }
        }
    }
    __threadfence();
    __syncthreads();
    if (tid == 0) fc->inbox_count = 0;
    const int n = min(ld_sys(&c->far_count[cur]), d.far_cap);
    const int *src = d.far + cur * d.far_cap;
    int m = INF_BITS;
    const bool keyed = d.fmode == 0;  // the invalidate mode has no order (and no keys): any parked tile counts as 0
    for (int i = tid; i < n; i += nthreads) {
        const int t = __ldcg(src + i);
        if (ld_sys(&d.state[t]) & ST_FAR) m = min(m, keyed ? ld_sys(&d.key[t]) : 0);
    }
    m = __reduce_min_sync(0xffffffffu, m);
    if ((tid & 31) == 0) s_red[tid >> 5] = m;
    __syncthreads();
    m = s_red[0];
    for (int w = 1; w < nthreads / 32; ++w) m = min(m, s_red[w]);
    __syncthreads();
    return m;
}

// new threshold: the parked tiles below it go to the queue (all threads of the CTA)
__device__ inline void fused_pile_release(const GridDesc &d, int nthr, int tid) {
    DevCtl *c = d.ctl;
    const int nthreads = blockDim.x;
    const int cur = ld_sys(&c->far_sel);
    const int n = min(ld_sys(&c->far_count[cur]), d.far_cap);
    const int *src = d.far + cur * d.far_cap;
    if (tid == 0) {
        fused_rank_wakes(d);
        atomicAdd(&c->pending, 1);  // hold the count while tiles are being pushed
        c->thr_bits = nthr;
        c->far_count[cur ^ 1] = 0;
        c->far_sel = cur ^ 1;  // parkings by the tiles pushed below go to the other list
        atomicAdd(&c->phases, 1);
        __threadfence();
    }
    __syncthreads();
    int *dst = d.far + (cur ^ 1) * d.far_cap;
    for (int i = tid; i < n; i += nthreads) {
        const int t = __ldcg(src + i);
        if (!(ld_sys(&d.state[t]) & ST_FAR)) continue;  // stale entry (tile was promoted)
        if (ld_sys(&d.key[t]) < nthr || nthr == INF_BITS) {
            if (atomicAnd(&d.state[t], ~ST_FAR) & ST_FAR) {
                q_push(d, t);
                atomicAdd(&c->pushes, 1ULL);
            }
        } else {
            const int slot = atomicAdd(&c->far_count[cur ^ 1], 1);
            if (slot < d.far_cap) dst[slot] = t;
            else if (atomicAnd(&d.state[t], ~ST_FAR) & ST_FAR) q_push(d, t);
        }
    }
    __syncthreads();
    if (tid == 0) {
        c->far_count[cur] = 0;
        if (atomicSub(&c->pending, 1) == 1) fused_rank_idles(d);  // nothing released here
    }
}

// wait until word `p` (written by another rank) equals `want`; false on watchdog
__device__ inline bool fused_wait_eq(const GridDesc &d, const int *p, int want) {
    const unsigned long long t0 = globaltimer_ns();
    while (ld_sys(p) != want) {
        if (ld_sys(&d.ctl->error) || globaltimer_ns() - t0 > d.ctl->timeout_ns) { q_fail(d); return false; }
        __nanosleep(64);
    }
    return true;
}

// CTA 0 of every rank, whenever q_pop_fused returns -2: serve the coordinator's command (ranks > 0) or run one phase
// boundary (rank 0). s_cmd: two shared ints.
__device__ inline void fused_service(const GridDesc &d, int *s_red, int *s_cmd, int tid) {
    const FusedTopo &ft = *d.ft;
    FusedCtl *me = ft.fc[ft.rank], *co = ft.fc[0];
    if (ft.rank == 0) {
        int seq = 0;
        if (tid == 0) {
            seq = ++me->seq;
            for (int r = 1; r < ft.world; ++r) { fence_sys(); *(volatile int *) &ft.fc[r]->cmd = (seq << 2) | FOP_MIN; }
        }
        int m = fused_pile_min(d, s_red, tid);
        if (tid == 0) {
            bool ok = true;
            for (int r = 1; r < ft.world && ok; ++r) {
                ok = fused_wait_eq(d, &me->ack[r], seq);
                m = min(m, ld_sys(&me->rmin[r]));
            }
            int op = FOP_DONE, nthr = INF_BITS;
            if (ok && m != INF_BITS) {
                op = FOP_RELEASE;
                if (d.fmode == 0 && d.delta > 0.0f) {
                    nthr = __float_as_int(__int_as_float(m) + d.delta);
                    if (nthr <= m) nthr = m + 1;  // delta lost to rounding: still let the lowest tile through
                }
            }
            seq = ++me->seq;
            for (int r = 1; r < ft.world; ++r) {
                *(volatile int *) &ft.fc[r]->cmd_thr_bits = nthr;
                fence_sys();
                *(volatile int *) &ft.fc[r]->cmd = (seq << 2) | op;
            }
            s_cmd[0] = op; s_cmd[1] = nthr;
            d.ctl->far_min_bits = m;
        }
        __syncthreads();
        const int op = s_cmd[0], nthr = s_cmd[1];
        if (op == FOP_RELEASE) {
            fused_pile_release(d, nthr, tid);
            if (tid == 0)
                for (int r = 1; r < ft.world; ++r)
                    if (!fused_wait_eq(d, &me->ack[r], seq)) break;
        } else if (tid == 0) {
            __threadfence();
            atomicExch(&d.ctl->done, 1);
        }
        __syncthreads();
        return;
    }
    if (tid == 0) {
        const int cmd = ld_sys(&me->cmd);
        fence_sys();
        s_cmd[0] = cmd; s_cmd[1] = ld_sys(&me->cmd_thr_bits);
    }
    __syncthreads();
    const int cmd = s_cmd[0], op = cmd & 3, seq = cmd >> 2;
    if (op == FOP_MIN) {
        const int m = fused_pile_min(d, s_red, tid);
        if (tid == 0) { *(volatile int *) &co->rmin[ft.rank] = m; d.ctl->far_min_bits = m; }
    } else if (op == FOP_RELEASE) {
        fused_pile_release(d, s_cmd[1], tid);
    } else if (tid == 0) {
        __threadfence();
        atomicExch(&d.ctl->done, 1);
    }
    __syncthreads();
    if (tid == 0) {
        me->seen_seq = seq;
        fence_sys();
        *(volatile int *) &co->ack[ft.rank] = seq;
    }
    __syncthreads();
}

// q_pop for CTA 0 in fused mode: also watches for the coordinator's command (rank 0: for the grid going quiescent).
// *ticket == ~0: take a new ticket. Returns the tile, -1 (launch over) or -2 (call fused_service, then come back
// with the same ticket).
__device__ inline int q_pop_fused(const GridDesc &d, unsigned long long *ticket) {
    const FusedTopo &ft = *d.ft;
    FusedCtl *me = ft.fc[ft.rank];
    if (*ticket == ~0ULL) *ticket = atomicAdd(&d.ctl->head, 1ULL);
    volatile int *slot = &d.queue[(unsigned) *ticket & d.qmask];
    volatile int *done = &d.ctl->done;
    unsigned long long t0 = 0, tail0 = 0;
    unsigned ns = 32;
    for (;;) {
        int v = *slot;
        if (v != 0) {
            *slot = 0;
            *ticket = ~0ULL;
            return v - 1;
        }
        if (*done) return -1;
        if (ft.rank == 0) {
            if (ld_sys(&me->active) == 0) return -2;
        } else if ((ld_sys(&me->cmd) >> 2) != me->seen_seq) return -2;
        if (t0 == 0) {
            t0 = globaltimer_ns();
            tail0 = *(volatile unsigned long long *) &d.ctl->tail;
        } else if (globaltimer_ns() - t0 > d.ctl->timeout_ns) {
            const unsigned long long tl = *(volatile unsigned long long *) &d.ctl->tail;
            if (tl == tail0) { q_fail(d); return -1; }
            tail0 = tl;
            t0 = globaltimer_ns();
        }
        __nanosleep(ns);
        if (ns < Q_SLEEP_MAX_NS) ns <<= 1;
    }
}

// q_pop of the bucketed near queue, called by the 32 lanes of warp 0 together: lane b looks at the counts of rings b and
// b + 31 (coalesced reads), the lowest ring with entries is tried: a unit of its count is taken with an unconditional
// atomic (given back if the count was already spent — no CAS loop, so hundreds of CTAs do not serialise on one word),
// then a ticket of its head, whose slot is filled or about to be. Nobody commits to a ring that may never be fed again.
// Lane 31 watches the plain ring (entries queued without an order before the launch). Returns the same value to every lane.
// fused0: CTA 0 of a fused multi-GPU launch, which also serves the coordinator (returns -2: call fused_service).
__device__ inline int q_pop_warp(const GridDesc &d, int lane, bool fused0, int *ring_out = nullptr) {
    DevCtl *c = d.ctl;
    const unsigned full = 0xffffffffu;
    const bool m0 = lane < 31 && lane < d.nb, m1 = !d.slide && lane < 31 && lane + 31 < d.nb;  // rings lane and lane + 31
    unsigned long long t0 = 0, sum0 = 0;
    unsigned ns = 32, polls = 0;
    for (;;) {
        int cnt0 = 0, cnt1 = 0, run0 = 0;
        unsigned long long ph = 0, pt = 0;
        // sliding window (nb <= 31): lane l looks at the ring l slices above the base, so the lowest lane is the lowest key
        int abase = 0, rot = lane;
        if (d.slide) {
            abase = __shfl_sync(full, *(volatile int *) &c->abase, 0);
            rot = (abase + lane) % d.nb;
            if (m0) run0 = *(volatile int *) &c->brun[rot];
        }
        if (m0) cnt0 = *(volatile int *) &c->bcount[rot];
        if (m1) cnt1 = *(volatile int *) &c->bcount[lane + 31];
        if (lane == 31) {
            ph = *(volatile unsigned long long *) &c->head;
            pt = *(volatile unsigned long long *) &c->tail;
        }
        const unsigned ma = __ballot_sync(full, cnt0 > 0), mb = __ballot_sync(full, cnt1 > 0);
        const unsigned mp = __ballot_sync(full, lane == 31 && ph < pt);
        if (d.slide) {
            // the window base follows the lowest ring that holds a queued tile OR a tile being visited (a visit pushes at
            // or above its own slice, so nothing can appear below such a ring); crossing a multiple of
            // SLIDE_SCAN_EVERY slices, this warp also looks through the far list
            const unsigned mr = __ballot_sync(full, cnt0 > 0 || run0 > 0);
            const int lowest = mr ? __ffs(mr) - 1 : 0;
            if (lowest > 0) {
                int scan = 0;
                if (lane == 0) {
                    const int old = atomicMax(&c->abase, abase + lowest);
                    scan = old < abase + lowest && (abase + lowest) / SLIDE_SCAN_EVERY > old / SLIDE_SCAN_EVERY &&
                           *(volatile int *) &c->far_count[0] > *(volatile int *) &c->far_scanned;
                }
                if (__shfl_sync(full, scan, 0)) slide_scan_warp(d, lane, false);
            }
        }
        if (ma | mb | mp) {
            const int src = ma ? __ffs(ma) - 1 : (mb ? __ffs(mb) - 1 : 31);
            const int b = ma ? (d.slide ? (abase + src) % d.nb : src) : src + 31;
            int v = 0;
            if (lane == src) {
                volatile int *slot = nullptr;
                if (!(ma | mb)) {
                    if (atomicCAS(&c->head, ph, ph + 1ULL) == ph) slot = &d.queue[(unsigned) ph & d.qmask];
                } else {
                    if (d.slide) atomicAdd(&c->brun[b], 1);  // before the count drops: the ring never looks idle in between
                    if (atomicSub(&c->bcount[b], 1) > 0) {
                        const unsigned long long h = atomicAdd(&c->bhead[b], 1ULL);
                        slot = &d.bq[(size_t) b * (d.bqmask + 1u) + ((unsigned) h & d.bqmask)];
                    } else {
                        atomicAdd(&c->bcount[b], 1);  // spent meanwhile: give the unit back
                        if (d.slide) atomicSub(&c->brun[b], 1);
                    }
                }
                if (slot) {
                    while ((v = *slot) == 0) {}
                    *slot = 0;
                }
            }
            v = __shfl_sync(full, v, src);
            if (v) {
                if (ring_out && lane == 0) *ring_out = (ma | mb) ? b : -1;
                return v - 1;
            }
            continue;  // somebody else took it: look again
        }
        int stop = 0;
        if (lane == 0) {
            if (*(volatile int *) &c->done) stop = 1;
            else if (fused0 && (d.ft->rank == 0 ? ld_sys(&d.ft->fc[0]->active) == 0
                                                : (ld_sys(&d.ft->fc[d.ft->rank]->cmd) >> 2) != d.ft->fc[d.ft->rank]->seen_seq))
                stop = 2;
            else if ((++polls & 63u) == 0) {
                // watchdog on GLOBAL progress: nothing pushed anywhere for timeout_ns
                unsigned long long sum = *(volatile unsigned long long *) &c->tail;
                for (int i = 0; i < d.nb; ++i) sum += *(volatile unsigned long long *) &c->btail[i];
                if (t0 == 0 || sum != sum0) { t0 = globaltimer_ns(); sum0 = sum; }
                else if (globaltimer_ns() - t0 > c->timeout_ns) { q_fail(d); stop = 1; }
            }
        }
        stop = __shfl_sync(full, stop, 0);
        if (stop) return -stop;
        __nanosleep(ns);
        if (ns < Q_SLEEP_MAX_NS) ns <<= 1;
    }
}

// ---- upwind deferral (TASI_DEFER; bucketed queue, unsharded grids) --------------------------------------------------------
// A tile visited while an upwind neighbour is still queued or running works on a border that is about to change, and is
// visited again when it has: that is most of the 8-17 visits a tile gets. Here a popped tile looks at its 8 neighbours
// first; if one of them is queued or running with a key lower than its own by more than `defer_margin` (about half a tile
// width of cost, so that tiles side by side on the same front do not wait for each other), it is set aside: ST_DEFER in
// its state word, in no ring, still counted as pending. It returns to the rings when (a) a neighbour leaves its visit (the
// leaving CTA releases every deferred neighbour) or (b) an activation lowers its key (what it waited for may now have to
// wait for it). "Lower by a margin" is acyclic and the tile with the lowest key never waits, so the schedule cannot stall;
// both hand-overs go through one atomic on the state word of the deferred tile (whoever clears the bit owns the push).
__device__ __forceinline__ bool defer_blocked(const GridDesc &d, int t, int lane) {
    const int mykey = *(volatile int *) &d.key[t];
    bool blocked = false;
    if (lane < 8 && mykey != INF_BITS) {
        const float lim = __int_as_float(mykey) - d.defer_margin;
        const int tx = t / d.TY, ty = t - tx * d.TY;
        const int dx = (0x31 >> lane & 1) ? -1 : ((0xc2 >> lane & 1) ? 1 : 0);
        const int dy = (0x54 >> lane & 1) ? -1 : ((0xa8 >> lane & 1) ? 1 : 0);
        const int nx = tx + dx, ny = ty + dy;
        if (lim > 0.0f && nx >= 0 && nx < d.TX && ny >= 0 && ny < d.TY && !dead_tile_row(d, nx)) {
            const int n = nx * d.TY + ny;
            const int st = *(volatile int *) &d.state[n];
            if ((st & (ST_DIRTY | ST_RUN)) && !(st & ST_FAR)) {
                int nk = INF_BITS;
                if (st & ST_RUN) nk = *(volatile int *) &d.rkey[n];
                if (st & ST_DIRTY) nk = min(nk, *(volatile int *) &d.key[n]);
                blocked = nk < __float_as_int(lim);
            }
        }
    }
    return __any_sync(0xffffffffu, blocked);
}

// The popped tile t: true = run it; false = it was set aside (or a releaser has taken it over): pop another. Warp 0.
__device__ inline bool defer_or_run(const GridDesc &d, int t, int lane) {
    if (!defer_blocked(d, t, lane)) return true;
    if (lane == 0) {
        atomicOr(&d.state[t], ST_DEFER);
        fence_gpu();
    }
    __syncwarp();
    int run = 0;
    if (!defer_blocked(d, t, lane)) {
        // whatever blocked it has left meanwhile (and may or may not have seen the bit): whoever clears it owns the tile
        if (lane == 0) run = (atomicAnd(&d.state[t], ~ST_DEFER) & ST_DEFER) ? 1 : 0;
        run = __shfl_sync(0xffffffffu, run, 0);
    } else if (lane == 0) {
        atomicAdd(&d.ctl->deferred, 1ULL);
    }
    return run != 0;
}

// A visit has ended (the tile is idle or parked again): release the deferred neighbours. Lanes 0..7 of one warp.
__device__ __forceinline__ void release_deferred(const GridDesc &d, int tile, int lane) {
    if (lane >= 8) return;
    const int tx = tile / d.TY, ty = tile - tx * d.TY;
    const int dx = (0x31 >> lane & 1) ? -1 : ((0xc2 >> lane & 1) ? 1 : 0);
    const int dy = (0x54 >> lane & 1) ? -1 : ((0xa8 >> lane & 1) ? 1 : 0);
    const int nx = tx + dx, ny = ty + dy;
    if (nx < 0 || nx >= d.TX || ny < 0 || ny >= d.TY || dead_tile_row(d, nx)) return;
    const int n = nx * d.TY + ny;
    if ((*(volatile int *) &d.state[n] & ST_DEFER) && (atomicAnd(&d.state[n], ~ST_DEFER) & ST_DEFER)) q_repush(d, n);
}

// ---- in-tile relaxation: strips with a register copy of their neighbourhood ------------------------------------------
// Thread t owns the 1x4 strip (row t>>3, columns 4*(t&7)..+3) of the staged tile and keeps, in registers, the 3x6
// window of values around it as it last saw them. A round = reload the window from shared memory; if any of the
// 18 values differs from the copy (some neighbour, or the strip itself, changed since this thread last looked)
// re-evaluate the 4 elements and store the ones that improved; __syncthreads_or of "somebody stored" ends the
// visit. There are no dirty bitmaps, lists or shared atomics on this path: "dirty" is the comparison, a round
// costs one window load (conflict-free: rows of the pitch-35 array are 3 banks apart, strips 4) plus, where
// something changed, four independent evaluations that overlap in the pipeline. Reads may race with the stores
// of the same round; whichever value a thread saw is what its copy records, so a missed store shows up as a
// difference in the next round. The visit ends after a round without stores, when every thread's copy equals the
// tile and has been evaluated: tile-local fixed point.
//
// Direction bits: which halo side of a tile changed (set by the neighbour that wrote it); they select the
// strips that evaluate in the first round (the others only take their copy).
constexpr int IN_N = 1, IN_S = 2, IN_W = 4, IN_E = 8, IN_NW = 16, IN_NE = 32, IN_SW = 64, IN_SE = 128, IN_ALL = 256;
constexpr int WIN_R = 3;          // a strip of SW elements reads a 3 x (SW + 2) window
constexpr int MAX_WARPS = 16;     // per CTA, over the kernel instances (SW = 2: 512 threads)
// Relaxation sweeps between two CTA barriers. Stores of a sweep are visible to every warp at once (the window loads are
// volatile), the barrier only decides termination: a super-round without a store anywhere means every thread has
// compared its window with the final tile and evaluated it. Two sweeps per barrier halve the barrier waits where the
// sweeps are Jacobi (MS-DFM 1024^2 3.35 -> 3.08 ms, 16384^2 153 -> 149 ms; FD* 4096^2 29.3 -> 28.3 ms); the
// Gauss-Seidel strips of the wide node instance do not gain (FD* 16384^2 251 -> 254 ms) and keep one.
#ifndef RPB_DFM
#define RPB_DFM 2
#endif
#ifndef UPWIND_FILTER
#define UPWIND_FILTER 1
#endif
template <int KIND, int SW>
struct SweepsPerBarrier { static constexpr int value = (KIND == TASI_KIND_DFM) ? RPB_DFM : (SW == 2 ? 2 : 1); };
constexpr int LEAVE_TID = 32;     // the thread that releases the tile after a visit (first lane of warp 1)

template <int KIND, int SW>
__device__ __forceinline__ float eval_strip_elem(const float (&w)[WIN_R][SW + 2], const float (&cw)[2][SW + 1], int j) {
    Nbr8 nb;
    nb.n = w[0][j + 1]; nb.s = w[2][j + 1]; nb.w = w[1][j]; nb.e = w[1][j + 2];
    nb.nw = w[0][j]; nb.ne = w[0][j + 2]; nb.sw = w[2][j]; nb.se = w[2][j + 2];
    if (KIND == TASI_KIND_DFM) return dfm_cell(nb, cw[0][j]);
    // cw[a][b] = cost of cell (row-1+a, col0-1+b): the four cells around node (row, col0+j)
    const float tl = cw[0][j], tr = cw[0][j + 1], bl = cw[1][j], br = cw[1][j + 1];
    return (KIND == TASI_KIND_FD) ? fd_node(nb, tl, tr, bl, br) : mfd_node(nb, tl, tr, bl, br);
}

// Relax (MODE 0) or invalidate (MODE 1) the staged tile to local convergence. w[1][1..4] holds the strip's final
// values on return; chg gets one bit per strip element that changed during the visit.
template <int KIND, int MODE, int SW>
__device__ __forceinline__ int relax_tile(float *sg, const float *sc, int inmask, int row, int col0, int rlim, int clim,
                                          int pin_r, int pin_c, float tol, float (&w)[WIN_R][SW + 2],
                                          unsigned &chg, unsigned &iters, unsigned &evals) {
    const float INF = __int_as_float(0x7f800000);
    constexpr int WIN_C = SW + 2;
    // which strip elements exist and may be written (the goal element is pinned, FieldDPlanner_impl.h:47,146)
    unsigned live = 0;
#pragma unroll
    for (int j = 0; j < SW; ++j)
        if (row < rlim && col0 + j < clim && !(row == pin_r && col0 + j == pin_c)) live |= 1u << j;
    // costs the strip reads, once per visit
    float cw[2][SW + 1];
    if (KIND == TASI_KIND_DFM) {
#pragma unroll
        for (int j = 0; j < SW; ++j) cw[0][j] = sc[row * CP + col0 + j];
    } else {
#pragma unroll
        for (int a = 0; a < 2; ++a)
#pragma unroll
            for (int b = 0; b < SW + 1; ++b) cw[a][b] = sc[(row + a) * CP + col0 + b];
    }
    // first round: strips that touch a changed halo side
    bool dirty = (inmask & IN_ALL) != 0;
    dirty |= (inmask & (IN_N | IN_NW | IN_NE)) && row == 0;
    dirty |= (inmask & (IN_S | IN_SW | IN_SE)) && row == TS - 1;
    dirty |= (inmask & (IN_W | IN_NW | IN_SW)) && col0 == 0;
    dirty |= (inmask & (IN_E | IN_NE | IN_SE)) && col0 == TS - SW;
    const float *base = sg + row * GP + col0;  // window origin: element (row-1, col0-1)
    int cnt_changed = 0;
    chg = 0;
    bool firstround = true, self_changed = false;
    for (;;) {
      bool stored_any = false;
#pragma unroll 1
      for (int rep = 0; rep < SweepsPerBarrier<KIND, SW>::value; ++rep) {
        bool diff = false;
#pragma unroll
        for (int a = 0; a < WIN_R; ++a)
#pragma unroll
            for (int b = 0; b < WIN_C; ++b) {
                const float x = *(const volatile float *) (base + a * GP + b);
                diff |= __float_as_int(x) != __float_as_int(w[a][b]);
                w[a][b] = x;
            }
        if (!firstround) dirty = diff || self_changed;
        firstround = false;
        self_changed = false;
        bool stored = false;
        if (dirty && live) {
            if (MODE == 0 && SW == 4 && KIND != TASI_KIND_DFM) {
                // Gauss-Seidel along the strip, direction alternating from round to round: an element reads the value its
                // strip-mate just took, so a change crosses the whole strip in one round instead of one element
                // (node planners with 4-element strips: 15.4 -> 13.7 rounds per visit, MFD* 16384^2 154 -> 134 ms;
                // measured slower for 2-element strips and for MS-DFM, which keep the Jacobi order below)
                if (iters & 1u) {
#pragma unroll
                    for (int j = SW - 1; j >= 0; --j) {
                        const float v = eval_strip_elem<KIND, SW>(w, cw, j);
                        if (v < w[1][j + 1] && ((live >> j) & 1u)) {
                            w[1][j + 1] = v;
                            sg[(row + 1) * GP + col0 + 1 + j] = v;
                            chg |= 1u << j;
                            ++cnt_changed;
                            stored = true;
                        }
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < SW; ++j) {
                        const float v = eval_strip_elem<KIND, SW>(w, cw, j);
                        if (v < w[1][j + 1] && ((live >> j) & 1u)) {
                            w[1][j + 1] = v;
                            sg[(row + 1) * GP + col0 + 1 + j] = v;
                            chg |= 1u << j;
                            ++cnt_changed;
                            stored = true;
                        }
                    }
                }
                self_changed = stored;
            } else {
                float v[SW];
#pragma unroll
                for (int j = 0; j < SW; ++j) v[j] = eval_strip_elem<KIND, SW>(w, cw, j);
#pragma unroll
                for (int j = 0; j < SW; ++j) {
                    const float cur = w[1][j + 1];
                    bool c;
                    if (MODE == 0) c = v[j] < cur;
                    else c = cur < INF && cur > 0.0f && v[j] > cur + cur * tol;  // 0 = a goal (costs are >= 1): pinned
                    c = c && ((live >> j) & 1u);
                    if (c) {
                        sg[(row + 1) * GP + col0 + 1 + j] = (MODE == 0) ? v[j] : INF;
                        chg |= 1u << j;
                        ++cnt_changed;
                        stored = true;
                    }
                }
            }
            evals += SW;
        }
        ++iters;
        stored_any |= stored;
      }
        if (!__syncthreads_or(stored_any)) break;
    }
    return cnt_changed;
}

// ---- in-tile relaxation, compacted (CMP = 1) --------------------------------------------------------------------------
// The strip relaxation above spends most of its issue slots on lanes that have nothing to do: a front crosses a tile
// diagonally, so in a typical round ~18 % of the elements are dirty but every warp holds a few of them and runs the
// whole SW-element evaluation (profiles/r01_ncu_solver_dfm4096_v12.txt: 19.8 of 32 lanes active). Here a
// round is split: (1) every thread compares the window of its strip with its register copy, as before, and derives
// WHICH of its elements have a changed neighbourhood; (2) the dirty elements of the whole tile are compacted into a
// dense list in shared memory (warp prefix sum + one shared atomic per warp); (3) thread i evaluates list element i —
// one element per thread, dense warps, Jacobi order. The owner of a strip learns what changed from the next round's
// window comparison (which is also what records the elements to write back), so there are no per-element flags.
// Two CTA barriers per round instead of one per 1-2 sweeps; the bet is on the issue slots.
template <int KIND, int SW>
__device__ __forceinline__ int relax_tile_compact(float *sg, const float *sc, unsigned short *s_list, int *s_n, int inmask,
                                                  int row, int col0, int rlim, int clim, int pin_r, int pin_c,
                                                  float (&w)[WIN_R][SW + 2], unsigned &chg, unsigned &iters,
                                                  unsigned &evals, int *dbg) {
    constexpr int WIN_C = SW + 2, NT = (TS * TS) / SW;
    const int tid = threadIdx.x, lane = tid & 31;
    unsigned live = 0;
#pragma unroll
    for (int j = 0; j < SW; ++j)
        if (row < rlim && col0 + j < clim && !(row == pin_r && col0 + j == pin_c)) live |= 1u << j;
    bool dirty0 = (inmask & IN_ALL) != 0;
    dirty0 |= (inmask & (IN_N | IN_NW | IN_NE)) && row == 0;
    dirty0 |= (inmask & (IN_S | IN_SW | IN_SE)) && row == TS - 1;
    dirty0 |= (inmask & (IN_W | IN_NW | IN_SW)) && col0 == 0;
    dirty0 |= (inmask & (IN_E | IN_NE | IN_SE)) && col0 == TS - SW;
    const float *base = sg + row * GP + col0;  // window origin: element (row-1, col0-1)
    int cnt_changed = 0;
    chg = 0;
    bool first = true;
    for (;;) {
        const int par = iters & 1;
        // (1) what changed around my strip since I last looked
        unsigned coldiff = 0;
        float cmn[WIN_C];  // per window column: the lowest value that changed since this thread last looked
#pragma unroll
        for (int b = 0; b < WIN_C; ++b) cmn[b] = __int_as_float(INF_BITS);
#pragma unroll
        for (int a = 0; a < WIN_R; ++a)
#pragma unroll
            for (int b = 0; b < WIN_C; ++b) {
                const float x = *(const volatile float *) (base + a * GP + b);
                if (__float_as_int(x) != __float_as_int(w[a][b])) {
                    coldiff |= 1u << b;
                    cmn[b] = fminf(cmn[b], x);
                    if (a == 1 && b >= 1 && b <= SW && !first) chg |= 1u << (b - 1);  // one of mine was lowered
                }
                w[a][b] = x;
            }
        unsigned dmask = 0;
        if (first) dmask = dirty0 ? (1u << SW) - 1u : 0u;
        else {
#pragma unroll
            for (int j = 0; j < SW; ++j) {
                // dirty = a neighbour changed to a value BELOW the element's own: a relaxation can only take an element
                // below its value through a neighbour that lies below it, so the changes on the downwind side of a front
                // (lowered again and again while it settles) do not make the elements behind it dirty. Bit-exact (245 GPU
                // tests); evaluations per element 12.6 -> 9.3 (FD* / MFD* 16384^2, -5 % time). The strips do not use
                // it: tracking the lowest changed value per window column costs every thread more than it saves there
                // (MS-DFM 4096^2 14.6 -> 16.7 ms)
                const bool d = UPWIND_FILTER ? fminf(fminf(cmn[j], cmn[j + 1]), cmn[j + 2]) < w[1][j + 1]
                                             : ((coldiff >> j) & 7u) != 0;
                if (d) dmask |= 1u << j;
            }
        }
        dmask &= live;
        first = false;
        // (2) compaction
        const int my = __popc(dmask);
        int incl = my;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        const int wtot = __shfl_sync(0xffffffffu, incl, 31);
        int wbase = 0;
        if (lane == 31 && wtot) wbase = atomicAdd(&s_n[par], wtot);
        wbase = __shfl_sync(0xffffffffu, wbase, 31);
        int pos = wbase + incl - my;
#pragma unroll
        for (int j = 0; j < SW; ++j)
            if ((dmask >> j) & 1u) s_list[pos++] = (unsigned short) (row * TS + col0 + j);
        __syncthreads();
        int n = *(volatile int *) &s_n[par];
        if (n == 0) {
            if (tid == 0) s_n[par ^ 1] = 0;  // last round's count: both counters are 0 when the next visit starts
            break;
        }
#ifdef TASI_DEBUG_COMPACT
        if (n > TS * TS || n < 0) { if (tid == 0) atomicExch(dbg, 100000 + n); n = TS * TS; }
#endif
        // (3) one element per thread
        for (int i = tid; i < n; i += NT) {
            int e = s_list[i];
#ifdef TASI_DEBUG_COMPACT
            if (e >= TS * TS) { atomicExch(dbg, 200000 + e); e &= TS * TS - 1; }
#endif
            const int r = e >> 5, c = e & 31;
            const volatile float *wb = sg + r * GP + c;
            Nbr8 nb;
            nb.nw = wb[0]; nb.n = wb[1]; nb.ne = wb[2];
            nb.w = wb[GP]; const float cur = wb[GP + 1]; nb.e = wb[GP + 2];
            nb.sw = wb[2 * GP]; nb.s = wb[2 * GP + 1]; nb.se = wb[2 * GP + 2];
            float v;
            if (KIND == TASI_KIND_DFM) v = dfm_cell(nb, sc[r * CP + c]);
            else {
                const float tl = sc[r * CP + c], tr = sc[r * CP + c + 1], bl = sc[(r + 1) * CP + c], br = sc[(r + 1) * CP + c + 1];
                v = (KIND == TASI_KIND_FD) ? fd_node(nb, tl, tr, bl, br) : mfd_node(nb, tl, tr, bl, br);
            }
            if (v < cur) {
                sg[(r + 1) * GP + c + 1] = v;
                ++cnt_changed;
            }
            ++evals;
        }
        if (tid == 0) s_n[par ^ 1] = 0;  // the next round's counter; nobody adds to it before the barrier below
        __syncthreads();
        ++iters;
    }
    return cnt_changed;
}

// ---- in-tile relaxation, compacted, dirty marks pushed by the evaluations (CMP = 3) ------------------------------------
// relax_tile_compact finds the dirty elements of a round by having EVERY thread re-read and compare the 3 x (SW + 2)
// window of its strip: about as many instructions per round as the evaluations themselves once the lists are short
// (FD* 16384^2: ~110 evaluations per round). Here the evaluation that lowers an element marks the neighbours the new
// value can help — those whose value it has just read as higher (it holds all eight in registers) — in a byte map in
// shared memory; the owner of a strip reads its SW bytes (one load), clears them and puts the marked elements on the
// round's list. A second byte map records which elements changed during the visit (what the owner writes back). Marks are
// plain byte stores of 1 (races are benign); a value read stale is read as higher, so a mark can only be superfluous,
// never missing: the fixed point is the one of the other relaxations (bit-exact). Two CTA barriers per round, as CMP = 1.
template <int KIND, int SW>
__device__ __forceinline__ int relax_tile_push(float *sg, const float *sc, unsigned short *s_list, int *s_n,
                                               unsigned char *s_dirty, unsigned char *s_chgmap, int inmask, int row, int col0,
                                               int rlim, int clim, int pin_r, int pin_c, float (&w)[WIN_R][SW + 2],
                                               unsigned &chg, unsigned &iters, unsigned &evals) {
    constexpr int NT = (TS * TS) / SW;
    const int tid = threadIdx.x, lane = tid & 31;
    unsigned live = 0;
#pragma unroll
    for (int j = 0; j < SW; ++j)
        if (row < rlim && col0 + j < clim && !(row == pin_r && col0 + j == pin_c)) live |= 1u << j;
    bool dirty0 = (inmask & IN_ALL) != 0;
    dirty0 |= (inmask & (IN_N | IN_NW | IN_NE)) && row == 0;
    dirty0 |= (inmask & (IN_S | IN_SW | IN_SE)) && row == TS - 1;
    dirty0 |= (inmask & (IN_W | IN_NW | IN_SW)) && col0 == 0;
    dirty0 |= (inmask & (IN_E | IN_NE | IN_SE)) && col0 == TS - SW;
    // this thread's SW bytes of the two maps (aligned: col0 is a multiple of SW)
    unsigned char *myd = s_dirty + row * TS + col0, *myc = s_chgmap + row * TS + col0;
    int cnt_changed = 0;
    bool first = true;
    for (;;) {
        const int par = iters & 1;
        // (1) my marked elements; the maps start clean with the first round
        unsigned dmask = 0;
        if (first) {
            dmask = dirty0 ? (1u << SW) - 1u : 0u;
            if (SW == 4) *(unsigned *) myc = 0u; else *(unsigned short *) myc = 0;
        } else {
            const unsigned bytes = (SW == 4) ? *(volatile unsigned *) myd : (unsigned) *(volatile unsigned short *) myd;
#pragma unroll
            for (int j = 0; j < SW; ++j)
                if ((bytes >> (8 * j)) & 0xffu) dmask |= 1u << j;
        }
        if (SW == 4) *(unsigned *) myd = 0u; else *(unsigned short *) myd = 0;
        dmask &= live;
        first = false;
        // (2) compaction (as relax_tile_compact); a warp without a marked element has nothing to add
        if (__any_sync(0xffffffffu, dmask != 0u)) {
            const int my = __popc(dmask);
            int incl = my;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += t;
            }
            const int wtot = __shfl_sync(0xffffffffu, incl, 31);
            int wbase = 0;
            if (lane == 31) wbase = atomicAdd(&s_n[par], wtot);
            wbase = __shfl_sync(0xffffffffu, wbase, 31);
            int pos = wbase + incl - my;
#pragma unroll
            for (int j = 0; j < SW; ++j)
                if ((dmask >> j) & 1u) s_list[pos++] = (unsigned short) (row * TS + col0 + j);
        }
        __syncthreads();
        const int n = *(volatile int *) &s_n[par];
        if (n == 0) {
            if (tid == 0) s_n[par ^ 1] = 0;
            break;
        }
        // (3) one element per thread; a lowered element marks the neighbours it has read as higher
        for (int i = tid; i < n; i += NT) {
            const int e = s_list[i];
            const int r = e >> 5, c = e & 31;
            const volatile float *wb = sg + r * GP + c;
            Nbr8 nb;
            nb.nw = wb[0]; nb.n = wb[1]; nb.ne = wb[2];
            nb.w = wb[GP]; const float cur = wb[GP + 1]; nb.e = wb[GP + 2];
            nb.sw = wb[2 * GP]; nb.s = wb[2 * GP + 1]; nb.se = wb[2 * GP + 2];
            float v;
            if (KIND == TASI_KIND_DFM) v = dfm_cell(nb, sc[r * CP + c]);
            else {
                const float tl = sc[r * CP + c], tr = sc[r * CP + c + 1], bl = sc[(r + 1) * CP + c], br = sc[(r + 1) * CP + c + 1];
                v = (KIND == TASI_KIND_FD) ? fd_node(nb, tl, tr, bl, br) : mfd_node(nb, tl, tr, bl, br);
            }
            if (v < cur) {
                sg[(r + 1) * GP + c + 1] = v;
                s_chgmap[e] = 1;
                ++cnt_changed;
                unsigned char *dm = s_dirty + e;
                const bool up = r > 0, dn = r < TS - 1, lf = c > 0, rt = c < TS - 1;
                if (up && lf && nb.nw > v) dm[-TS - 1] = 1;
                if (up && nb.n > v) dm[-TS] = 1;
                if (up && rt && nb.ne > v) dm[-TS + 1] = 1;
                if (lf && nb.w > v) dm[-1] = 1;
                if (rt && nb.e > v) dm[1] = 1;
                if (dn && lf && nb.sw > v) dm[TS - 1] = 1;
                if (dn && nb.s > v) dm[TS] = 1;
                if (dn && rt && nb.se > v) dm[TS + 1] = 1;
            }
            ++evals;
        }
        if (tid == 0) s_n[par ^ 1] = 0;  // the next round's counter; nobody adds to it before the barrier below
        __syncthreads();
        ++iters;
    }
    // what the owner writes back: the final values of its strip and which of them changed during the visit
    __syncthreads();
    const unsigned cb = (SW == 4) ? *(volatile unsigned *) myc : (unsigned) *(volatile unsigned short *) myc;
    chg = 0;
#pragma unroll
    for (int j = 0; j < SW; ++j) {
        if ((cb >> (8 * j)) & 0xffu) chg |= 1u << j;
        w[1][j + 1] = *(const volatile float *) (sg + (row + 1) * GP + col0 + 1 + j);
    }
    return cnt_changed;
}

// ---- in-tile relaxation, compacted inside each warp (CMP = 2) ------------------------------------------------------------
// As relax_tile_compact, but the dirty elements are compacted per WARP (a warp owns SW whole rows of the tile): ballot-free
// prefix sum over the lanes, a per-warp list in shared memory, lane i evaluates the warp's i-th dirty element. No CTA
// barrier is added (the round barrier stays the only one, and two sweeps still share it); what is given up against the
// CTA-wide list is the balance between warps.
template <int KIND, int SW>
__device__ __forceinline__ int relax_tile_wcompact(float *sg, const float *sc, unsigned short *s_wl, int inmask, int row,
                                                   int col0, int rlim, int clim, int pin_r, int pin_c,
                                                   float (&w)[WIN_R][SW + 2], unsigned &chg, unsigned &iters,
                                                   unsigned &evals) {
    constexpr int WIN_C = SW + 2, WELEMS = 32 * SW;  // elements (= SW rows) per warp
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned short *wl = s_wl + warp * WELEMS;
    const int wrow0 = warp * SW, lrow = row - wrow0;
    unsigned live = 0;
#pragma unroll
    for (int j = 0; j < SW; ++j)
        if (row < rlim && col0 + j < clim && !(row == pin_r && col0 + j == pin_c)) live |= 1u << j;
    bool dirty0 = (inmask & IN_ALL) != 0;
    dirty0 |= (inmask & (IN_N | IN_NW | IN_NE)) && row == 0;
    dirty0 |= (inmask & (IN_S | IN_SW | IN_SE)) && row == TS - 1;
    dirty0 |= (inmask & (IN_W | IN_NW | IN_SW)) && col0 == 0;
    dirty0 |= (inmask & (IN_E | IN_NE | IN_SE)) && col0 == TS - SW;
    const float *base = sg + row * GP + col0;
    int cnt_changed = 0;
    chg = 0;
    bool first = true;
    for (;;) {
        bool stored_any = false;
#pragma unroll 1
        for (int rep = 0; rep < 2; ++rep) {
            unsigned coldiff = 0;
#pragma unroll
            for (int a = 0; a < WIN_R; ++a)
#pragma unroll
                for (int b = 0; b < WIN_C; ++b) {
                    const float x = *(const volatile float *) (base + a * GP + b);
                    if (__float_as_int(x) != __float_as_int(w[a][b])) {
                        coldiff |= 1u << b;
                        if (a == 1 && b >= 1 && b <= SW && !first) chg |= 1u << (b - 1);
                    }
                    w[a][b] = x;
                }
            unsigned dmask = 0;
            if (first) dmask = dirty0 ? (1u << SW) - 1u : 0u;
            else {
#pragma unroll
                for (int j = 0; j < SW; ++j)
                    if ((coldiff >> j) & 7u) dmask |= 1u << j;
            }
            dmask &= live;
            first = false;
            const int my = __popc(dmask);
            int incl = my;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int t = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += t;
            }
            const int total = __shfl_sync(0xffffffffu, incl, 31);
            if (total) {  // uniform over the warp
                int pos = incl - my;
#pragma unroll
                for (int j = 0; j < SW; ++j)
                    if ((dmask >> j) & 1u) wl[pos++] = (unsigned short) (lrow * TS + col0 + j);
                __syncwarp();
                for (int i = lane; i < total; i += 32) {
                    const int e = wl[i], r = wrow0 + (e >> 5), c = e & 31;
                    const volatile float *wb = sg + r * GP + c;
                    Nbr8 nb;
                    nb.nw = wb[0]; nb.n = wb[1]; nb.ne = wb[2];
                    nb.w = wb[GP]; const float cur = wb[GP + 1]; nb.e = wb[GP + 2];
                    nb.sw = wb[2 * GP]; nb.s = wb[2 * GP + 1]; nb.se = wb[2 * GP + 2];
                    float v;
                    if (KIND == TASI_KIND_DFM) v = dfm_cell(nb, sc[r * CP + c]);
                    else {
                        const float tl = sc[r * CP + c], tr = sc[r * CP + c + 1], bl = sc[(r + 1) * CP + c],
                                    br = sc[(r + 1) * CP + c + 1];
                        v = (KIND == TASI_KIND_FD) ? fd_node(nb, tl, tr, bl, br) : mfd_node(nb, tl, tr, bl, br);
                    }
                    if (v < cur) {
                        sg[(r + 1) * GP + c + 1] = v;
                        ++cnt_changed;
                        stored_any = true;
                    }
                    ++evals;
                }
                __syncwarp();
            }
            ++iters;
        }
        if (!__syncthreads_or(stored_any)) break;
    }
    return cnt_changed;
}

// ---- TMA staging (measured alternative, TASI_TMA=1) ----------------------------------------------------------------------
// The tile's 34 x 34 value window and its 33 x 33 cost window are boxes of two 2-D tensor maps over the padded arrays
// (rows are 128-byte multiples): one elected thread issues cp.async.bulk.tensor.2d for each, the CTA waits on an mbarrier
// for the bytes, then re-pitches the landing buffers into the conflict-free pitch-35 arrays the relaxation reads. The
// inner box origin must be 16-byte aligned (an unaligned inner coordinate raises "illegal instruction" on sm_100a,
// scripts/probes/tma_probe.cu) and the inner extent a multiple of 16 bytes, so the boxes start at column y0 - 4 (values,
// 40 floats) and y0 - 16 (costs, 48 bytes), as the vector loads do; with pitch 40 the strip windows would collide. Values written by other CTAs reach L2 through
// the generic proxy; the issuing thread orders them before its bulk read with fence.proxy.async.
constexpr int TMA_GW = 40, TMA_CW = 48;  // landing pitches: floats per value row, bytes per cost row
constexpr int TMA_GOFF = 4, TMA_COFF = 16;  // columns the boxes start left of the tile

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t) __cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    uint32_t ok = 0;
    const uint32_t a = smem_u32(bar);
    do {
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                     : "=r"(ok) : "r"(a), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void tma_load_2d(void *dst, const void *tmap, uint64_t *bar, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(smem_u32(dst)), "l"(tmap), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async;" ::: "memory"); }

// ---- the persistent solver kernel ------------------------------------------------------------------------------
// Staging geometry: values are fetched as 16-byte vectors from column y0-4 (rows are 128-byte aligned and the
// arrays carry GMARGIN columns of margin), 10 vectors cover columns y0-4..y0+35 of one row; costs likewise as
// 4-byte vectors. Every thread issues all of its loads before the first use, so a tile is staged in about one
// L2 round trip.
constexpr int GV_PER_ROW = 10;
constexpr int G_VECS = (TS + 2) * GV_PER_ROW;                                  // 340 float4 per tile
constexpr int C_ROWS_NODE = TS + 1, C_ROWS_CELL = TS;

// SW = strip width: 4 (256 threads, two CTAs per SM) or 2 (512 threads, one CTA per SM: half the serial work per
// thread in a round, for the node planners on grids whose front never feeds more than one CTA per SM anyway).
// CMP = 1 / 2: the relax visits use relax_tile_compact / relax_tile_wcompact (MODE 0 only).
// TMA = 1: the windows are staged by cp.async.bulk.tensor.2d (see "TMA staging" above) instead of vector loads.
template <int KIND, int MODE, int SW, int CMP = 0, int TMA = 0>
#ifndef CMP1_BLOCKS
#define CMP1_BLOCKS 3  // resident CTAs per SM the CTA-list instances with 4-element strips are compiled for (80 registers,
                       // no spills in the pushed-mark instances; run_solver launches 3 per SM on large grids only)
#endif
__global__ void __launch_bounds__(1024 / SW, SW == 4 ? ((CMP == 3 && MODE == 0) ? CMP1_BLOCKS : 2) : 1)
solver_kernel(GridDesc d, const __grid_constant__ TmaMaps tm) {
    constexpr int NT = (TS * TS) / SW, WIN_C = SW + 2;
    constexpr int G_ITERS = (G_VECS + NT - 1) / NT, C_ITERS = (C_ROWS_NODE * GV_PER_ROW + NT - 1) / NT;
    __shared__ float sg[(TS + 2) * GP];   // values, rows/cols -1..32
    __shared__ float sc[(TS + 1) * CP];   // cell costs as float, INF = obstacle
    __shared__ int s_tile, s_inmask, s_again, s_adv;
    __shared__ int s_chain, s_chain_key;  // chaining: the neighbour this CTA goes on with (-1: none)
    __shared__ int s_ring;                // sliding window: the ring the tile being visited was claimed from (-1: none)
    if (threadIdx.x == 0) s_chain = -1;
    __shared__ int s_kmin[8];             // per neighbour direction: lowest changed value it will read (bits)
    __shared__ int s_red[MAX_WARPS];
    __shared__ int s_cmd[2];              // fused mode: command being served by CTA 0
    __shared__ unsigned short s_list[CMP ? TS * TS : 1];  // compacted relaxation: the dirty elements of a round (CMP = 2:
                                                          // one list of 32 * SW entries per warp)
    __shared__ int s_n[2];                // ... and how many (two counters, alternating by round)
    __shared__ __align__(16) unsigned char s_dirty[CMP == 3 ? TS * TS : 4];   // CMP = 3: elements marked for the next round
    __shared__ __align__(16) unsigned char s_chgmap[CMP == 3 ? TS * TS : 4];  // ... and elements changed during the visit
    __shared__ __align__(128) float s_land[TMA ? (TS + 2) * TMA_GW : 1];        // TMA landing buffers
    __shared__ __align__(128) unsigned char s_cland[TMA ? (TS + 1) * TMA_CW : 16];
    __shared__ __align__(8) uint64_t s_bar;
    unsigned tma_phase = 0;
    if (TMA) {
        if (threadIdx.x == 0) mbar_init(&s_bar, 1);
        __syncthreads();
    }
    if (CMP && threadIdx.x == 0) s_n[0] = s_n[1] = 0;
    const bool fused = d.ft != nullptr;
    unsigned long long ticket = ~0ULL;    // fused mode, CTA 0: the queue ticket it is waiting on

    const float INF = __int_as_float(0x7f800000);
    const int tid = threadIdx.x, lane = tid & 31;
    const int row = tid / (TS / SW), col0 = (tid % (TS / SW)) * SW;  // this thread's strip
    unsigned long long n_visits = 0, n_iters = 0, n_changed = 0, n_evals = 0;
    long long c_pop = 0, c_stage = 0, c_relax = 0, c_post = 0, c0 = clock64(), c1;
    const unsigned long long t_launch = (blockIdx.x == 0 && tid == 0) ? globaltimer_ns() : 0ULL;

    // resumed band solve with nothing queued: arm_solver left one virtual pending tile for CTA 0 to retire, which
    // takes the count to 0 and starts the threshold advance
    if (blockIdx.x == 0 && !fused) {
        if (tid == 0) {
            s_adv = 0;
            if (*(volatile int *) &d.ctl->kick) {
                d.ctl->kick = 0;
                s_adv = (atomicSub(&d.ctl->pending, 1) == 1) ? 1 : 0;
            }
        }
        drain_and_advance(d, s_red, &s_adv, tid);
    }

    for (;;) {
        if (d.nb > 0 && tid < 32) {  // bucketed near queue: warp 0 pops together
            int t = -1;  // kept from the last visit (chaining)? (reset after the barrier below: the other threads still read it)
            if (tid == 0 && s_chain >= 0) t = s_chain;
            t = __shfl_sync(0xffffffffu, t, 0);
            if (t >= 0 && d.defer_margin >= 0.0f && !defer_or_run(d, t, tid)) t = -1;
            if (t < 0) {
                do {
                    if (tid == 0) s_ring = -1;
                    t = q_pop_warp(d, tid, fused && blockIdx.x == 0, &s_ring);
                } while (t >= 0 && d.defer_margin >= 0.0f && !defer_or_run(d, t, tid));
            }
            if (tid == 0) {
                if (t >= 0) {
                    if (d.defer_margin >= 0.0f) d.rkey[t] = *(volatile int *) &d.key[t];
                    const int old = atomicExch(&d.state[t], ST_RUN);
                    s_inmask = old >> ST_SHIFT;
                    atomicExch(&d.key[t], INF_BITS);
                    fence_gpu();
                }
                s_tile = t;
            }
        } else if (d.nb <= 0 && tid == 0) {
            int t = (fused && blockIdx.x == 0) ? q_pop_fused(d, &ticket) : q_pop(d);
            if (t >= 0) {
                // claim the tile: the word we take away carries the direction bits of every activation so far;
                // their values were fenced before the bits were set, the fence below orders our loads after it
                const int old = atomicExch(&d.state[t], ST_RUN);
                s_inmask = old >> ST_SHIFT;
                atomicExch(&d.key[t], INF_BITS);
                fence_gpu();
            }
            s_tile = t;
        }
        if (tid < 8) s_kmin[tid] = INF_BITS;
        __syncthreads();
        if (tid == 0) s_chain = -1;
        const int tile = s_tile;
        if (tile == -2) {  // fused mode, CTA 0: a phase boundary / a command of the coordinator
            fused_service(d, s_red, s_cmd, tid);
            continue;
        }
        if (tile < 0) break;
        c1 = clock64(); c_pop += c1 - c0; c0 = c1;
        const int tx = tile / d.TY, ty = tile - tx * d.TY;
        const int x0 = tx * TS, y0 = ty * TS;
        const int rlim = min(TS, d.EX - x0), clim = min(TS, d.EY - y0);
        const int pin_r = d.pin_x - x0, pin_c = d.pin_y - y0;

        // cell costs around the tile (immutable while the solver runs): issue the loads now, convert after the
        // value loads of the first staging are in flight too
        constexpr int c_r0 = (KIND == TASI_KIND_DFM) ? 0 : -1;  // first cost row / column relative to the tile
        constexpr int c_rows = (KIND == TASI_KIND_DFM) ? C_ROWS_CELL : C_ROWS_NODE;
        uchar4 cv[C_ITERS];
        if (!TMA) {
#pragma unroll
            for (int k = 0; k < C_ITERS; ++k) {
                const int i = tid + k * NT;
                const int r = i / GV_PER_ROW, v = i - r * GV_PER_ROW;
                if (i < c_rows * GV_PER_ROW)
                    cv[k] = __ldg((const uchar4 *) (d.cost + (int64_t) (x0 + c_r0 + r) * d.cpitch + (y0 - 4 + 4 * v)));
            }
        }
        bool first = true;

        bool again;
        do {
            if (TMA) {
                // coordinates in the padded arrays: element (ex, ey) sits at row ex + 1, column GMARGIN + ey
                if (tid == 0) {
                    fence_proxy_async();
                    mbar_expect_tx(&s_bar, (unsigned) ((TS + 2) * TMA_GW * 4 + (first ? (TS + 1) * TMA_CW : 0)));
                    tma_load_2d(s_land, &tm.g, &s_bar, GMARGIN + y0 - TMA_GOFF, x0);
                    if (first) tma_load_2d(s_cland, &tm.c, &s_bar, GMARGIN + y0 - TMA_COFF, x0 + c_r0 + 1);
                }
                mbar_wait(&s_bar, tma_phase);
                tma_phase ^= 1u;
                for (int i = tid; i < (TS + 2) * (TS + 2); i += NT) {
                    const int r = i / (TS + 2), c = i - r * (TS + 2);
                    sg[r * GP + c] = s_land[r * TMA_GW + c + TMA_GOFF - 1];
                }
                if (first) {
                    first = false;
                    for (int i = tid; i < c_rows * c_rows; i += NT) {
                        const int r = i / c_rows, c = i - r * c_rows;
                        const int b = s_cland[r * TMA_CW + c + TMA_COFF + c_r0];
                        sc[r * CP + c] = (b >= d.thr) ? INF : (float) b;
                    }
                }
            } else {
            // stage values + halo ring; L1 is bypassed (other CTAs write these lines)
            float4 gv[G_ITERS];
#pragma unroll
            for (int k = 0; k < G_ITERS; ++k) {
                const int i = tid + k * NT;
                const int r = i / GV_PER_ROW, v = i - r * GV_PER_ROW;
                if (i < G_VECS)
                    gv[k] = __ldcg((const float4 *) (d.g + (int64_t) (x0 - 1 + r) * d.gpitch + (y0 - 4 + 4 * v)));
            }
            if (first) {
                first = false;
#pragma unroll
                for (int k = 0; k < C_ITERS; ++k) {
                    const int i = tid + k * NT;
                    const int r = i / GV_PER_ROW, v = i - r * GV_PER_ROW;
                    if (i < c_rows * GV_PER_ROW) {
                        const unsigned char bb[4] = {cv[k].x, cv[k].y, cv[k].z, cv[k].w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const int c = 4 * v + e - 4 - c_r0;  // column index inside sc
                            if (c >= 0 && c < c_rows) sc[r * CP + c] = (bb[e] >= d.thr) ? INF : (float) bb[e];
                        }
                    }
                }
            }
#pragma unroll
            for (int k = 0; k < G_ITERS; ++k) {
                const int i = tid + k * NT;
                const int r = i / GV_PER_ROW, v = i - r * GV_PER_ROW;
                if (i < G_VECS) {
                    const float ff[4] = {gv[k].x, gv[k].y, gv[k].z, gv[k].w};
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const int c = 4 * v + e - 3;  // sg column: element column y0-1 is column 0
                        if (c >= 0 && c < TS + 2) sg[r * GP + c] = ff[e];
                    }
                }
            }
            }
            __syncthreads();
            ++n_visits;
            c1 = clock64(); c_stage += c1 - c0; c0 = c1;

            unsigned it = 0, ev = 0, chg = 0;
            float w[WIN_R][WIN_C];
#pragma unroll
            for (int a2 = 0; a2 < WIN_R; ++a2)
#pragma unroll
                for (int b2 = 0; b2 < WIN_C; ++b2) w[a2][b2] = 0.0f;
            if (CMP == 3 && MODE == 0)
                n_changed += relax_tile_push<KIND, SW>(sg, sc, s_list, s_n, s_dirty, s_chgmap, s_inmask, row, col0, rlim, clim,
                                                       pin_r, pin_c, w, chg, it, ev);
            else if (CMP == 2 && MODE == 0)
                n_changed += relax_tile_wcompact<KIND, SW>(sg, sc, s_list, s_inmask, row, col0, rlim, clim, pin_r, pin_c, w,
                                                           chg, it, ev);
            else if (CMP == 1 && MODE == 0)
                n_changed += relax_tile_compact<KIND, SW>(sg, sc, s_list, s_n, s_inmask, row, col0, rlim, clim, pin_r, pin_c,
                                                          w, chg, it, ev, &d.ctl->kick);
            else
                n_changed += relax_tile<KIND, MODE, SW>(sg, sc, s_inmask, row, col0, rlim, clim, pin_r, pin_c, d.inval_tol, w,
                                                        chg, it, ev);
            n_iters += it;   // counted once per CTA below (every thread sees the same number of rounds)
            n_evals += ev;
            c1 = clock64(); c_relax += c1 - c0; c0 = c1;

            // write back the strip elements that changed; per neighbour direction keep the lowest changed value
            // it reads
            if (chg) {
                float *gp = d.g + (int64_t) (x0 + row) * d.gpitch + (y0 + col0);
                if (SW == 4 && chg == 0xfu) {
                    *(float4 *) gp = make_float4(w[1][1], w[1][2], w[1][3], w[1][SW]);
                } else {
#pragma unroll
                    for (int j = 0; j < SW; ++j)
                        if ((chg >> j) & 1u) gp[j] = w[1][j + 1];
                }
                if (fused && (row == 0 || row == TS - 1)) {
                    // fused mode: a band border row goes straight into the halo row of the rank across the edge
                    const FusedTopo &ft = *d.ft;
                    const int bp = d.band_rt + 1, bt = tx % bp, jb = tx / bp, gb = jb * ft.world + ft.rank;
                    float *rp = nullptr;
                    if (row == 0 && bt == 0 && gb > 0) {
                        const int ju = ft.rank > 0 ? jb : jb - 1;  // the band above is the neighbour's band ju
                        rp = ft.nbr[0].g + (int64_t) ((ju * bp + d.band_rt) * TS) * ft.nbr[0].gpitch;
                    } else if (row == TS - 1 && bt == d.band_rt - 1 && gb + 1 < ft.parts) {
                        const int jd = ft.rank < ft.world - 1 ? jb : jb + 1;
                        rp = ft.nbr[1].g + ((int64_t) (jd * bp) * TS - 1) * ft.nbr[1].gpitch;
                    }
                    if (rp) {
                        rp += y0 + col0;
#pragma unroll
                        for (int j = 0; j < SW; ++j)
                            if ((chg >> j) & 1u) __stcg(rp + j, w[1][j + 1]);
                    }
                }
                int kb = INF_BITS, kfirst = INF_BITS, klast = INF_BITS;
#pragma unroll
                for (int j = 0; j < SW; ++j)
                    if ((chg >> j) & 1u) {
                        const int bits = (MODE == 0) ? __float_as_int(w[1][j + 1]) : 0;
                        kb = min(kb, bits);
                        if (j == 0) kfirst = bits;
                        if (j == SW - 1) klast = bits;
                    }
                const bool wcol = col0 == 0 && kfirst != INF_BITS, ecol = col0 == TS - SW && klast != INF_BITS;
                if (row == 0) {
                    atomicMin(&s_kmin[0], kb);
                    if (wcol) atomicMin(&s_kmin[4], kfirst);
                    if (ecol) atomicMin(&s_kmin[5], klast);
                }
                if (row == TS - 1) {
                    atomicMin(&s_kmin[1], kb);
                    if (wcol) atomicMin(&s_kmin[6], kfirst);
                    if (ecol) atomicMin(&s_kmin[7], klast);
                }
                if (wcol) atomicMin(&s_kmin[2], kfirst);
                if (ecol) atomicMin(&s_kmin[3], klast);
            }
            __syncthreads();
            // the CTA's write-back (ordered by the barrier) before the flags below; the leaving thread sits in another
            // warp than the activating ones, so its atomics overlap theirs instead of following them
            if (tid < 8 || tid == LEAVE_TID) {
                // fused mode: only a tile on a band border stores into another GPU's memory and raises its flags; the
                // values of an interior tile are read by CTAs of this GPU alone, for which the device-scope fence does
                if (fused && (tx % (d.band_rt + 1) == 0 || tx % (d.band_rt + 1) == d.band_rt - 1)) fence_sys();
                else fence_gpu();
            }

            // activate the neighbours that read a changed border, one thread per direction
            // (0 N, 1 S, 2 W, 3 E, 4 NW, 5 NE, 6 SW, 7 SE); our north border is the south halo of the tile above
            if (tid < 8) {
                const int k = s_kmin[tid];
                s_kmin[tid] = INF_BITS;
                const int dx = (0x31 >> tid & 1) ? -1 : ((0xc2 >> tid & 1) ? 1 : 0);
                const int dy = (0x54 >> tid & 1) ? -1 : ((0xa8 >> tid & 1) ? 1 : 0);
                const int seen_as = (IN_S << 0) * (tid == 0) + IN_N * (tid == 1) + IN_E * (tid == 2) + IN_W * (tid == 3) +
                                    IN_SE * (tid == 4) + IN_SW * (tid == 5) + IN_NE * (tid == 6) + IN_NW * (tid == 7);
                const int nx = tx + dx, ny = ty + dy;
                const bool chain = MODE == 0 && d.chain && d.nb > 0 && !fused && d.defer_margin < 0.0f;
                bool held = false;
                if (k != INF_BITS && nx >= 0 && nx < d.TX && ny >= 0 && ny < d.TY && !dead_tile_row(d, nx))
                    held = activate(d, nx * d.TY + ny, k, seen_as, chain);
                if (chain) {
                    // chaining: of the neighbours this visit woke from idle, the CTA keeps the one with the lowest key for
                    // itself (it is not queued: no push, no pop; the pending count of the ending visit passes to it); the
                    // others go to the rings
                    const unsigned hm = __ballot_sync(0xffu, held);
                    int best = -1, bk = INF_BITS;
                    for (int l = 0; l < 8; ++l) {
                        const int kl = __shfl_sync(0xffu, k, l);
                        if (((hm >> l) & 1u) && kl < bk) { bk = kl; best = l; }
                    }
                    if (held && tid != best) {
                        q_push_key(d, nx * d.TY + ny, k);
                        atomicAdd(&d.ctl->pushes, 1ULL);
                    }
                    if (tid == best) { s_chain = nx * d.TY + ny; s_chain_key = k; }
                } else if (fused && k != INF_BITS && dx != 0 && ny >= 0 && ny < d.TY) {
                    // across a band edge: the tile belongs to a neighbour rank
                    const FusedTopo &ft = *d.ft;
                    const int bp = d.band_rt + 1, bt = tx % bp, jb = tx / bp, gb = jb * ft.world + ft.rank;
                    if (dx < 0 && bt == 0 && gb > 0) {
                        const int ju = ft.rank > 0 ? jb : jb - 1;
                        activate_remote(d, ft.nbr[0], ft.fc[(ft.rank + ft.world - 1) % ft.world],
                                        (ju * bp + d.band_rt - 1) * d.TY + ny, k, seen_as);
                    } else if (dx > 0 && bt == d.band_rt - 1 && gb + 1 < ft.parts) {
                        const int jd = ft.rank < ft.world - 1 ? jb : jb + 1;
                        activate_remote(d, ft.nbr[1], ft.fc[(ft.rank + 1) % ft.world], (jd * bp) * d.TY + ny, k, seen_as);
                    }
                }
            }
            // ... while thread 32 leaves the tile (a neighbour may re-queue it at once: we no longer touch its values)
            if (tid == LEAVE_TID) {
                if (MODE == 1) d.seed[tile] = 1;  // phase 2 revisits every tile phase 1 looked at
                // leave the tile, unless a neighbour changed our halo meanwhile
                int again_f = 0;
                int old = atomicCAS(&d.state[tile], ST_RUN, 0);
                if (old != ST_RUN) {  // RUN|DIRTY + direction bits
                    const int kk = atomicMin(&d.key[tile], INF_BITS);  // atomic read
                    if (kk < cur_thr_bits(d) || kk == INF_BITS) {
                        // due now (or its key was consumed with the previous staging): go round again
                        if (d.nb > 0 && d.defer_margin >= 0.0f && kk != INF_BITS) d.rkey[tile] = kk;
                        old = atomicExch(&d.state[tile], ST_RUN);
                        s_inmask = old >> ST_SHIFT;
                        atomicExch(&d.key[tile], INF_BITS);
                        fence_gpu();
                        again_f = 1;
                    } else {
                        atomicXor(&d.state[tile], ST_RUN | ST_FAR);  // RUN|DIRTY -> DIRTY|FAR, direction bits stay
                        far_append(d, tile);
                    }
                }
                if (*(volatile int *) &d.ctl->error) again_f = 0;
                s_again = again_f;
            }
            __syncthreads();
            again = s_again != 0;
            if (again && tid == 0 && s_chain >= 0) {  // the visit goes on: the kept neighbour goes to the rings after all
                q_push_key(d, s_chain, s_chain_key);
                atomicAdd(&d.ctl->pushes, 1ULL);
                s_chain = -1;
            }
            if (!again && d.nb > 0 && d.defer_margin >= 0.0f && tid >= 32 && tid < 40) {
                fence_gpu();  // our state word left RUN before we look at the neighbours' (see defer_or_run)
                release_deferred(d, tile, tid - 32);
            }
            c1 = clock64(); c_post += c1 - c0; c0 = c1;
        } while (again);

        // the tile is no longer pending; whoever takes the count to 0 advances the threshold (fused mode: tells the
        // coordinator that this rank has gone idle; thresholds advance globally, fused_service)
        if (fused) {
            if (tid == 0 && atomicSub(&d.ctl->pending, 1) == 1) fused_rank_idles(d);
            continue;
        }
        if (s_chain >= 0) continue;  // chaining: the pending count of this visit passes to the kept neighbour
        if (tid == 0) {
            // sliding window: the visit is over and its activations are queued — its ring may be passed by the base
            if (d.slide && s_ring >= 0) atomicSub(&d.ctl->brun[s_ring], 1);
            s_adv = (atomicSub(&d.ctl->pending, 1) == 1) ? 1 : 0;
        }
        drain_and_advance(d, s_red, &s_adv, tid);
    }

    // per-CTA counters -> global statistics
    for (int o = 16; o > 0; o >>= 1) {
        n_changed += __shfl_down_sync(0xffffffffu, n_changed, o);
    }
    if (lane == 0 && n_changed) atomicAdd(MODE == 0 ? &d.ctl->lowered : &d.ctl->invalidated, n_changed);
    for (int o = 16; o > 0; o >>= 1) n_evals += __shfl_down_sync(0xffffffffu, n_evals, o);
    if (lane == 0 && n_evals) atomicAdd(&d.ctl->evals, n_evals);
    if (tid == 0) {
        if (blockIdx.x == 0) {
            atomicAdd(&d.ctl->busy_ns, globaltimer_ns() - t_launch);
            atomicAdd(&d.ctl->launches, 1);
        }
        atomicAdd(&d.ctl->visits, n_visits);
        atomicAdd(&d.ctl->rounds, n_iters);
        atomicAdd(&d.ctl->cyc_pop, (unsigned long long) c_pop);
        atomicAdd(&d.ctl->cyc_stage, (unsigned long long) c_stage);
        atomicAdd(&d.ctl->cyc_relax, (unsigned long long) c_relax);
        atomicAdd(&d.ctl->cyc_post, (unsigned long long) c_post);
    }
}

// ---- auxiliary kernels -----------------------------------------------------------------------------------------
// Fill `n4` float4 (the whole padded value array, margins included) with INF.
__global__ void fill_inf_kernel(float4 *p, int64_t n4) {
    const float INF = __int_as_float(0x7f800000);
    const float4 v = make_float4(INF, INF, INF, INF);
    for (int64_t i = blockIdx.x * (int64_t) blockDim.x + threadIdx.x; i < n4; i += (int64_t) gridDim.x * blockDim.x)
        p[i] = v;
}

// init() of the three planners (FieldDPlanner_impl.h:15-21, DynamicFastMarching_impl.h:6-11,
// ShiftedGridPlanner_impl.h:9-15): goal value 0, its tile is the first active tile.
__global__ void seed_goal_kernel(GridDesc d) {
    if (d.pin_x >= 0 && d.pin_x < d.EX && d.pin_y >= 0 && d.pin_y < d.EY) {
        d.g[(int64_t) d.pin_x * d.gpitch + d.pin_y] = 0.0f;
        d.seed[(d.pin_x / TS) * d.TY + (d.pin_y / TS)] = 1;
    }
    // further goals: in the relax mode nothing can lower a 0, in the invalidate mode a 0 is never reset, so they
    // need no per-tile pin
    for (int i = 0; i < d.n_xgoals; ++i) {
        const int2 q = d.xgoals[i];
        if (q.x >= 0 && q.x < d.EX && q.y >= 0 && q.y < d.EY) {
            d.g[(int64_t) q.x * d.gpitch + q.y] = 0.0f;
            d.seed[(q.x / TS) * d.TY + (q.y / TS)] = 1;
        }
    }
}

// Move the tiles flagged in seed[] into the work queue (runs alone, before the solver).
// One warp per flagged tile: its key becomes (at most) the lowest value it can READ — the 34x34 window of the tile
// plus its one-element halo ring — so that seeded tiles enter the value-ordered schedule where they belong instead
// of all running at once with stale inputs (a replan seeds thousands of tiles; the ones nearer the goal must settle
// first). The halo ring matters: a key has to be a lower bound on anything the visit can write (the focus test of
// advance_threshold and the band limit rely on it), and after a cost-lowering patch a tile whose interior is all
// INF (a block of believed obstacles that opened up) takes its new values from its halo.
__global__ void seed_keys_kernel(GridDesc d) {
    const int n = d.TX * d.TY, lane = threadIdx.x & 31;
    const int nwarps = (gridDim.x * blockDim.x) >> 5;
    for (int t = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; t < n; t += nwarps) {
        if (!d.seed[t]) continue;
        const int tx = t / d.TY, ty = t - tx * d.TY;
        // rows -1..32 of the window exist for every tile (margin / halo rows), columns -1 and 32 too (GMARGIN)
        const float *base = d.g + (int64_t) (tx * TS - 1) * d.gpitch + ty * TS;
        int m = INF_BITS;
        for (int r = 0; r < TS + 2; ++r) {
            const float *rp = base + (int64_t) r * d.gpitch;
            m = min(m, __float_as_int(__ldcg(rp + lane)));  // values >= 0: int order == float order
            if (lane < 2) m = min(m, __float_as_int(__ldcg(rp + (lane ? TS : -1))));
        }
        m = __reduce_min_sync(0xffffffffu, m);
        if (lane == 0) atomicMin(&d.key[t], m);
    }
}

// Move the tiles flagged in seed[] into the schedule (runs alone, before the solver): into the work queue, or —
// `ordered` — into the far pile, from where the first threshold advance releases them by key.
__global__ void enqueue_seeds_kernel(GridDesc d, int ordered) {
    const int n = d.TX * d.TY;
    DevCtl *c = d.ctl;
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x) {
        if (d.seed[t]) {
            d.seed[t] = 0;
            // every element of a seeded tile is evaluated (all direction bits)
            const int bits = ST_DIRTY | (0x1ff << ST_SHIFT) | (ordered ? ST_FAR : 0);
            const int old = atomicOr(&d.state[t], bits) & ST_MASK;
            if (!ordered && (old & ST_FAR)) {
                // parked by an earlier focused plan: this launch (the invalidate phase) runs it now, its pile
                // entry goes stale
                atomicAnd(&d.state[t], ~ST_FAR);
                atomicAdd(&c->pending, 1);
                unsigned long long s = atomicAdd(&c->tail, 1ULL);
                d.queue[(unsigned) s & d.qmask] = t + 1;
                atomicAdd(&c->pushes, 1ULL);
            } else if (old == 0) {
                if (ordered) {
                    const int sel = c->far_sel;
                    const int slot = atomicAdd(&c->far_count[sel], 1);
                    if (slot < d.far_cap) { d.far[sel * d.far_cap + slot] = d.slide ? t + 1 : t; continue; }
                    atomicAnd(&d.state[t], ~ST_FAR);  // pile full: run it now
                }
                atomicAdd(&c->pending, 1);
                unsigned long long s = atomicAdd(&c->tail, 1ULL);
                d.queue[(unsigned) s & d.qmask] = t + 1;
                atomicAdd(&c->pushes, 1ULL);
            }
        }
    }
}

// Nothing queued => the solver must see done == 1 or its CTAs would wait for work forever.
__global__ void arm_solver_kernel(GridDesc d) {
    DevCtl *c = d.ctl;
    const bool parked = c->far_count[c->far_sel] > 0;
    c->kick = 0;
    if (c->pending == 0 && parked) {  // only parked tiles: a virtual pending tile makes CTA 0 advance the threshold
        c->pending = 1;
        c->kick = 1;
    }
    c->done = (c->pending == 0) ? 1 : 0;
}

// Graph::update (ProjectToolkit/Graph.cpp:36-51) on the device cost grid: rows x..x+h-1, cols y..y+w-1 take
// the patch bytes; chg[] records which cells changed and whether the traversal cost went up (2) or down (1).
// Changes are counted for the rows cnt_lo <= x + i < cnt_hi only (a band of a sharded grid also stores a halo copy of the
// cell row above it, which its owner counts).
__global__ void patch_apply_kernel(GridDesc d, const uint8_t *patch, uint8_t *chg, int x, int y, int w, int h, int cnt_lo,
                                   int cnt_hi) {
    const int64_t n = (int64_t) w * h;
    unsigned long long c_all = 0, c_up = 0, c_dn = 0;
    for (int64_t i = blockIdx.x * (int64_t) blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        const int pi = (int) (i / w), pj = (int) (i - (int64_t) pi * w);
        uint8_t *mv = d.cost + (int64_t) (x + pi) * d.cpitch + (y + pj);
        const uint8_t ov = *mv, nv = patch[i];
        uint8_t f = 0;
        if (ov != nv) {
            // effective traversal cost: bytes >= thr are all the same obstacle
            const int oe = ov >= d.thr ? 256 : ov, ne = nv >= d.thr ? 256 : nv;
            f = ne > oe ? 2 : (ne < oe ? 1 : 3);  // 3: byte changed, cost did not
            *mv = nv;
            if (x + pi >= cnt_lo && x + pi < cnt_hi) {
                ++c_all;
                if (f == 2) ++c_up;
                if (f == 1) ++c_dn;
            }
        }
        chg[i] = f;
    }
    // one atomic per warp and counter: per-thread atomics on three addresses were ~1 ms of a 2.7 M-cell patch
    const unsigned w_all = __reduce_add_sync(0xffffffffu, (unsigned) c_all);
    const unsigned w_up = __reduce_add_sync(0xffffffffu, (unsigned) c_up);
    const unsigned w_dn = __reduce_add_sync(0xffffffffu, (unsigned) c_dn);
    if ((threadIdx.x & 31) == 0) {
        if (w_all) atomicAdd(&d.ctl->n_changed, (unsigned long long) w_all);
        if (w_up) atomicAdd(&d.ctl->n_raised, (unsigned long long) w_up);
        if (w_dn) atomicAdd(&d.ctl->n_lowered_cells, (unsigned long long) w_dn);
    }
}

// update() of the planners: FD*/MFD* seed the 4 corner nodes of every changed cell
// (FieldDPlanner_impl.h:127-136, ShiftedGridPlanner_impl.h:172-195), MS-DFM the changed cell itself
// (DynamicFastMarching_impl.h:106-112). Counts the distinct seeded elements (num_nodes_updated) and flags
// their tiles. Elements of the patch range: nodes (x..x+h, y..y+w), cells (x..x+h-1, y..y+w-1).
template <int KIND>
// Only elements of the rows seed_lo <= row < seed_hi are seeded and counted (the rows this grid owns).
__global__ void patch_seed_kernel(GridDesc d, const uint8_t *chg, int x, int y, int w, int h, int seed_lo, int seed_hi) {
    const int ew = (KIND == TASI_KIND_DFM) ? w : w + 1, eh = (KIND == TASI_KIND_DFM) ? h : h + 1;
    const int64_t n = (int64_t) ew * eh;
    unsigned long long cnt = 0;
    for (int64_t i = blockIdx.x * (int64_t) blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        const int ei = (int) (i / ew), ej = (int) (i - (int64_t) ei * ew);
        bool hit;
        if (KIND == TASI_KIND_DFM) {
            hit = chg[(int64_t) ei * w + ej] != 0;
        } else {
            hit = false;
            for (int di = -1; di <= 0; ++di)
                for (int dj = -1; dj <= 0; ++dj) {
                    const int ci = ei + di, cj = ej + dj;
                    if (ci >= 0 && ci < h && cj >= 0 && cj < w && chg[(int64_t) ci * w + cj] != 0) hit = true;
                }
        }
        if (hit && x + ei >= seed_lo && x + ei < seed_hi) {
            ++cnt;
            const int gx = x + ei, gy = y + ej;
            if (gx < d.EX && gy < d.EY) d.seed[(gx / TS) * d.TY + (gy / TS)] = 1;
        }
    }
    const unsigned w_cnt = __reduce_add_sync(0xffffffffu, (unsigned) cnt);
    if ((threadIdx.x & 31) == 0 && w_cnt) atomicAdd(&d.ctl->n_seed_elems, (unsigned long long) w_cnt);
}

// Row-band sharding: merge a received boundary row into the halo row (`hrow` = -1 or EX) and flag the
// tiles that read the elements that got lower.
__global__ void band_import_kernel(GridDesc d, const float *stage, int hrow, int trow) {
    unsigned long long cnt = 0;
    for (int yy = blockIdx.x * blockDim.x + threadIdx.x; yy < d.EY; yy += gridDim.x * blockDim.x) {
        float *hp = d.g + (int64_t) hrow * d.gpitch + yy;
        const float nv = stage[yy];
        if (nv < *hp) {
            *hp = nv;
            ++cnt;
            const int ty = yy / TS;
            const int kb = __float_as_int(nv);
            atomicMin(&d.key[trow * d.TY + ty], kb);  // a parked tile must be released when the band reaches nv
            d.seed[trow * d.TY + ty] = 1;
            if ((yy & (TS - 1)) == 0 && ty > 0) {
                atomicMin(&d.key[trow * d.TY + ty - 1], kb);
                d.seed[trow * d.TY + ty - 1] = 1;
            }
            if ((yy & (TS - 1)) == TS - 1 && ty + 1 < d.TY) {
                atomicMin(&d.key[trow * d.TY + ty + 1], kb);
                d.seed[trow * d.TY + ty + 1] = 1;
            }
        }
    }
    if (cnt) atomicAdd(&d.ctl->halo_lowered, cnt);
}

// Sum and count of the free-cell costs (sizes the value band of the near/far schedule).
__global__ void cost_stats_kernel(const uint8_t *cost, int64_t cpitch, int rows, int W, int thr,
                                  unsigned long long *out) {
    const int64_t n = (int64_t) rows * W;
    unsigned long long sum = 0, cnt = 0;
    for (int64_t i = blockIdx.x * (int64_t) blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        const int r = (int) (i / W), c = (int) (i - (int64_t) r * W);
        const uint8_t b = cost[(int64_t) r * cpitch + c];
        if (b < thr) { sum += b; ++cnt; }
    }
    for (int o = 16; o > 0; o >>= 1) {
        sum += __shfl_down_sync(0xffffffffu, sum, o);
        cnt += __shfl_down_sync(0xffffffffu, cnt, o);
    }
    if ((threadIdx.x & 31) == 0 && cnt) {
        atomicAdd(&out[0], sum);
        atomicAdd(&out[1], cnt);
    }
}

// Deterministic synthetic traversability grid (same function as synth.py):
// h = splitmix64(seed ^ (gx*W + gy)); obstacle (255) if h % 1000 < 10*p_obs else 1 + ((h >> 16) % 64);
// 9x9 blocks of cost 1 around (sx,sy) and (gx,gy) (global cell coordinates, negative = none).
__device__ __forceinline__ uint64_t splitmix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

__global__ void synth_map_kernel(uint8_t *cost, int64_t cpitch, int rows, int W, int64_t row0, int dst_row0,
                                 uint64_t seed, int p_obs, int sx, int sy, int gx, int gy) {
    const int64_t n = (int64_t) rows * W;
    for (int64_t i = blockIdx.x * (int64_t) blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        const int r = (int) (i / W), c = (int) (i - (int64_t) r * W);
        const int64_t x = row0 + r;
        const uint64_t hsh = splitmix64(seed ^ ((uint64_t) x * (uint64_t) W + (uint64_t) c));
        uint8_t v = (hsh % 1000ULL) < (uint64_t) (10 * p_obs) ? 255 : (uint8_t) (1 + ((hsh >> 16) % 64ULL));
        if (sx >= 0 && x >= sx - 4 && x <= sx + 4 && c >= sy - 4 && c <= sy + 4) v = 1;
        if (gx >= 0 && x >= gx - 4 && x <= gx + 4 && c >= gy - 4 && c <= gy + 4) v = 1;
        cost[(int64_t) (dst_row0 + r) * cpitch + c] = v;
    }
}

}  // namespace tasi
