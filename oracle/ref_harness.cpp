// TEST INFRASTRUCTURE — not product code.
//
// Thin C-ABI harness around the UNMODIFIED reference planners. It is compiled by
// oracle/Makefile against the sources where they lie under /root/reference (nothing is
// copied); outputs go to oracle/_ref/libref_{fd,dfm,sg}.so. One planner family per
// translation unit because the three reference headers share an include guard and the
// global aliases key_type/elem_type/info_type (FieldDPlanner.h:1,19-27,
// DynamicFastMarching.h:1,19-27, ShiftedGridPlanner.h:19-27).
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg
// may load these libraries.
//
// Select the family with -DREF_FAMILY_FD / -DREF_FAMILY_DFM / -DREF_FAMILY_SG.
#include <cstdint>
#include <cstring>
#include <cmath>
#include <memory>
#include <vector>
#include <limits>
#include <iostream>
#include <chrono>
#include <tuple>

// Operator-level taps (ref_op_*) call the planners' private compute_optimal_cost: open the access
// specifiers for THIS translation unit only, after every standard header the reference pulls in has been
// included (so the library headers are untouched); the reference sources themselves stay unmodified.
#include <algorithm>
#include <array>
#include <cassert>
#include <functional>
#include <map>
#include <numeric>
#include <queue>
#include <set>
#include <string>
#include <unordered_map>
#include <unordered_set>
#include <utility>
#define private public
#define protected public
#if defined(REF_FAMILY_FD)
#include "FieldDPlanner.h"
template<int O> using PlannerT = FieldDPlanner<O>;
static const int kMaxLevel = 1;
static const bool kNodeBased = true;
#elif defined(REF_FAMILY_DFM)
#include "DynamicFastMarching.h"
template<int O> using PlannerT = DFMPlanner<O>;
static const int kMaxLevel = 1;
static const bool kNodeBased = false;
#elif defined(REF_FAMILY_SG)
#include "ShiftedGridPlanner.h"
template<int O> using PlannerT = ShiftedGridPlanner<O>;
static const int kMaxLevel = 2;
static const bool kNodeBased = true;
#else
#error "define REF_FAMILY_FD, REF_FAMILY_DFM or REF_FAMILY_SG"
#endif
#include "LinearInterpolationPathExtractor.h"

namespace {

struct IRef {
    virtual ~IRef() {}
    virtual void set_map(const uint8_t *m, int w, int l, float occ) = 0;
    virtual void patch_map(const uint8_t *p, int x, int y, int w, int h) = 0;
    virtual void set_start(float x, float y) = 0;
    virtual void set_goal(float x, float y) = 0;
    virtual void reset() = 0;
    virtual int step() = 0;
    virtual void times(float *u, float *p) = 0;
    virtual void counters(uint64_t *upd, uint64_t *exp) = 0;
    virtual uint64_t map_size() = 0;
    virtual float top_key(int *empty) = 0;
    virtual void dump(float *g, float *rhs) = 0;
    virtual int step_costs(float *out, int cap) = 0;
    virtual int dump_info(int32_t *out) = 0;
    virtual int extract(int use_field, const float *field, int max_steps, int lookahead,
                        int allow_indirect, float *xy, int cap, float *cost, float *dist,
                        float *e_ms) = 0;
    virtual const uint8_t *map_bytes() = 0;
    virtual float op_tri(int nx, int ny, int ax, int ay, int bx, int by, float ga, float gb) = 0;
    virtual float op_stencil(float ga, float gb, float tau, float h) = 0;
};

// INFO of an expanded element (the level-1 back-pointers): a Node for FD*/MFD* (the first node of the triangle
// (b, ccw(b)) that gave rhs, FieldDPlanner_impl.h:91-95), a pair of Cells for MS-DFM (DynamicFastMarching_impl.h:
// 322-342); level 0 keeps none (two-element tuples: the variadic overload answers 0)
static void put_info(const Node &n, int32_t *o) { o[0] = n.x; o[1] = n.y; o[2] = o[3] = INT32_MIN; }
static void put_info(const std::pair<Cell, Cell> &p, int32_t *o) {
    o[0] = p.first.x; o[1] = p.first.y; o[2] = p.second.x; o[3] = p.second.y;
}
template<typename Tuple>
static typename std::enable_if<std::tuple_size<Tuple>::value == 3, int>::type info_of(const Tuple &t, int32_t *o) {
    put_info(std::get<2>(t), o);
    return 1;
}
template<typename Tuple>
static typename std::enable_if<std::tuple_size<Tuple>::value != 3, int>::type info_of(const Tuple &, int32_t *) {
    return 0;
}

template<int O>
struct Ref : IRef {
    typedef PlannerT<O> P;
    typedef typename P::Map Map;
    typedef typename P::Map::ElemType Elem;
    typedef typename P::Base::Info Info;
    P planner;
    std::shared_ptr<uint8_t> data;
    int W = 0, L = 0;

    int ex() const { return kNodeBased ? L + 1 : L; }
    int ey() const { return kNodeBased ? W + 1 : W; }

    void set_map(const uint8_t *m, int w, int l, float occ) override {
        W = w; L = l;
        data = std::shared_ptr<uint8_t>(new uint8_t[(size_t) w * l], std::default_delete<uint8_t[]>());
        std::memcpy(data.get(), m, (size_t) w * l);
        planner.set_occupancy_threshold(occ);
        planner.set_map(data, w, l);
    }
    void patch_map(const uint8_t *p, int x, int y, int w, int h) override {
        std::shared_ptr<uint8_t> patch(new uint8_t[(size_t) w * h], std::default_delete<uint8_t[]>());
        std::memcpy(patch.get(), p, (size_t) w * h);
        planner.patch_map(patch, x, y, w, h);
    }
    void set_start(float x, float y) override { planner.set_start(Position(x, y)); }
    void set_goal(float x, float y) override { planner.set_goal(Position(x, y)); }
    void reset() override { planner.reset(); }
    int step() override { return planner.step(); }
    void times(float *u, float *p) override { *u = planner.u_time; *p = planner.p_time; }
    void counters(uint64_t *upd, uint64_t *exp) override {
        *upd = planner.num_nodes_updated; *exp = planner.num_nodes_expanded;
    }
    uint64_t map_size() override { return planner.map.size(); }
    float top_key(int *empty) override {
        *empty = planner.priority_queue.empty() ? 1 : 0;
        if (*empty) return std::numeric_limits<float>::infinity();
#ifdef NO_HEURISTIC
        return planner.priority_queue.top_key();
#else
        return planner.priority_queue.top_key().first;
#endif
    }
    void dump(float *g, float *rhs) override {
        const float inf = std::numeric_limits<float>::infinity();
        size_t n = (size_t) ex() * ey();
        for (size_t i = 0; i < n; ++i) { g[i] = inf; rhs[i] = inf; }
        for (auto &b : planner.map.buckets)
            for (auto &kv : b) {
                const Elem &e = kv.first;
                if (e.x < 0 || e.y < 0 || e.x >= ex() || e.y >= ey()) continue;
                size_t i = (size_t) e.x * ey() + e.y;
                g[i] = std::get<0>(kv.second);
                rhs[i] = std::get<1>(kv.second);
            }
    }
    int dump_info(int32_t *out) override {
        size_t n = (size_t) ex() * ey();
        for (size_t i = 0; i < 4 * n; ++i) out[i] = INT32_MIN;
        int have = 0;
        for (auto &b : planner.map.buckets)
            for (auto &kv : b) {
                const Elem &e = kv.first;
                if (e.x < 0 || e.y < 0 || e.x >= ex() || e.y >= ey()) continue;
                have |= info_of(kv.second, out + 4 * ((size_t) e.x * ey() + e.y));
            }
        return have;
    }
    int extract(int use_field, const float *field, int max_steps, int lookahead, int allow_indirect,
                float *xy, int cap, float *cost, float *dist, float *e_ms) override {
        Map local;
        const Map *src = &planner.get_expanded_map();
        if (use_field) {
            local.init(planner.grid.width_, planner.grid.length_, 8);
            for (int x = 0; x < ex(); ++x)
                for (int y = 0; y < ey(); ++y) {
                    float v = field[(size_t) x * ey() + y];
                    if (v < std::numeric_limits<float>::infinity())
                        local.insert_or_assign(Elem(x, y), v, v);
                }
            src = &local;
        }
        LinearInterpolationPathExtractor<Elem, Info> extractor(*src, planner.get_grid());
        extractor.max_steps = max_steps;
        extractor.lookahead = lookahead != 0;
        extractor.allow_indirect_traversals = allow_indirect != 0;
        extractor.extract_path();
        int n = (int) extractor.path_.size();
        for (int i = 0; i < n && i < cap; ++i) {
            xy[2 * i] = extractor.path_[i].x;
            xy[2 * i + 1] = extractor.path_[i].y;
        }
        *cost = extractor.total_cost;
        *dist = extractor.total_dist;
        *e_ms = extractor.e_time;
        last_costs = extractor.cost_;
        return n;
    }
    std::vector<float> last_costs;
    int step_costs(float *out, int cap) override {
        for (int i = 0; i < (int) last_costs.size() && i < cap; ++i) out[i] = last_costs[i];
        return (int) last_costs.size();
    }
    const uint8_t *map_bytes() override { return data.get(); }
    // the planner's own local update on caller-given neighbour values (costs come from the planner's map)
    float op_tri(int nx, int ny, int ax, int ay, int bx, int by, float ga, float gb) override {
#if defined(REF_FAMILY_DFM)
        (void) nx; (void) ny; (void) ax; (void) ay; (void) bx; (void) by; (void) ga; (void) gb;
        return std::numeric_limits<float>::quiet_NaN();
#else
        return planner.compute_optimal_cost(Node(nx, ny), Node(ax, ay), Node(bx, by), ga, gb);
#endif
    }
    float op_stencil(float ga, float gb, float tau, float h) override {
#if defined(REF_FAMILY_DFM)
        Cell ca(0, 0), cb(1, 1);
        return planner.compute_optimal_cost(ca, cb, ga, gb, tau, h).first;
#else
        (void) ga; (void) gb; (void) tau; (void) h;
        return std::numeric_limits<float>::quiet_NaN();
#endif
    }
};

}  // namespace

extern "C" {

int ref_family_node_based() { return kNodeBased ? 1 : 0; }
int ref_max_level() { return kMaxLevel; }

void *ref_create(int level) {
    switch (level) {
        case 0: return new Ref<0>();
        case 1: return new Ref<1>();
#if defined(REF_FAMILY_SG)
        case 2: return new Ref<2>();
#endif
        default: return nullptr;
    }
}
void ref_destroy(void *h) { delete static_cast<IRef *>(h); }
void ref_set_map(void *h, const uint8_t *m, int w, int l, float occ) { static_cast<IRef *>(h)->set_map(m, w, l, occ); }
void ref_patch_map(void *h, const uint8_t *p, int x, int y, int w, int hh) { static_cast<IRef *>(h)->patch_map(p, x, y, w, hh); }
void ref_set_start(void *h, float x, float y) { static_cast<IRef *>(h)->set_start(x, y); }
void ref_set_goal(void *h, float x, float y) { static_cast<IRef *>(h)->set_goal(x, y); }
void ref_reset(void *h) { static_cast<IRef *>(h)->reset(); }
int ref_step(void *h) { return static_cast<IRef *>(h)->step(); }
void ref_times(void *h, float *u, float *p) { static_cast<IRef *>(h)->times(u, p); }
void ref_counters(void *h, uint64_t *u, uint64_t *e) { static_cast<IRef *>(h)->counters(u, e); }
uint64_t ref_map_size(void *h) { return static_cast<IRef *>(h)->map_size(); }
float ref_top_key(void *h, int *empty) { return static_cast<IRef *>(h)->top_key(empty); }
void ref_dump(void *h, float *g, float *rhs) { static_cast<IRef *>(h)->dump(g, rhs); }
int ref_extract(void *h, int use_field, const float *field, int max_steps, int lookahead, int allow_indirect,
                float *xy, int cap, float *cost, float *dist, float *e_ms) {
    return static_cast<IRef *>(h)->extract(use_field, field, max_steps, lookahead, allow_indirect, xy, cap, cost,
                                           dist, e_ms);
}
// INFO of every expanded element as 4 int32 (INT32_MIN = absent); returns 0 when the level keeps no INFO
int ref_dump_info(void *h, int32_t *out) { return static_cast<IRef *>(h)->dump_info(out); }
// cost_ of the last ref_extract (one entry per path piece)
int ref_step_costs(void *h, float *out, int cap) { return static_cast<IRef *>(h)->step_costs(out, cap); }
float ref_op_tri(void *h, int nx, int ny, int ax, int ay, int bx, int by, float ga, float gb) {
    return static_cast<IRef *>(h)->op_tri(nx, ny, ax, ay, bx, by, ga, gb);
}
float ref_op_stencil(void *h, float ga, float gb, float tau, float hh) {
    return static_cast<IRef *>(h)->op_stencil(ga, gb, tau, hh);
}
const uint8_t *ref_map_bytes(void *h) { return static_cast<IRef *>(h)->map_bytes(); }

}  // extern "C"
