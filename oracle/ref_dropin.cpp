// TEST INFRASTRUCTURE — not product code.
//
// Drop-in check: the UNMODIFIED reference path extractor (PathExtraction/LinearInterpolationPathExtractor.h) and
// the reference's demo loop (Tests/Planners/FDSTAR/main.cpp:77-88, 111-166) run twice on the same map and the
// same sequence of sensor patches — once over the reference's own CPU planner, once over the GPU planner class of
// unige_tasi_path_planners_b200/host/GpuPlanners.h built with -DTASI_WITH_REFERENCE_TYPES (its `grid` is the
// reference's Graph, its `map` the reference's ExpandedMap refilled from the device field after every step).
// Per step it prints one JSON line with both paths' lengths, the largest point difference, both costs.
//
// Compiled by oracle/Makefile against the reference sources where they lie; outputs oracle/_ref/ref_dropin_{fd,dfm,sg}.
// One planner family per translation unit (the reference headers share an include guard).
#include <array>
#include <cinttypes>
#include <functional>
#include <iostream>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <vector>

#if defined(REF_FAMILY_FD)
#include "FieldDPlanner.h"
template <int O> using RefPlannerT = FieldDPlanner<O>;
#elif defined(REF_FAMILY_DFM)
#include "DynamicFastMarching.h"
template <int O> using RefPlannerT = DFMPlanner<O>;
#elif defined(REF_FAMILY_SG)
#include "ShiftedGridPlanner.h"
template <int O> using RefPlannerT = ShiftedGridPlanner<O>;
#else
#error "define REF_FAMILY_FD, REF_FAMILY_DFM or REF_FAMILY_SG"
#endif
#include "LinearInterpolationPathExtractor.h"

#define TASI_WITH_REFERENCE_TYPES
#include "GpuPlanners.h"

#if defined(REF_FAMILY_FD)
template <int O> using GpuPlannerT = GpuFieldDPlanner<O>;
#elif defined(REF_FAMILY_DFM)
template <int O> using GpuPlannerT = GpuDFMPlanner<O>;
#else
template <int O> using GpuPlannerT = GpuShiftedGridPlanner<O>;
#endif

static uint64_t splitmix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

// "true" world and what the rover knows: the known map starts as a blurred guess (every free cell at cost 20,
// no obstacles) and a square sensor patch of the true world is revealed around the rover at every step — the
// role Simulator/simulator/run_simulator.py:9-28 (round_patch_update) plays for the demos.
static void make_world(std::vector<uint8_t> &w, int L, int W, uint64_t seed, int p_obs) {
    w.resize((size_t) L * W);
    for (int x = 0; x < L; ++x)
        for (int y = 0; y < W; ++y) {
            const uint64_t h = splitmix64(seed ^ ((uint64_t) x * (uint64_t) W + (uint64_t) y));
            w[(size_t) x * W + y] = (h % 1000ULL) < (uint64_t) (10 * p_obs) ? 255 : (uint8_t) (1 + ((h >> 16) % 64ULL));
        }
}

template <typename Planner>
struct Drive {
    typedef LinearInterpolationPathExtractor<typename Planner::Map::ElemType, typename Planner::Base::Info> Extractor;
    Planner planner;
    Extractor extractor;
    std::shared_ptr<uint8_t> data;
    Position next_point, goal;
    std::function<void()> on_extract;  // runs right after extractor.extract_path(), before the rover advances
    Drive(const std::vector<uint8_t> &known, int L, int W, Position start, Position goal_, bool indirect)
        : extractor(planner.get_expanded_map(), planner.get_grid()), next_point(start), goal(goal_) {
        data = std::shared_ptr<uint8_t>(new uint8_t[known.size()], std::default_delete<uint8_t[]>());
        std::memcpy(data.get(), known.data(), known.size());
        extractor.allow_indirect_traversals = indirect;
        planner.reset();
        planner.set_occupancy_threshold(1);
        planner.set_heuristic_multiplier(1);
        planner.set_map(data, W, L);
        planner.set_start(next_point);
        planner.set_goal(goal);
    }
    // one iteration of the demo loop: patch, step, extract, advance <= 5 cells (main.cpp:111-166)
    bool iterate(const std::vector<uint8_t> &world, int L, int W, int radius) {
        Cell c(next_point);
        const int top = std::max(0, c.x - radius), left = std::max(0, c.y - radius);
        const int h = std::min(L, c.x + radius + 1) - top, w = std::min(W, c.y + radius + 1) - left;
        std::shared_ptr<uint8_t> patch(new uint8_t[(size_t) h * w], std::default_delete<uint8_t[]>());
        for (int i = 0; i < h; ++i)
            std::memcpy(patch.get() + (size_t) i * w, world.data() + (size_t) (top + i) * W + left, w);
        planner.patch_map(patch, top, left, w, h);
        planner.step();
        extractor.extract_path();
        if (on_extract) on_extract();
        if (extractor.path_.empty()) return false;
        Position prev = next_point;
        for (unsigned i = 1; i < extractor.path_.size(); i++) {
            next_point = {extractor.path_[i].x, extractor.path_[i].y};
            if (Cell(next_point).distance(Cell(prev)) > 5) break;
        }
        if (next_point == goal) return false;
        planner.set_start(next_point);
        return true;
    }
};

template <int O>
static int run(int L, int W, uint64_t seed, int max_iter, int radius, bool focus) {
    std::vector<uint8_t> world, known((size_t) L * W, 20);
    make_world(world, L, W, seed, 10);
    const Position start(5.0f, 5.0f), goal((float) (L - 6), (float) (W - 6));
    for (int x = 0; x < L; ++x)
        for (int y = 0; y < W; ++y)
            if ((std::abs(x - 5) <= 4 && std::abs(y - 5) <= 4) || (std::abs(x - (L - 6)) <= 4 && std::abs(y - (W - 6)) <= 4))
                world[(size_t) x * W + y] = 1;
#if defined(REF_FAMILY_SG)
    const bool indirect = false;  // SGDFM/main.cpp:97
#else
    const bool indirect = true;   // FDSTAR/main.cpp:82, DFM/main.cpp:80
#endif
    Drive<RefPlannerT<O>> ref(known, L, W, start, goal, indirect);
    Drive<GpuPlannerT<O>> gpu(known, L, W, start, goal, indirect);
    // third arm: the device walk (tasi_extract_paths behind GpuPathExtractor) over the same GPU planner, to be
    // compared with the reference extractor's walk over that planner's hand-off window
    gpu.planner.set_focus(focus);  // the reference always stops at the start (end_condition); ours on request
    tasi_host::GpuPathExtractor<GpuPlannerT<O>> native(gpu.planner);
    native.allow_indirect_traversals = indirect;
    gpu.on_extract = [&native]() { native.extract_path(); };
    int bad = 0;
    for (int it = 0; it < max_iter; ++it) {
        const bool more_r = ref.iterate(world, L, W, radius);
        const bool more_g = gpu.iterate(world, L, W, radius);
        const auto &pr = ref.extractor.path_, &pg = gpu.extractor.path_;
        double maxd = 0;
        const size_t n = std::min(pr.size(), pg.size());
        for (size_t i = 0; i < n; ++i)
            maxd = std::max(maxd, (double) std::max(std::fabs(pr[i].x - pg[i].x), std::fabs(pr[i].y - pg[i].y)));
        const bool same = pr.size() == pg.size() && maxd == 0.0 && more_r == more_g;
        const double cr = ref.extractor.total_cost, cg = gpu.extractor.total_cost;
        const bool cost_ok = (cr == cg) || std::fabs(cr - cg) <= 1e-3 * std::fabs(cr);
        bool native_same = native.path_.size() == pg.size() && native.total_cost == gpu.extractor.total_cost &&
                           native.total_dist == gpu.extractor.total_dist && native.cost_ == gpu.extractor.cost_;
        for (size_t i = 0; native_same && i < pg.size(); ++i)
            native_same = native.path_[i].x == pg[i].x && native.path_[i].y == pg[i].y;
        if (!same || !cost_ok || !native_same) ++bad;
        printf("{\"iter\": %d, \"device_walk_identical\": %s, \"device_walk_ms\": %.3f, \"ref_walk_ms\": %.3f,",
               it, native_same ? "true" : "false", native.e_time, ref.extractor.e_time);
        printf(" \"it\": %d, \"ref_points\": %zu, \"gpu_points\": %zu, \"max_point_diff\": %.9g, \"ref_cost\": %.9g, "
               "\"gpu_cost\": %.9g, \"ref_pos\": [%.6g, %.6g], \"gpu_pos\": [%.6g, %.6g], \"identical\": %s, "
               "\"ref_ms\": [%.3f, %.3f], \"gpu_ms\": [%.3f, %.3f], \"ref_expanded\": %lu, \"gpu_updated\": %lu}\n",
               it, pr.size(), pg.size(), maxd, cr, cg, ref.next_point.x, ref.next_point.y, gpu.next_point.x,
               gpu.next_point.y, same ? "true" : "false", ref.planner.u_time, ref.planner.p_time, gpu.planner.u_time,
               gpu.planner.p_time, ref.planner.num_nodes_expanded, gpu.planner.num_nodes_updated);
        if (!more_r || !more_g) break;
    }
    printf("{\"mismatching_steps\": %d}\n", bad);
    return bad == 0 ? 0 : 1;
}

int main(int argc, char **argv) {
    if (argc < 6) {
        fprintf(stderr, "usage: %s <level> <length> <width> <seed> <max_iter> [radius] [focus 0|1]\n", argv[0]);
        return 64;
    }
    const int level = atoi(argv[1]), L = atoi(argv[2]), W = atoi(argv[3]);
    const uint64_t seed = strtoull(argv[4], nullptr, 10);
    const int max_iter = atoi(argv[5]), radius = argc > 6 ? atoi(argv[6]) : 5;
    const bool focus = argc > 7 && atoi(argv[7]) != 0;
    try {
        if (level == 0) return run<0>(L, W, seed, max_iter, radius, focus);
        return run<1>(L, W, seed, max_iter, radius, focus);
    } catch (const std::exception &e) {
        fprintf(stderr, "ref_dropin: %s\n", e.what());
        return 70;
    }
}
