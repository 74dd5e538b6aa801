// The planner process of the reference's demos over the GPU planners: same FIFO wire protocol
// (Tests/Planners/FDSTAR/main.cpp:49-175, DFM/main.cpp:44-96, SGDFM/main.cpp:97), so the Python simulator side
// (Simulator/simulator/run_simulator.py:38-113) can drive it unchanged.
//
//   fifo_demo <fd|mfd> <level> <fifo_in> <fifo_out> <from_x> <from_y> <to_x> <to_y> [tof] [--focus] [--device N]
//   fifo_demo dfm      <level> <fifo_in> <fifo_out>                                      [--focus] [--device N]
//   --max-iter N ends the drive after N steps (the MS-DFM walk circles its goal cell without landing on it)
//
// Wire format (little endian, as the reference writes raw host memory):
//   handshake   planner -> 0 ; simulator -> 0, int32 width, int32 height, width*height map bytes, int32 min
//               (DFM variant only: simulator -> float from_x, from_y, to_x, to_y, byte tof)
//   every step  planner -> 1, float x, float y, float step_cost      (DFM variant reports x + 0.5, y + 0.5)
//               simulator -> 1, int32 top, left, height, width, height*width patch bytes, int32 min
//               planner -> 3, int32 n, n * (float x, float y), n-1 float piece costs (cost_, zero padded),
//                          float total_dist, float total_cost, float u_time, float p_time, float e_time
//               planner -> 4, int64 count, count * (int32 x, int32 y, float g, float rhs)         (tof only)
//   end         planner -> 2 ; simulator -> 2
// Differences from the reference process: no CPU pinning / SCHED_FIFO (main.cpp:36-47 — the work is on the GPU);
// the planner is picked by argv instead of by executable; the walk runs on the device (GpuPathExtractor); the tof
// record lists every reached element of the field (g == rhs: the field is converged or, with --focus, settled up
// to the start). No CPU fallback: without a CUDA device the process exits 70.
#include <cinttypes>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fcntl.h>
#include <limits>
#include <memory>
#include <string>
#include <unistd.h>
#include <vector>

#include "GpuPlanners.h"

namespace {

struct Wire {
    int in = -1, out = -1;
    std::vector<char> obuf;
    bool get(void *p, size_t n) {
        char *c = (char *) p;
        while (n) {
            const ssize_t r = read(in, c, n);
            if (r <= 0) return false;
            c += r;
            n -= (size_t) r;
        }
        return true;
    }
    template <typename T> void put(const T &v) { obuf.insert(obuf.end(), (const char *) &v, (const char *) &v + sizeof(T)); }
    bool flush() {
        const char *c = obuf.data();
        size_t n = obuf.size();
        while (n) {
            const ssize_t w = write(out, c, n);
            if (w <= 0) return false;
            c += w;
            n -= (size_t) w;
        }
        obuf.clear();
        return true;
    }
    bool wait_for(char code) {  // the reference spins until the expected byte arrives (main.cpp:59-61)
        char a = -1;
        while (a != code)
            if (!get(&a, 1)) return false;
        return true;
    }
};

}  // namespace

template <typename Planner>
static int serve(Wire &w, bool dfm_variant, bool indirect, float from_x, float from_y, float to_x, float to_y, bool tof,
                 bool focus, int device, int max_iter) {
    using tasi_host::Cell;
    using tasi_host::Position;
    int32_t width = 0, height = 0, top = 0, left = 0, min = 0;
    w.put<char>(0);
    if (!w.flush() || !w.wait_for(0)) return 3;
    if (!w.get(&width, 4) || !w.get(&height, 4) || width <= 0 || height <= 0) return 3;
    std::shared_ptr<uint8_t> data(new uint8_t[(size_t) width * height], std::default_delete<uint8_t[]>());
    if (!w.get(data.get(), (size_t) width * height) || !w.get(&min, 4)) return 3;
    if (dfm_variant) {
        char t = 0;
        if (!w.get(&from_x, 4) || !w.get(&from_y, 4) || !w.get(&to_x, 4) || !w.get(&to_y, 4) || !w.get(&t, 1)) return 3;
        tof = t != 0;
    }
    Position next_point(from_x, from_y), goal(to_x, to_y);
    float next_step_cost = 0;

    Planner planner(device);
    tasi_host::GpuPathExtractor<Planner> extractor(planner);
    extractor.allow_indirect_traversals = indirect;
    planner.reset();
    planner.set_occupancy_threshold(1);
    planner.set_heuristic_multiplier((float) min);
    planner.set_focus(focus);
    planner.set_map(data, width, height);
    planner.set_start(next_point);
    planner.set_goal(goal);

    int steps = 0;
    while (true) {
        w.put<char>(1);
        w.put<float>(next_point.x + (dfm_variant ? 0.5f : 0.0f));
        w.put<float>(next_point.y + (dfm_variant ? 0.5f : 0.0f));
        w.put<float>(next_step_cost);
        if (!w.flush() || !w.wait_for(1)) return 4;
        if (!w.get(&top, 4) || !w.get(&left, 4) || !w.get(&height, 4) || !w.get(&width, 4)) return 4;
        if (width <= 0 || height <= 0) return 4;
        std::shared_ptr<uint8_t> patch(new uint8_t[(size_t) width * height], std::default_delete<uint8_t[]>());
        if (!w.get(patch.get(), (size_t) width * height) || !w.get(&min, 4)) return 4;
        planner.patch_map(patch, top, left, width, height);
        planner.set_heuristic_multiplier((float) min);

        const int rc = planner.step();
        if (rc != LOOP_OK) { fprintf(stderr, "fifo_demo: step %d: %s\n", rc, planner.last_error()); return 5; }
        extractor.extract_path();
        ++steps;

        w.put<char>(3);
        w.put<int32_t>((int32_t) extractor.path_.size());
        for (const auto &pose : extractor.path_) { w.put<float>(pose.x); w.put<float>(pose.y); }
        // the simulator reads exactly n - 1 piece costs (run_simulator.py receive_path); cost_ is shorter when a
        // step added three positions for two pieces (Type I from an opposite edge): pad, so the stream stays aligned
        for (size_t i = 0; i + 1 < extractor.path_.size(); ++i)
            w.put<float>(i < extractor.cost_.size() ? extractor.cost_[i] : 0.0f);
        w.put<float>(extractor.total_dist);
        w.put<float>(extractor.total_cost);
        w.put<float>(planner.u_time);
        w.put<float>(planner.p_time);
        w.put<float>(extractor.e_time);
        if (tof) {
            int ex = 0, ey = 0;
            planner.field_dims(&ex, &ey);
            const std::vector<float> f = planner.read_field();
            int64_t count = 0;
            for (const float v : f) count += v < std::numeric_limits<float>::infinity();
            w.put<char>(4);
            w.put<int64_t>(count);
            for (int x = 0; x < ex; ++x)
                for (int y = 0; y < ey; ++y) {
                    const float v = f[(size_t) x * ey + y];
                    if (!(v < std::numeric_limits<float>::infinity())) continue;
                    w.put<int32_t>(x); w.put<int32_t>(y); w.put<float>(v); w.put<float>(v);
                }
        }
        if (!w.flush()) return 4;

        if (extractor.path_.size() < 2) break;  // nowhere to go (the reference would report the same place for ever)
        const Position prev_point = next_point;
        for (size_t i = 1; i < extractor.path_.size(); ++i) {  // advance at most 5 cells (main.cpp:157-163)
            next_point = Position(extractor.path_[i].x, extractor.path_[i].y);
            if (i - 1 < extractor.cost_.size()) next_step_cost = extractor.cost_[i - 1];
            const Cell a(next_point), b(prev_point);
            if (std::hypot((float) (a.x - b.x), (float) (a.y - b.y)) > 5) break;
        }
        if (next_point.x == goal.x && next_point.y == goal.y) break;
        if (max_iter > 0 && steps >= max_iter) break;
        if (next_point.x == prev_point.x && next_point.y == prev_point.y) break;  // stuck (run_test.py:127-131 kills it)
        planner.set_start(next_point);
    }
    w.put<char>(2);
    if (!w.flush() || !w.wait_for(2)) return 6;
    fprintf(stderr, "fifo_demo: %d steps, last position [%g, %g]\n", steps, next_point.x, next_point.y);
    return 0;
}

int main(int argc, char **argv) {
    std::vector<std::string> a;
    bool focus = false;
    int device = 0, max_iter = 0;
    for (int i = 1; i < argc; ++i) {
        if (!strcmp(argv[i], "--focus")) focus = true;
        else if (!strcmp(argv[i], "--max-iter") && i + 1 < argc) max_iter = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--device") && i + 1 < argc) device = atoi(argv[++i]);
        else a.push_back(argv[i]);
    }
    const bool dfm = !a.empty() && a[0] == "dfm";
    if (a.size() < 4 || (!dfm && a.size() < 8)) {
        fprintf(stderr, "usage: %s <fd|mfd> <level> <fifo_in> <fifo_out> <from_x> <from_y> <to_x> <to_y> [tof] [--focus] "
                        "[--device N]\n       %s dfm <level> <fifo_in> <fifo_out> [--focus] [--device N]\n", argv[0], argv[0]);
        return 64;
    }
    const int level = atoi(a[1].c_str());
    float fx = 0, fy = 0, tx = 0, ty = 0;
    bool tof = false;
    if (!dfm) {
        fx = strtof(a[4].c_str(), nullptr); fy = strtof(a[5].c_str(), nullptr);
        tx = strtof(a[6].c_str(), nullptr); ty = strtof(a[7].c_str(), nullptr);
        tof = a.size() > 8 && atoi(a[8].c_str()) != 0;
    }
    {   // fail before touching the FIFOs when there is no device: this process has no CPU path
        tasi_planner *probe = nullptr;
        const int rc = tasi_create(TASI_FD, 0, device, &probe);
        if (rc != TASI_OK) {
            fprintf(stderr, "fifo_demo: no CUDA device or library (code %d): this path has no CPU fallback\n", rc);
            return 70;
        }
        tasi_destroy(probe);
    }
    Wire w;
    w.in = open(a[2].c_str(), O_RDONLY);   // same order as the reference: input first, then output (main.cpp:52-53)
    w.out = open(a[3].c_str(), O_WRONLY);
    if (w.in < 0 || w.out < 0) { perror("fifo_demo: open"); return 66; }
#define SERVE(P, DV, IND) return serve<P>(w, DV, IND, fx, fy, tx, ty, tof, focus, device, max_iter)
    try {
        if (a[0] == "fd") { if (level == 0) SERVE(GpuFieldDPlanner<0>, false, true); SERVE(GpuFieldDPlanner<1>, false, true); }
        if (a[0] == "dfm") { if (level == 0) SERVE(GpuDFMPlanner<0>, true, true); SERVE(GpuDFMPlanner<1>, true, true); }
        if (a[0] == "mfd") {
            if (level == 0) SERVE(GpuShiftedGridPlanner<0>, false, false);
            if (level == 1) SERVE(GpuShiftedGridPlanner<1>, false, false);
            SERVE(GpuShiftedGridPlanner<2>, false, false);
        }
    } catch (const tasi_host::GpuError &e) {
        fprintf(stderr, "fifo_demo: %s (code %d)\n", e.what(), e.code);
        return 70;
    }
    fprintf(stderr, "unknown planner %s\n", a[0].c_str());
    return 64;
}
