// plan_demo — drives the C++ planner classes of GpuPlanners.h the way the reference's demo mains drive theirs
// (Tests/Planners/FDSTAR/main.cpp:77-88, 111-115): reset, set_occupancy_threshold, set_heuristic_multiplier,
// set_map, set_start, set_goal, step; then patch_map + set_start + step for a replan.
//
//   plan_demo <fd|dfm|mfd> <level> <length> <width> <seed> [--patch x y h w pseed] [--dump FILE] [--device N]
//             [--devices 0,1,...]     the same planner object over several GPUs (the sharded planner of the C ABI)
//
// The map is the deterministic synthetic grid of SURVEY.md 8(d) (same generator as synth.py and the CUDA
// synth kernel). Prints one JSON line per step with the counters and a checksum of the field; --dump writes
// the converged field (float32, row-major) so a test can compare it element by element.
#include <cinttypes>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "GpuPlanners.h"

namespace {

uint64_t splitmix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ULL;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

void synth(uint8_t *m, uint64_t seed, int length, int width, int p_obs) {
    for (int x = 0; x < length; ++x)
        for (int y = 0; y < width; ++y) {
            const uint64_t h = splitmix64(seed ^ ((uint64_t) x * (uint64_t) width + (uint64_t) y));
            m[(size_t) x * width + y] = (h % 1000ULL) < (uint64_t) (10 * p_obs) ? 255 : (uint8_t) (1 + ((h >> 16) % 64ULL));
        }
}

void clear_block(uint8_t *m, int length, int width, int cx, int cy) {
    for (int x = cx - 4; x <= cx + 4; ++x)
        for (int y = cy - 4; y <= cy + 4; ++y)
            if (x >= 0 && x < length && y >= 0 && y < width) m[(size_t) x * width + y] = 1;
}

template <typename Planner>
void report(Planner &pl, const char *what, const std::string &dump) {
    std::vector<float> f = pl.read_field();
    double sum = 0;
    uint64_t fin = 0;
    for (float v : f)
        if (v < std::numeric_limits<float>::infinity()) { sum += v; ++fin; }
    const tasi_stats st = pl.stats();
    const auto &grid = pl.get_grid();
    // the demo loop's extractor.extract_path() (FDSTAR/main.cpp:116), on the device
    tasi_host::GpuPathExtractor<Planner> ex(pl);
    ex.extract_path();
    typename Planner::Elem se(grid.start_pos_);
    printf("{\"step\": \"%s\", \"u_time\": %.4f, \"p_time\": %.4f, \"num_nodes_updated\": %lu, \"num_nodes_expanded\": %lu, "
           "\"finite\": %" PRIu64 ", \"sum\": %.9g, \"start_value\": %.9g, \"start_interp_rhs\": %.9g, "
           "\"tile_visits\": %" PRIu64 ", \"kernel_ms_solve\": %.4f, \"kernel_ms_inval\": %.4f, "
           "\"path_points\": %zu, \"path_cost\": %.9g, \"path_dist\": %.9g, \"path_status\": %d, \"e_time\": %.4f}\n",
           what, pl.u_time, pl.p_time, pl.num_nodes_updated, pl.num_nodes_expanded, fin, sum,
           pl.get_expanded_map().get_rhs(se),
           pl.get_expanded_map().get_interp_rhs(tasi_host::Node(grid.start_pos_)), st.tile_visits,
           st.kernel_ms_solve, st.kernel_ms_inval, ex.path_.size(), ex.total_cost, ex.total_dist, ex.status, ex.e_time);
    if (!dump.empty()) {
        FILE *fp = fopen((dump + "." + what).c_str(), "wb");
        if (fp) { fwrite(f.data(), sizeof(float), f.size(), fp); fclose(fp); }
        fp = fopen((dump + "." + what + ".path").c_str(), "wb");
        if (fp) {
            for (const auto &q : ex.path_) { fwrite(&q.x, sizeof(float), 1, fp); fwrite(&q.y, sizeof(float), 1, fp); }
            fclose(fp);
        }
    }
}

template <typename Planner>
int run(int length, int width, uint64_t seed, bool patch, int px, int py, int ph, int pw, uint64_t pseed,
        const std::string &dump, const std::vector<int> &devices) {
    std::shared_ptr<uint8_t> data(new uint8_t[(size_t) length * width], std::default_delete<uint8_t[]>());
    synth(data.get(), seed, length, width, 10);
    const tasi_host::Position start(5.0f, 5.0f), goal((float) (length - 6), (float) (width - 6));
    clear_block(data.get(), length, width, 5, 5);
    clear_block(data.get(), length, width, length - 6, width - 6);

    Planner planner(devices);
    if (planner.step() != LOOP_FAILURE_NO_GRAPH) return 2;  // ReplannerBase.h:44
    planner.reset();
    planner.set_occupancy_threshold(1);
    planner.set_heuristic_multiplier(1);
    planner.set_map(data, width, length);
    if (planner.step() != LOOP_FAILURE_NO_GOAL) return 3;  // ReplannerBase.h:45
    planner.set_start(start);
    planner.set_goal(goal);
    int rc = planner.step();
    if (rc != LOOP_OK) { fprintf(stderr, "step: %d %s\n", rc, planner.last_error()); return 4; }
    report(planner, "plan", dump);
    if (patch) {
        std::shared_ptr<uint8_t> pb(new uint8_t[(size_t) ph * pw], std::default_delete<uint8_t[]>());
        synth(pb.get(), pseed, ph, pw, 15);
        planner.patch_map(pb, px, py, pw, ph);  // (patch, top row, left column, columns, rows)
        planner.set_start(tasi_host::Position(8.0f, 8.0f));
        rc = planner.step();
        if (rc != LOOP_OK) { fprintf(stderr, "replan: %d %s\n", rc, planner.last_error()); return 5; }
        report(planner, "replan", dump);
        // the host buffer is shared and patched in place, as Graph::update does (Graph.cpp:45-48)
        if (data.get()[(size_t) px * width + py] != pb.get()[0]) return 6;
    }
    return 0;
}

}  // namespace

int main(int argc, char **argv) {
    if (argc < 6) {
        fprintf(stderr, "usage: %s <fd|dfm|mfd> <level> <length> <width> <seed> [--patch x y h w pseed] [--dump FILE] "
                        "[--device N | --devices 0,1,...]\n", argv[0]);
        return 64;
    }
    const std::string kind = argv[1];
    const int level = atoi(argv[2]), length = atoi(argv[3]), width = atoi(argv[4]);
    const uint64_t seed = strtoull(argv[5], nullptr, 10);
    bool patch = false;
    int px = 0, py = 0, ph = 0, pw = 0, device = 0;
    uint64_t pseed = 0;
    std::string dump;
    std::vector<int> devices;
    for (int i = 6; i < argc; ++i) {
        if (!strcmp(argv[i], "--patch") && i + 5 < argc) {
            patch = true;
            px = atoi(argv[i + 1]); py = atoi(argv[i + 2]); ph = atoi(argv[i + 3]); pw = atoi(argv[i + 4]);
            pseed = strtoull(argv[i + 5], nullptr, 10);
            i += 5;
        } else if (!strcmp(argv[i], "--dump") && i + 1 < argc) dump = argv[++i];
        else if (!strcmp(argv[i], "--device") && i + 1 < argc) device = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--devices") && i + 1 < argc) {
            for (const char *c = argv[++i]; *c;) {
                devices.push_back((int) strtol(c, (char **) &c, 10));
                if (*c == ',') ++c;
            }
        }
    }
    if (devices.empty()) devices.push_back(device);
    try {
#define RUN(P) return run<P>(length, width, seed, patch, px, py, ph, pw, pseed, dump, devices)
        if (kind == "fd") { if (level == 0) RUN(GpuFieldDPlanner<0>); else RUN(GpuFieldDPlanner<1>); }
        if (kind == "dfm") { if (level == 0) RUN(GpuDFMPlanner<0>); else RUN(GpuDFMPlanner<1>); }
        if (kind == "mfd") {
            if (level == 0) RUN(GpuShiftedGridPlanner<0>);
            else if (level == 1) RUN(GpuShiftedGridPlanner<1>);
            else RUN(GpuShiftedGridPlanner<2>);
        }
#undef RUN
    } catch (const tasi_host::GpuError &e) {
        fprintf(stderr, "plan_demo: %s (code %d)\n", e.what(), e.code);
        return 70;
    }
    fprintf(stderr, "unknown planner %s\n", kind.c_str());
    return 64;
}
