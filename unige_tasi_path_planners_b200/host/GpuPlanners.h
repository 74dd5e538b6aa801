// GpuPlanners.h — the C++ host side of the B200 cost-propagation core.
//
// Three class templates with the public surface of the reference's planners, on top of the C ABI of
// include/tasi_gpu.h (libtasi_gpu.so). There is no CPU fallback: constructing a planner without a CUDA
// device (or without the library) throws.
//
//   GpuFieldDPlanner<O>        stands in for FieldDPlanner<O>        (FieldDStar/FieldDPlanner.h:29-52)
//   GpuDFMPlanner<O>           stands in for DFMPlanner<O>           (DynamicFastMarching/DynamicFastMarching.h:29-52)
//   GpuShiftedGridPlanner<O>   stands in for ShiftedGridPlanner<O>   (ShiftedGridFastMarching/ShiftedGridPlanner.h:29-52)
//
// The mirrored interface is ReplannerBase<Derived,Elem,Info,Key> (ProjectToolkit/include/ReplannerBase.h:29-161):
// reset, step (LOOP_OK / LOOP_FAILURE_NO_GRAPH / LOOP_FAILURE_NO_GOAL), set_occupancy_threshold,
// set_heuristic_multiplier, set_map, patch_map(patch, x, y, w, h), set_start, set_goal, get_expanded_map,
// get_grid, and the public data u_time, p_time, num_nodes_updated, num_nodes_expanded, heuristic_multiplier,
// the five state flags, grid and map.
//
//   tasi_host::GpuPathExtractor<Planner>   stands in for LinearInterpolationPathExtractor<E,T>
//                                          (PathExtraction/LinearInterpolationPathExtractor.h:9-22), walk on the device
// Beyond the reference: set_focus (its end_condition as a switch), read_parents (its level-1 INFO on demand), stats.
//
// Two build modes:
//   * default (stand-alone): Position / Node / Cell / HostGrid / FieldMap below are this repo's own small types
//     with the reference's member names; FieldMap is a dense view of the device field that fetches windows on
//     demand (get_g / get_rhs / get_interp_rhs follow ExpandedMap_impl.h:65-97).
//   * -DTASI_WITH_REFERENCE_TYPES (with the reference's include paths): grid is the reference's own Graph and map
//     its own ExpandedMap<Elem,void>, refilled after every step() from a window of the device field around the
//     start (ExpandedMap::insert_or_assign, ExpandedMap_impl.h:16-28), so the UNMODIFIED
//     LinearInterpolationPathExtractor<Elem,void> and the demo mains (Tests/Planners/*/main.cpp) compile against
//     these classes after changing one #include and one type name.
#ifndef TASI_GPU_PLANNERS_H
#define TASI_GPU_PLANNERS_H

#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <memory>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

#include "tasi_gpu.h"

#ifdef TASI_WITH_REFERENCE_TYPES
#include "Graph.h"
#include "ExpandedMap.h"
#endif

#ifndef LOOP_OK
#define LOOP_OK 0
#define LOOP_FAILURE_NO_GRAPH -1
#define LOOP_FAILURE_NO_GOAL -2
#endif

namespace tasi_host {

// A CUDA / argument failure reported by the C ABI from a member the reference declares void.
class GpuError : public std::runtime_error {
 public:
    GpuError(int code_, const std::string &what_) : std::runtime_error(what_), code(code_) {}
    int code;
};

#ifdef TASI_WITH_REFERENCE_TYPES
using ::Cell;
using ::Node;
using ::Position;
typedef ::Graph HostGrid;
#else
// ---- stand-alone geometry types (member names as in ProjectToolkit/include/{Position,Node,Cell}.h) -------------
struct Position {
    float x = 0, y = 0;
    Position() = default;
    Position(float x_, float y_) : x(x_), y(y_) {}
    bool operator==(const Position &o) const { return x == o.x && y == o.y; }
    bool operator!=(const Position &o) const { return !(*this == o); }
};

struct Node {  // grid vertex; Node(Position) rounds (Node.cpp:14-17)
    int x = 0, y = 0;
    Node() = default;
    Node(int x_, int y_) : x(x_), y(y_) {}
    explicit Node(const Position &p) : x((int) roundf(p.x)), y((int) roundf(p.y)) {}
    bool operator==(const Node &o) const { return x == o.x && y == o.y; }
    bool operator!=(const Node &o) const { return !(*this == o); }
};

struct Cell {  // grid cell; Cell(Position) rounds as well (Cell.cpp:20-21)
    int x = 0, y = 0;
    Cell() = default;
    Cell(int x_, int y_) : x(x_), y(y_) {}
    explicit Cell(const Position &p) : x((int) roundf(p.x)), y((int) roundf(p.y)) {}
    bool operator==(const Cell &o) const { return x == o.x && y == o.y; }
    bool operator!=(const Cell &o) const { return !(*this == o); }
    // the four corner nodes (Cell.cpp:48-60)
    std::vector<Node> corners() const { return {Node(x, y), Node(x + 1, y), Node(x, y + 1), Node(x + 1, y + 1)}; }
};

// Host mirror of the cost grid: the members of Graph the hot path and the extractor read
// (ProjectToolkit/include/Graph.h:20-66, Graph.cpp:6-63, 262-268).
class HostGrid {
 public:
    std::shared_ptr<uint8_t> map_;
    std::vector<Cell> updated_cells_;
    Cell start_cell_, goal_cell_;
    Node start_node_, goal_node_;
    Position start_pos_, goal_pos_;
    int length_ = 0, width_ = 0;
    float flength_ = 0, fwidth_ = 0;
    int size_ = 0;
    int occupancy_threshold_uchar_ = 254;

    void set_start(const Position &s) { start_pos_ = s; start_cell_ = Cell(s); start_node_ = Node(s); }
    void set_goal(const Position &g) { goal_pos_ = g; goal_cell_ = Cell(g); goal_node_ = Node(g); }
    void set_occupancy_threshold(float t) { occupancy_threshold_uchar_ = (int) (t * 255.0f); }
    void init(std::shared_ptr<uint8_t> image, int width, int length) {
        length_ = length; width_ = width; flength_ = (float) length; fwidth_ = (float) width;
        size_ = length * width;
        map_ = std::move(image);  // shared with the caller, as the reference does
    }
    // rows x..x+h-1, columns y..y+w-1 take the patch; the changed cells are listed
    void update(const std::shared_ptr<uint8_t> &patch, int x, int y, int w, int h) {
        updated_cells_.clear();
        for (int i = 0; i < h; ++i)
            for (int j = 0; j < w; ++j) {
                uint8_t &mv = map_.get()[(size_t) (x + i) * width_ + (y + j)];
                const uint8_t pv = patch.get()[(size_t) i * w + j];
                if (mv != pv) updated_cells_.emplace_back(x + i, y + j);
                mv = pv;
            }
    }
    bool is_valid(const Node &s) const { return s.x <= length_ && s.y <= width_ && s.x >= 0 && s.y >= 0; }
    bool is_valid(const Cell &c) const { return c.x >= 0 && c.x < length_ && c.y >= 0 && c.y < width_; }
    bool is_valid(const Position &p) const { return p.x >= 0.0f && p.x <= flength_ && p.y >= 0.0f && p.y <= fwidth_; }
    float get_cost(const Cell &c) const {
        if (!is_valid(c)) return std::numeric_limits<float>::infinity();
        const uint8_t b = map_.get()[(size_t) c.x * width_ + c.y];
        return b >= occupancy_threshold_uchar_ ? std::numeric_limits<float>::infinity() : (float) b;
    }
};
#endif  // TASI_WITH_REFERENCE_TYPES

// ---- dense view of the device field: what stands where `const ExpandedMap<E,T>&` stands ------------------------
// Elements are fetched in 64x64 blocks on first touch (tasi_read_window) and dropped at the next step().
template <typename ElemT>
class FieldMap {
 public:
    typedef ElemT ElemType;
    typedef void InfoType;
    static constexpr int BLK = 64;

    void bind(tasi_planner *h, bool cell_based) { h_ = h; cell_based_ = cell_based; }
    void clear() noexcept { blocks_.clear(); }
    size_t size() const {  // elements held (the reference counts expanded elements; here the field is dense)
        int ex = 0, ey = 0;
        return (h_ && tasi_field_dims(h_, &ex, &ey) == TASI_OK) ? (size_t) ex * (size_t) ey : 0;
    }
    float get_g(const ElemT &s) const { return at(s.x, s.y); }
    float get_rhs(const ElemT &s) const { return at(s.x, s.y); }  // converged: g == rhs
    std::pair<float, float> get_g_rhs(const ElemT &s) const { const float v = at(s.x, s.y); return {v, v}; }
    bool consistent(const ElemT &) const { return true; }
    // node value read by the extractor: the node itself, or the mean of the 4 cells around it for a cell map
    float get_interp_rhs(const Node &s) const {
        if (!cell_based_) return at(s.x, s.y);
        return (at(s.x - 1, s.y - 1) + at(s.x - 1, s.y) + at(s.x, s.y - 1) + at(s.x, s.y)) / 4.0f;
    }
    // rows [x0,x0+nx) x columns [y0,y0+ny) into out (row-major), INF outside the grid
    void window(int x0, int y0, int nx, int ny, float *out) const {
        const int rc = tasi_read_window(h_, x0, y0, nx, ny, out);
        if (rc != TASI_OK) throw GpuError(rc, std::string("tasi_read_window: ") + tasi_last_error(h_));
    }

 private:
    float at(int x, int y) const {
        if (x < 0 || y < 0) return std::numeric_limits<float>::infinity();
        const int bx = x / BLK, by = y / BLK;
        const uint64_t key = ((uint64_t) (uint32_t) bx << 32) | (uint32_t) by;
        auto it = blocks_.find(key);
        if (it == blocks_.end()) {
            std::vector<float> blk((size_t) BLK * BLK);
            window(bx * BLK, by * BLK, BLK, BLK, blk.data());
            it = blocks_.emplace(key, std::move(blk)).first;
        }
        return it->second[(size_t) (x - bx * BLK) * BLK + (y - by * BLK)];
    }
    tasi_planner *h_ = nullptr;
    bool cell_based_ = false;
    mutable std::unordered_map<uint64_t, std::vector<float>> blocks_;
};

// ---- the planner surface ---------------------------------------------------------------------------------------
template <int KIND, typename ElemT, int LEVEL>
class GpuReplanner {
 public:
    typedef float Key;
    typedef ElemT Elem;
    typedef void Info;
#ifdef TASI_WITH_REFERENCE_TYPES
    typedef ::ExpandedMap<ElemT, void> Map;
#else
    typedef FieldMap<ElemT> Map;
#endif
    // the demos name the extractor's second template argument `Planner::Base::Info` (FDSTAR/main.cpp:78-81)
    struct Base { typedef void Info; typedef ElemT Elem; typedef float Key; };

    float u_time = 0, p_time = 0;

    explicit GpuReplanner(int device = 0) {
        const int rc = tasi_create(KIND, LEVEL, device, &h_);
        if (rc != TASI_OK)
            throw GpuError(rc, "tasi_create failed: no CUDA device or library (this path has no CPU fallback)");
#ifndef TASI_WITH_REFERENCE_TYPES
        map.bind(h_, KIND == TASI_DFM);
#endif
    }
    // One grid over several GPUs (tasi_create_sharded): same members, same results; the element rows are cut into bands
    // dealt out cyclically over `devices`, each GPU runs one persistent solver over its bands and the boundary rows
    // travel through peer-mapped memory during the tile visits (include/tasi_gpu.h, csrc/tasi_shard.cuh).
    explicit GpuReplanner(const std::vector<int> &devices) {
        const int rc = devices.size() == 1 ? tasi_create(KIND, LEVEL, devices[0], &h_)
                                           : tasi_create_sharded(KIND, LEVEL, devices.data(), (int) devices.size(), &h_);
        if (rc != TASI_OK)
            throw GpuError(rc, std::string(devices.size() == 1 ? "tasi_create" : "tasi_create_sharded") + " failed: " +
                                   tasi_last_error(nullptr) + " (no CUDA device or library? this path has no CPU fallback)");
#ifndef TASI_WITH_REFERENCE_TYPES
        map.bind(h_, KIND == TASI_DFM);
#endif
    }
    ~GpuReplanner() { tasi_destroy(h_); }
    GpuReplanner(const GpuReplanner &) = delete;
    GpuReplanner &operator=(const GpuReplanner &) = delete;

    void reset() {  // ReplannerBase.h:39-41
        initialize_search = true;
        check(tasi_reset(h_), "tasi_reset");
    }

    int step() {  // ReplannerBase.h:43-75; the flag logic itself runs inside tasi_step
        uint64_t nu = 0, ne = 0;
        const int rc = tasi_step(h_, &u_time, &p_time, &nu, &ne);
        if (rc != TASI_OK) return rc;  // -1 / -2 as the reference; < -2 = device error (see last_error())
        num_nodes_updated = (unsigned long) nu;
        num_nodes_expanded = (unsigned long) ne;
        if (initialize_search || new_goal) grid.updated_cells_.clear();  // ReplannerBase.h:52
        else if (new_start) new_start = false;
        new_goal = initialize_search = false;
        hand_off();
        return LOOP_OK;
    }

    void set_occupancy_threshold(float threshold) {
        grid.set_occupancy_threshold(threshold);
        check(tasi_set_occupancy_threshold(h_, threshold), "tasi_set_occupancy_threshold");
    }

    void set_heuristic_multiplier(float mult) {
        heuristic_multiplier = mult;
        check(tasi_set_heuristic_multiplier(h_, mult), "tasi_set_heuristic_multiplier");
    }

    // w = columns, h = rows (ReplannerBase.h:85-88); the host buffer stays shared with the caller and is what
    // get_grid().get_cost() reads, the device gets its own copy during the call
    void set_map(const std::shared_ptr<uint8_t> &new_map, int w, int h) {
        grid.init(new_map, w, h);
        check(tasi_set_map(h_, new_map.get(), w, h), "tasi_set_map");
        initialize_graph = false;
    }

    // x = top row, y = left column, w = columns, h = rows (ReplannerBase.h:90-92, Graph.cpp:36-51)
    void patch_map(const std::shared_ptr<uint8_t> &patch, int x, int y, int w, int h) {
        check(tasi_patch_map(h_, patch.get(), x, y, w, h, nullptr), "tasi_patch_map");
        grid.update(patch, x, y, w, h);
    }

    void set_start(const Position &pos) {
        grid.set_start(pos);
        new_start = true;
        check(tasi_set_start(h_, pos.x, pos.y), "tasi_set_start");
    }

    void set_goal(const Position &point) {  // ReplannerBase.h:99-108
        new_goal = goal_set ? (ElemT(point) != goal_elem_) : true;
        goal_elem_ = ElemT(point);
        grid.set_goal(point);
        goal_set = true;
        check(tasi_set_goal(h_, point.x, point.y), "tasi_set_goal");
    }

    const Map &get_expanded_map() { return map; }
    const HostGrid &get_grid() { return grid; }

    // the reference's plan() always stops once the start is settled (end_condition); here it is a switch, off by
    // default (the whole field converges). slack < 0 = default
    void set_focus(bool enabled, float slack = -1.0f) { check(tasi_set_focus(h_, enabled ? 1 : 0, slack), "tasi_set_focus"); }

    // ---- beyond the reference: device-side evidence and bulk access ------------------------------------------
    tasi_stats stats() const { tasi_stats s; tasi_get_stats(h_, &s); return s; }
    const char *last_error() const { return tasi_last_error(h_); }
    tasi_planner *handle() const { return h_; }
    void field_dims(int *ex, int *ey) const { check(tasi_field_dims(h_, ex, ey), "tasi_field_dims"); }
    std::vector<float> read_field() const {
        int ex = 0, ey = 0;
        field_dims(&ex, &ey);
        std::vector<float> f((size_t) ex * ey);
        check(tasi_read_field(h_, f.data()), "tasi_read_field");
        return f;
    }
    // INFO of the reference's level-1 planners, on demand: 4 signed bytes per element of the window, the offsets of
    // its two parents ((0, 0) = absent); FD*/MFD*: b and ccw(b), MS-DFM: the lower and the higher stencil cell
    std::vector<signed char> read_parents(int x0, int y0, int nx, int ny) const {
        std::vector<signed char> out((size_t) nx * ny * 4);
        check(tasi_read_parents(h_, x0, y0, nx, ny, out.data()), "tasi_read_parents");
        return out;
    }
    // half-side of the window exported to the reference's ExpandedMap after a step (reference-types mode);
    // the extractor walks at most max_steps cells from the start (default 20), <= 0 exports the whole field
    int handoff_radius = 48;

 public:
    unsigned long num_nodes_updated = 0;
    unsigned long num_nodes_expanded = 0;
    float heuristic_multiplier = 1;
    bool initialize_graph = true;
    bool initialize_search = true;
    bool goal_set = false;
    bool new_goal = false;
    bool new_start = false;

    HostGrid grid;
    Map map;

 private:
    void check(int rc, const char *what) const {
        if (rc != TASI_OK) throw GpuError(rc, std::string(what) + ": " + tasi_last_error(h_));
    }

    void hand_off() {
#ifdef TASI_WITH_REFERENCE_TYPES
        // what ReplannerBase::step does to the map on init (ReplannerBase.h:53-54), then the settled values
        map.clear();
        map.init(grid.width_, grid.length_, 8);
        int ex = 0, ey = 0;
        field_dims(&ex, &ey);
        int x0 = 0, y0 = 0, nx = ex, ny = ey;
        if (handoff_radius > 0) {
            const ElemT c(grid.start_pos_);
            x0 = c.x - handoff_radius < 0 ? 0 : c.x - handoff_radius;
            y0 = c.y - handoff_radius < 0 ? 0 : c.y - handoff_radius;
            nx = (c.x + handoff_radius + 1 > ex ? ex : c.x + handoff_radius + 1) - x0;
            ny = (c.y + handoff_radius + 1 > ey ? ey : c.y + handoff_radius + 1) - y0;
        }
        if (nx <= 0 || ny <= 0) return;
        std::vector<float> w((size_t) nx * ny);
        check(tasi_read_window(h_, x0, y0, nx, ny, w.data()), "tasi_read_window");
        for (int i = 0; i < nx; ++i)
            for (int j = 0; j < ny; ++j) {
                const float v = w[(size_t) i * ny + j];
                if (v < std::numeric_limits<float>::infinity()) map.insert_or_assign(ElemT(x0 + i, y0 + j), v, v);
            }
#else
        map.clear();  // cached blocks belong to the previous field
#endif
    }

    tasi_planner *h_ = nullptr;
    ElemT goal_elem_;
};

// ---- the extractor surface -------------------------------------------------------------------------------------
// LinearInterpolationPathExtractor (PathExtraction/LinearInterpolationPathExtractor.h:9-22) with the walk on the
// device: same public members; bound to the planner instead of to its (map, grid) pair, because the field never
// leaves the GPU. One extract_path() = one launch of the warp-per-walk kernel (tasi_extract_paths) and one D2H of
// the positions; the hand-off window of reference-types mode is not read on this route.
template <typename Planner>
class GpuPathExtractor {
 public:
    explicit GpuPathExtractor(Planner &planner) : planner_(planner) {}

    void extract_path() {
        const int cap = 1 + 3 * max_steps, capc = 2 * max_steps;
        std::vector<float> xy((size_t) cap * 2), sc((size_t) capc);
        int np = 0, nc = 0;
        float totals[2] = {0, 0};
        const int rc = tasi_extract_paths(planner_.handle(), nullptr, nullptr, nullptr, 1, max_steps, lookahead ? 1 : 0,
                                          allow_indirect_traversals ? 1 : 0, xy.data(), cap, &np, sc.data(), capc, &nc,
                                          totals, &status, &e_time);
        if (rc != TASI_OK) throw GpuError(rc, std::string("tasi_extract_paths: ") + planner_.last_error());
        path_.clear();
        cost_.assign(sc.begin(), sc.begin() + nc);
        for (int i = 0; i < np; ++i) path_.push_back(Position(xy[2 * i], xy[2 * i + 1]));
        total_cost = totals[0];
        total_dist = totals[1];
    }

    std::vector<Position> path_{};
    std::vector<float> cost_{};
    float total_cost = 0;
    float total_dist = 0;
    bool lookahead = true;
    int max_steps = 20;
    float e_time = 0;
    bool allow_indirect_traversals = true;
    int status = TASI_WALK_GOAL;  // extra: why the walk ended (TASI_WALK_*)

 private:
    Planner &planner_;
};

}  // namespace tasi_host

template <int OptimizationLevel>
class GpuFieldDPlanner : public tasi_host::GpuReplanner<TASI_FD, tasi_host::Node, OptimizationLevel> {
    static_assert(OptimizationLevel == 0 || OptimizationLevel == 1, "FD* has optimisation levels 0 and 1");
 public:
    using tasi_host::GpuReplanner<TASI_FD, tasi_host::Node, OptimizationLevel>::GpuReplanner;
};

template <int OptimizationLevel>
class GpuDFMPlanner : public tasi_host::GpuReplanner<TASI_DFM, tasi_host::Cell, OptimizationLevel> {
    static_assert(OptimizationLevel == 0 || OptimizationLevel == 1, "MS-DFM has optimisation levels 0 and 1");
 public:
    using tasi_host::GpuReplanner<TASI_DFM, tasi_host::Cell, OptimizationLevel>::GpuReplanner;
};

template <int OptimizationLevel>
class GpuShiftedGridPlanner : public tasi_host::GpuReplanner<TASI_MFD, tasi_host::Node, OptimizationLevel> {
    static_assert(OptimizationLevel >= 0 && OptimizationLevel <= 2, "MFD* has optimisation levels 0, 1 and 2");
 public:
    using tasi_host::GpuReplanner<TASI_MFD, tasi_host::Node, OptimizationLevel>::GpuReplanner;
};

#endif  // TASI_GPU_PLANNERS_H
