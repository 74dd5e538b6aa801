// Path extraction over the converged field, as __host__ __device__ functions.
//
// Restates what LinearInterpolationPathExtractor does with one position of the path
// (PathExtraction/LinearInterpolationPathExtractor_impl.h:60-213): every edge of the ring of nodes around the
// position is a candidate to cross; the interpolated traversal model (ProjectToolkit/InterpolatedTraversal.cpp:
// Type I :6-124, II :125-235, III :236-322, A :324-401, B :403-478, selection :482-778) gives the cost to the goal
// through that edge and the points the path gains. Float32 in the reference's operation order, no FMA
// contraction (nvcc -fmad=false, gcc -ffp-contract=off); hypotf as glibc computes it (in double).
//
// One candidate edge = one call of hop_over_edge(); the device kernel at the end of this file gives each lane of a
// warp one candidate, the host build (tasi_hostops.cpp, tests only) loops.
#pragma once
#include "tasi_update.cuh"

#ifdef __CUDACC__
#define TASI_XD __host__ __device__
#else
#define TASI_XD
#endif

namespace tasi {

// What the extractor reads: the field, the cost map and the map bounds (Graph::is_valid, Graph.cpp:53-63).
struct FieldView {
    const float *g;        // element (0,0)
    int64_t gpitch;
    const uint8_t *cost;   // cell (0,0)
    int64_t cpitch;
    int L, W;              // cell rows / columns
    int thr;               // Graph.cpp:262-268
    int cells;             // 1: the field lives on cells (MS-DFM), node values are 4-cell means

    TASI_XD bool node_ok(int x, int y) const { return x >= 0 && y >= 0 && x <= L && y <= W; }
    TASI_XD float cell_cost(int x, int y) const {
        if (x < 0 || y < 0 || x >= L || y >= W) return inf_f();
        const int v = cost[(int64_t) x * cpitch + y];
        return v >= thr ? inf_f() : (float) v;
    }
    TASI_XD float cell_val(int x, int y) const {
        if (x < 0 || y < 0 || x >= L || y >= W) return inf_f();
        return g[(int64_t) x * gpitch + y];
    }
    // ExpandedMap::get_interp_rhs (ExpandedMap_impl.h:87-97); node (x,y) is known to be valid
    TASI_XD float node_val(int x, int y) const {
        if (!cells) return g[(int64_t) x * gpitch + y];
        const float a = cell_val(x, y - 1), b = cell_val(x - 1, y - 1), c = cell_val(x, y), d = cell_val(x - 1, y);
        return (a + b + c + d) * 0.25f;
    }
};

// hypotf as glibc evaluates it (sysdeps/ieee754/flt-32/e_hypotf.c: double products, one double sqrt, one narrowing)
TASI_XD inline float hyp(float a, float b) {
    const double x = (double) a, y = (double) b;
    return (float) sqrt(x * x + y * y);
}
TASI_XD inline float cathf(float hy, float leg) { return sqrtf(hy * hy - leg * leg); }  // Macros.h CATH

// What crossing one edge adds to the path: up to 3 points, up to 2 piece costs, the cost to the goal from here.
struct Hop {
    int n;         // points gained; 0 = the edge cannot be crossed
    float x[3], y[3];
    int nc;
    float sc[2];
    float to_goal;
    float step;    // sum of sc, accumulated from 0.0f as std::accumulate does
};

// One triangle (p0, p1, p2) of the traversal model: p0 on the line of p1, p2 the far end of the edge p1-p2;
// c = cost of the cell the triangle lies in, b = cost of the cell on the other side of the line p0-p1.
struct Tri {
    float p0x, p0y;
    int p1x, p1y, p2x, p2y;
    float b, c, g1, g2, f;
    float p, q;   // offset of the position along its own edge (Opposite: p, Contiguous: q)
};

// LinearInterpolationPathExtractor::fill_traversal_costs (_impl.h:221-235)
TASI_XD inline void tri_costs(const FieldView &v, Tri &t) {
    if (t.p0x == (float) t.p1x) {
        const int cy = t.p1y - (t.p0y > (float) t.p1y ? 0 : 1);
        t.b = v.cell_cost(t.p1x - (t.p2x > t.p1x ? 1 : 0), cy);
        t.c = v.cell_cost(t.p1x - (t.p2x < t.p1x ? 1 : 0), cy);
    } else {
        const int cx = t.p1x - (t.p0x < (float) t.p1x ? 1 : 0);
        t.b = v.cell_cost(cx, t.p1y - (t.p2y < t.p1y ? 0 : 1));
        t.c = v.cell_cost(cx, t.p1y - (t.p2y > t.p1y ? 0 : 1));
    }
}

// a point on the line p0 -> p1 at offset d from p0 (Macros.h INTERP_1 / INTERP_ABS with a float origin)
TASI_XD inline void on_line(const Tri &t, float d, bool unit, float &x, float &y) {
    if (t.p0x == (float) t.p1x) {
        const float span = (float) t.p1y - t.p0y;
        x = t.p0x;
        y = unit ? t.p0y + span * d : t.p0y + span / fabsf(span) * d;
    } else {
        const float span = (float) t.p1x - t.p0x;
        x = unit ? t.p0x + span * d : t.p0x + span / fabsf(span) * d;
        y = t.p0y;
    }
}
// a point on the edge p1 -> p2 at offset d from p1 (integer ends: the span is +-1 either way)
TASI_XD inline void on_edge(const Tri &t, float d, float &x, float &y) {
    if (t.p0x == (float) t.p1x) {
        x = (float) t.p1x + (float) (t.p2x - t.p1x) * d;
        y = (float) t.p1y;
    } else {
        x = (float) t.p1x;
        y = (float) t.p1y + (float) (t.p2y - t.p1y) * d;
    }
}


enum { SIT_CORNER = 0, SIT_CONTIG = 1, SIT_OPPOSITE = 2 };
enum { TY_I = 0, TY_II = 1, TY_III = 2, TY_A = 3, TY_B = 4 };

// Is a traversal type admissible from this situation (the cond() of each family)? The corner situation has its
// own decision tree (pick_corner) and never asks.
TASI_XD inline bool type_ok(int ty, int sit, const Tri &t) {
    switch (ty) {
        case TY_I:
            return sit == SIT_CONTIG ? t.c > t.b * hyp(1.0f, 1.0f / (1.0f - t.q)) : t.c > t.b * hyp(1.0f, 1.0f + t.p);
        case TY_II:
            if (!(t.f > 0.0f)) return false;
            return sit == SIT_CONTIG ? t.c > t.f * hyp(1.0f, 1.0f - t.q) : t.c > t.f * hyp(1.0f, 1.0f / (1.0f - t.p));
        case TY_III:
            return sit == SIT_CONTIG ? t.c > t.b : t.c > t.b * hyp(1.0f, t.p);
        default:
            return true;
    }
}

// Cost to the goal through the edge with a given traversal type (the cost() of each family)
TASI_XD inline float type_cost(int ty, int sit, const Tri &t) {
    switch (ty) {
        case TY_I:
            if (sit == SIT_CORNER) return t.g2 + t.b + cathf(t.c, t.b);
            if (sit == SIT_CONTIG) return t.g2 + (1.0f - t.q) * t.b + cathf(t.c, t.b);
            return t.g2 + t.b + (1.0f + t.p) * cathf(t.c, t.b);
        case TY_II:
            if (sit == SIT_CORNER) return t.g1 + cathf(t.c, t.f);
            if (sit == SIT_CONTIG) return t.g1 + (1.0f - t.q) * cathf(t.c, t.f);
            return t.g2 + cathf(t.c, t.f) + (1.0f - t.p) * t.f;
        case TY_III:
            if (sit == SIT_CORNER) return t.g1 + t.b;
            if (sit == SIT_CONTIG) return t.g1 + (1.0f - t.q) * t.b;
            return t.g1 + t.b + t.p * cathf(t.c, t.b);
        case TY_A:
            if (sit == SIT_CORNER) return t.g2 + t.c * TASI_SQRT2F;
            if (sit == SIT_CONTIG) return t.g2 + t.c * hyp(1.0f, 1.0f - t.q);
            return t.g2 + t.c * hyp(1.0f - t.p, 1.0f);
        default:  // TY_B (never asked from an opposite edge)
            if (sit == SIT_CORNER) return t.g1 + t.c;
            return t.g1 + t.c * (1.0f - t.q);
    }
}

TASI_XD inline void hop_end(Hop &h, int px, int py) {
    h.x[h.n] = (float) px;
    h.y[h.n] = (float) py;
    ++h.n;
}

// The points and piece costs of a traversal type (the additions() and stepcosts() of each family)
TASI_XD inline void type_pieces(int ty, int sit, const Tri &t, Hop &h) {
    h.n = 0;
    h.nc = 1;
    h.sc[1] = 0.0f;
    switch (ty) {
        case TY_I: {  // along the line p0-p1 through the cheap neighbour cell, then across to p2
            const float ca = cathf(t.c, t.b);
            float x, far;
            if (sit == SIT_CORNER) {
                x = 1.0f - t.b / ca;
                far = hyp(1.0f - x, 1.0f);
                on_line(t, x, true, h.x[0], h.y[0]);
                h.n = 1;
            } else if (sit == SIT_CONTIG) {
                x = 1.0f - t.q - t.b / ca;
                far = hyp(1.0f - t.q - x, 1.0f);
                on_line(t, x, false, h.x[0], h.y[0]);
                h.n = 1;
            } else {
                x = 1.0f - (1.0f + t.p) * t.b / ca;
                far = hyp(1.0f - x, 1.0f + t.p);
                const float v = (1.0f - x) * t.p / (t.p + 1.0f);
                on_line(t, v, true, h.x[0], h.y[0]);
                on_line(t, v + x, true, h.x[1], h.y[1]);
                h.n = 2;
            }
            hop_end(h, t.p2x, t.p2y);
            h.sc[0] = x * t.b;
            h.sc[1] = far * t.c;
            h.nc = 2;
            break;
        }
        case TY_II: {  // straight through the cell to a point of the edge p1-p2
            const float ca = cathf(t.c, t.f);
            float y, run;
            if (sit == SIT_CORNER) {
                y = t.f / ca;
                run = hyp(1.0f, y);
            } else if (sit == SIT_CONTIG) {
                y = (1.0f - t.q) * t.f / ca;
                run = hyp(1.0f - t.q, y);
            } else {
                y = t.p + t.f / ca;
                run = hyp(1.0f, y - t.p);
            }
            on_edge(t, y, h.x[0], h.y[0]);
            h.n = 1;
            h.sc[0] = run * t.c;
            break;
        }
        case TY_III:  // to p1, along the line p0-p1
            if (sit == SIT_OPPOSITE) {
                const float x = t.p * t.b / cathf(t.c, t.b);
                on_line(t, x, true, h.x[0], h.y[0]);
                h.n = 1;
                h.sc[0] = hyp(x, t.p) * t.c;
                h.sc[1] = (1.0f - x) * t.b;
                h.nc = 2;
            } else {
                h.sc[0] = sit == SIT_CORNER ? t.b : (1.0f - t.q) * t.b;
            }
            hop_end(h, t.p1x, t.p1y);
            break;
        case TY_A:  // straight to p2
            h.sc[0] = sit == SIT_CORNER ? t.c * TASI_SQRT2F
                    : sit == SIT_CONTIG ? t.c * hyp(1.0f, 1.0f - t.q) : t.c * hyp(1.0f - t.p, 1.0f);
            hop_end(h, t.p2x, t.p2y);
            break;
        default:    // TY_B: straight to p1
            h.sc[0] = sit == SIT_CORNER ? t.c : (1.0f - t.q) * t.c;
            hop_end(h, t.p1x, t.p1y);
            break;
    }
    float acc = 0.0f;
    acc = acc + h.sc[0];
    if (h.nc == 2) acc = acc + h.sc[1];
    h.step = acc;
}

// Corner: the decision tree of InterpolatedTraversal::traversalFromCorner (:495-520) / directTraversalFromCorner
// (:667-676); the f*f <= sqrt(c*c - b*b) test is the reference's
TASI_XD inline int pick_corner(const Tri &t, bool indirect) {
    if (indirect && t.c > t.b) {
        if (t.f <= 0.0f || t.f * t.f <= cathf(t.c, t.b)) return TY_III;
        if (t.f <= t.b && t.c > t.f * TASI_SQRT2F) return TY_II;
        if (t.f > t.b && t.c > t.b * TASI_SQRT2F) return TY_I;
        return TY_A;
    }
    if (t.f <= 0.0f) return TY_B;
    if (t.f * TASI_SQRT2F < t.c) return TY_II;
    return TY_A;
}

// the first lowest admissible cost wins (std::min_element over the condcost() array, :558-566, :605-620)
struct Pick {
    int ty, side;
    float cost;
    bool any;
    TASI_XD void offer(int ty_, int side_, float c) {
        if (!any || c < cost) {
            ty = ty_;
            side = side_;
            cost = c;
            any = true;
        }
    }
};

// The ring of nodes around a position in the reference's order (Graph::consecutive_neighbors, Graph.cpp:151-200).
// Returns the ring length (6 on an edge, 8 on a vertex); candidate k is the node pair (k, (k+1) % len).
TASI_XD inline int ring_of(float px, float py, int *rx, int *ry) {
    float fx, fy;
    const float dx = modff(px, &fx), dy = modff(py, &fy);
    const int x = (int) fx, y = (int) fy;
    if (0.0f < dx && dx < 1.0f) {
        rx[0] = x;     ry[0] = y;
        rx[1] = x;     ry[1] = y - 1;
        rx[2] = x + 1; ry[2] = y - 1;
        rx[3] = x + 1; ry[3] = y;
        rx[4] = x + 1; ry[4] = y + 1;
        rx[5] = x;     ry[5] = y + 1;
        return 6;
    }
    if (0.0f < dy && dy < 1.0f) {
        rx[0] = x;     ry[0] = y;
        rx[1] = x + 1; ry[1] = y;
        rx[2] = x + 1; ry[2] = y + 1;
        rx[3] = x;     ry[3] = y + 1;
        rx[4] = x - 1; ry[4] = y + 1;
        rx[5] = x - 1; ry[5] = y;
        return 6;
    }
    const int ox[8] = {1, 1, 0, -1, -1, -1, 0, 1}, oy[8] = {0, 1, 1, 1, 0, -1, -1, -1};
    for (int k = 0; k < 8; ++k) {
        rx[k] = x + ox[k];
        ry[k] = y + oy[k];
    }
    return 8;
}

// Graph::is_valid_vertex (Graph.cpp:65-69)
TASI_XD inline bool is_vertex(const FieldView &v, float px, float py) {
    return ceilf(px) == px && ceilf(py) == py && px >= 0.0f && py >= 0.0f && px <= (float) v.L && py <= (float) v.W;
}

// Crossing the edge (a, b) of the ring from the position (px, py): LinearInterpolationPathExtractor::
// traversalFromCorner / traversalFromContiguousEdge / traversalFromOppositeEdge (_impl.h:60-145) followed by the
// matching InterpolatedTraversal selection. h.n == 0 when the edge cannot be crossed.
TASI_XD inline void hop_over_edge(const FieldView &v, float px, float py, bool at_vertex, int ax, int ay, int bx,
                                  int by, bool indirect, Hop &h) {
    h.n = 0;
    h.nc = 0;
    h.to_goal = inf_f();
    h.step = 0.0f;
    const bool a_in_line = px == (float) ax || py == (float) ay;
    const bool b_in_line = px == (float) bx || py == (float) by;
    if (at_vertex || a_in_line || b_in_line) {
        Tri t;
        t.p0x = px;
        t.p0y = py;
        t.p1x = a_in_line ? ax : bx;
        t.p1y = a_in_line ? ay : by;
        t.p2x = a_in_line ? bx : ax;
        t.p2y = a_in_line ? by : ay;
        t.g1 = v.node_val(t.p1x, t.p1y);
        t.g2 = v.node_val(t.p2x, t.p2y);
        tri_costs(v, t);
        t.p = 0.0f;
        t.q = at_vertex ? 0.0f : 1.0f - fabsf((float) t.p1y - py) - fabsf((float) t.p1x - px);
        if ((t.g1 == inf_f() && t.g2 == inf_f()) || t.c == inf_f()) return;
        t.f = t.g1 - t.g2;
        const int sit = at_vertex ? SIT_CORNER : SIT_CONTIG;
        int ty;
        if (at_vertex) {
            ty = pick_corner(t, indirect);
            h.to_goal = type_cost(ty, sit, t);
        } else {
            Pick pk = {TY_I, 0, inf_f(), false};
            if (indirect) pk.offer(TY_I, 0, type_ok(TY_I, sit, t) ? type_cost(TY_I, sit, t) : inf_f());
            pk.offer(TY_II, 0, type_ok(TY_II, sit, t) ? type_cost(TY_II, sit, t) : inf_f());
            if (indirect) pk.offer(TY_III, 0, type_ok(TY_III, sit, t) ? type_cost(TY_III, sit, t) : inf_f());
            pk.offer(TY_A, 0, type_cost(TY_A, sit, t));
            pk.offer(TY_B, 0, type_cost(TY_B, sit, t));
            ty = pk.ty;
            h.to_goal = pk.cost;
        }
        type_pieces(ty, sit, t, h);
        return;
    }
    // the edge opposite to the position's own: two mirrored triangles, one per end of the position's edge
    Tri t[2];
    const bool same_x = ax == bx;
    for (int s = 0; s < 2; ++s) {
        t[s].p1x = s ? bx : ax;
        t[s].p1y = s ? by : ay;
        t[s].p2x = s ? ax : bx;
        t[s].p2y = s ? ay : by;
        t[s].p0x = same_x ? px : (float) t[s].p1x;
        t[s].p0y = same_x ? (float) t[s].p1y : py;
        t[s].q = 0.0f;
    }
    t[0].g1 = t[1].g2 = v.node_val(ax, ay);
    t[0].g2 = t[1].g1 = v.node_val(bx, by);
    tri_costs(v, t[0]);
    tri_costs(v, t[1]);
    t[0].p = fabsf(py - t[0].p0y) + fabsf(px - t[0].p0x);
    t[1].p = 1.0f - t[0].p;
    // traversalFromOppositeEdge tests g(a) twice (cell1.g1 and cell2.g2 are the same node, :591)
    if (t[0].g1 == inf_f() || t[0].c == inf_f()) return;
    t[0].f = t[0].g1 - t[0].g2;
    t[1].f = -t[0].f;
    Pick pk = {TY_I, 0, inf_f(), false};
    for (int ty = indirect ? TY_I : TY_II; ty <= TY_A; ++ty) {
        if (!indirect && ty == TY_III) continue;
        for (int s = 0; s < 2; ++s)
            pk.offer(ty, s, type_ok(ty, SIT_OPPOSITE, t[s]) ? type_cost(ty, SIT_OPPOSITE, t[s]) : inf_f());
    }
    h.to_goal = pk.cost;
    type_pieces(pk.ty, SIT_OPPOSITE, t[pk.side], h);
}

// LinearInterpolationPathExtractor::getPathAdditions with do_lookahead = false (_impl.h:166-213): the lowest cost
// to the goal over candidates k0..k1-1 of the ring of (px, py); INF when none can be crossed (the caller turns that
// into 0, the cost of a value-initialised PathAdditions).
TASI_XD inline float ring_floor(const FieldView &v, float px, float py, bool indirect, int k0, int k1) {
    int rx[8], ry[8];
    const int len = ring_of(px, py, rx, ry);
    const bool vert = is_vertex(v, px, py);
    float best = inf_f();
    Hop h;
    for (int k = k0; k < k1 && k < len; ++k) {
        const int j = k + 1 == len ? 0 : k + 1;
        if (!v.node_ok(rx[k], ry[k]) || !v.node_ok(rx[j], ry[j])) continue;
        hop_over_edge(v, px, py, vert, rx[k], ry[k], rx[j], ry[j], indirect, h);
        if (h.n && h.to_goal < best) best = h.to_goal;
    }
    return best;
}

// Sequential chooser (host tests, single-thread device use): the full getPathAdditions with the one-step
// look-ahead of Perkins et al. (_impl.h:188-194). Returns false when no edge qualifies.
struct SerialChooser {
    TASI_XD bool operator()(const FieldView &v, float px, float py, bool lookahead, bool indirect, Hop &out) const {
        int rx[8], ry[8];
        const int len = ring_of(px, py, rx, ry);
        const bool vert = is_vertex(v, px, py);
        float best = inf_f();
        bool found = false;
        Hop h;
        for (int k = 0; k < len; ++k) {
            const int j = k + 1 == len ? 0 : k + 1;
            if (!v.node_ok(rx[k], ry[k]) || !v.node_ok(rx[j], ry[j])) continue;
            hop_over_edge(v, px, py, vert, rx[k], ry[k], rx[j], ry[j], indirect, h);
            if (!h.n) continue;
            const float ex = h.x[h.n - 1], ey = h.y[h.n - 1];
            if (lookahead && !is_vertex(v, ex, ey)) {
                float ahead = ring_floor(v, ex, ey, indirect, 0, 8);
                if (ahead == inf_f()) ahead = 0.0f;
                if (ahead > h.to_goal) continue;
            }
            if (h.to_goal < best) {
                best = h.to_goal;
                out = h;
                found = true;
            }
        }
        return found;
    }
};

enum { WALK_GOAL = 0, WALK_MAX_STEPS = 1, WALK_STUCK = 2, WALK_FULL = 3 };

struct PathOut {
    float *xy;       // cap_points x 2
    int cap_points;
    float *costs;    // cap_costs piece costs, may be null
    int cap_costs;
    int n_points, n_costs;
    float total_cost, total_dist;
};

// LinearInterpolationPathExtractor::extract_path (_impl.h:12-58). `writer` = this caller stores the output (the
// device kernel walks with a whole warp in lock step and lane 0 writes). A position with no crossable edge ends
// the walk (WALK_STUCK): the reference would repeat the same empty step max_steps times.
template <class Chooser>
TASI_XD inline int walk_path(const FieldView &v, const Chooser &choose, float sx, float sy, float gx, float gy,
                             int max_steps, bool lookahead, bool indirect, bool writer, PathOut &o) {
    o.n_points = o.n_costs = 0;
    o.total_cost = o.total_dist = 0.0f;
    if (o.cap_points < 1) return WALK_FULL;
    if (writer) {
        o.xy[0] = sx;
        o.xy[1] = sy;
    }
    o.n_points = 1;
    float lx = sx, ly = sy;
    int steps = 0;
    Hop h;
    do {
        if (!choose(v, lx, ly, lookahead, indirect, h)) return WALK_STUCK;
        if (o.n_points + h.n > o.cap_points || (o.costs && o.n_costs + h.nc > o.cap_costs)) return WALK_FULL;
        float step_dist = 0.0f;
        for (int i = 0; i < h.n; ++i) {
            step_dist += hyp(lx - h.x[i], ly - h.y[i]);
            lx = h.x[i];
            ly = h.y[i];
            if (writer) {
                o.xy[2 * (o.n_points + i)] = lx;
                o.xy[2 * (o.n_points + i) + 1] = ly;
            }
        }
        if (writer && o.costs)
            for (int i = 0; i < h.nc; ++i) o.costs[o.n_costs + i] = h.sc[i];
        o.n_points += h.n;
        o.n_costs += h.nc;
        o.total_cost += h.step;
        o.total_dist += step_dist;
        ++steps;
    } while (!(lx == gx && ly == gy) && steps < max_steps);
    return (lx == gx && ly == gy) ? WALK_GOAL : WALK_MAX_STEPS;
}

#ifdef __CUDACC__
// One warp per path. A position has at most 8 candidate edges and each candidate's look-ahead ring at most 8
// more: lanes 0-7 cross the candidates, then all 32 lanes share the 64 look-ahead crossings (lane = candidate * 4
// + quarter of its ring), and the first lowest admissible candidate is broadcast to every lane, so the warp walks
// in lock step and lane 0 stores the path.
struct WarpChooser {
    __device__ bool operator()(const FieldView &v, float px, float py, bool lookahead, bool indirect, Hop &out) const {
        const unsigned full = 0xffffffffu;
        const int lane = threadIdx.x & 31;
        int rx[8], ry[8];
        const int len = ring_of(px, py, rx, ry);
        const bool vert = is_vertex(v, px, py);
        Hop h;
        h.n = 0;
        h.nc = 0;
        h.to_goal = inf_f();
        h.step = 0.0f;
#pragma unroll
        for (int i = 0; i < 3; ++i) h.x[i] = h.y[i] = 0.0f;
        h.sc[0] = h.sc[1] = 0.0f;
        if (lane < len) {
            const int j = lane + 1 == len ? 0 : lane + 1;
            if (v.node_ok(rx[lane], ry[lane]) && v.node_ok(rx[j], ry[j]))
                hop_over_edge(v, px, py, vert, rx[lane], ry[lane], rx[j], ry[j], indirect, h);
        }
        const float ex = h.n ? h.x[h.n - 1] : 0.0f, ey = h.n ? h.y[h.n - 1] : 0.0f;
        const int need = lookahead && h.n && !is_vertex(v, ex, ey);
        const int cand = lane >> 2, quarter = lane & 3;
        const float qx = __shfl_sync(full, ex, cand), qy = __shfl_sync(full, ey, cand);
        const int qneed = __shfl_sync(full, need, cand);
        float ahead = inf_f();
        if (qneed) ahead = ring_floor(v, qx, qy, indirect, 2 * quarter, 2 * quarter + 2);
#pragma unroll
        for (int off = 1; off <= 2; off <<= 1) {
            const float o = __shfl_xor_sync(full, ahead, off);
            ahead = o < ahead ? o : ahead;
        }
        ahead = __shfl_sync(full, ahead, (lane * 4) & 31);
        if (ahead == inf_f()) ahead = 0.0f;
        const bool ok = lane < 8 && h.n && !(need && ahead > h.to_goal) && h.to_goal < inf_f();
        const float mine = ok ? h.to_goal : inf_f();
        float low = mine;
#pragma unroll
        for (int off = 16; off; off >>= 1) {
            const float o = __shfl_xor_sync(full, low, off);
            low = o < low ? o : low;
        }
        if (!(low < inf_f())) return false;
        const int win = __ffs(__ballot_sync(full, ok && mine == low)) - 1;
        out.n = __shfl_sync(full, h.n, win);
        out.nc = __shfl_sync(full, h.nc, win);
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            out.x[i] = __shfl_sync(full, h.x[i], win);
            out.y[i] = __shfl_sync(full, h.y[i], win);
        }
        out.sc[0] = __shfl_sync(full, h.sc[0], win);
        out.sc[1] = __shfl_sync(full, h.sc[1], win);
        out.to_goal = __shfl_sync(full, h.to_goal, win);
        out.step = __shfl_sync(full, h.step, win);
        return true;
    }
};

// row0 / rows: the walk sees only cell rows row0 .. row0 + rows - 1 as its map, in coordinates relative to row0
// (rows == 0: the whole map) — one block of a stack of independent maps (tasi_set_goals)
struct PathQuery { float sx, sy, gx, gy; int row0, rows; };

// q == nullptr: one walk, its query travels in the launch arguments (q0)
__global__ void __launch_bounds__(128) extract_paths_kernel(FieldView v, const PathQuery *q, PathQuery q0, int n,
                                                            int max_steps,
                                                            int lookahead, int indirect, float *xy, int cap_points,
                                                            float *costs, int cap_costs, int *n_points, int *n_costs,
                                                            float *totals, int *status) {
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (w >= n) return;
    const bool writer = (threadIdx.x & 31) == 0;
    PathOut o = {xy + (int64_t) w * cap_points * 2, cap_points, costs ? costs + (int64_t) w * cap_costs : nullptr,
                 cap_costs, 0, 0, 0.0f, 0.0f};
    const PathQuery pq = q ? q[w] : q0;
    if (pq.rows > 0) {
        v.g += (int64_t) pq.row0 * v.gpitch;
        v.cost += (int64_t) pq.row0 * v.cpitch;
        v.L = pq.rows;
    }
    const int rc = walk_path(v, WarpChooser(), pq.sx, pq.sy, pq.gx, pq.gy, max_steps, lookahead != 0, indirect != 0,
                             writer, o);
    if (writer) {
        n_points[w] = o.n_points;
        n_costs[w] = o.n_costs;
        totals[2 * w] = o.total_cost;
        totals[2 * w + 1] = o.total_dist;
        status[w] = rc;
    }
}
#endif  // __CUDACC__
}  // namespace tasi
