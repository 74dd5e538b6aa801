// Parent (back-pointer) field of a converged value field, as __host__ __device__ functions.
//
// The reference's level-1 planners keep, next to (g, rhs), the INFO of every element: which neighbours gave its rhs.
//   FD* / MFD*   a Node b: rhs came from the triangle (b, ccw(b)) (FieldDPlanner_impl.h:91-95, min_rhs<1> :196-208,
//                ShiftedGridPlanner_impl.h:139-148; ccw ring Graph.cpp:232-245)
//   MS-DFM       a pair of Cells: the lower and the higher cell of the winning two-neighbour stencil, the second
//                absent when the update was one-sided (DynamicFastMarching_impl.h:212-268, 322-342, best_cell :344-351)
// Here the parents are not stored while solving (the solver is label-correcting, a stored pointer would be rewritten
// many times); they are recomputed from the converged field on demand: the same operators the solver applies
// (tasi_update.cuh), evaluated once per element, with the reference's loop order and tie rules (min_rhs<1>: the last
// neighbour whose triangle equals the minimum wins; best_cell: ties go to the second cell).
//
// Output per element: 4 signed bytes (dx1, dy1, dx2, dy2), offsets of the two parents; (0, 0) = absent. Goals and
// unreached elements have no parents.
#pragma once
#include "tasi_update.cuh"

namespace tasi {

struct Parents {
    signed char dx1, dy1, dx2, dy2;
};

// FD* / MFD* node. g = the 8 neighbour values, tl..br = the 4 cell costs around the node, cur = its own value.
TASI_HD Parents node_parents(int kind, const Nbr8 &g, float tl, float tr, float bl, float br, float cur) {
    const Parents none = {0, 0, 0, 0};
    const float INF = inf_f();
    if (!(cur < INF) || cur == 0.0f) return none;
    // the 8 neighbours b in the order of Graph::neighbors_8 (top, top-left, left, bottom-left, bottom, bottom-right,
    // right, top-right; Graph.cpp:71-76) with the triangle (b, ccw(b)): p1 = its orthogonal node, p2 = its diagonal one
    float v[8];
    if (kind == TASI_KIND_FD) {
        v[0] = fd_tri(g.n, g.ne, tl, tr);
        v[1] = fd_tri(g.n, g.nw, tr, tl);
        v[2] = fd_tri(g.w, g.nw, bl, tl);
        v[3] = fd_tri(g.w, g.sw, tl, bl);
        v[4] = fd_tri(g.s, g.sw, br, bl);
        v[5] = fd_tri(g.s, g.se, bl, br);
        v[6] = fd_tri(g.e, g.se, tr, br);
        v[7] = fd_tri(g.e, g.ne, br, tr);
    } else {
        v[0] = mfd_tri(g.n, g.ne, tr);
        v[1] = mfd_tri(g.n, g.nw, tl);
        v[2] = mfd_tri(g.w, g.nw, tl);
        v[3] = mfd_tri(g.w, g.sw, bl);
        v[4] = mfd_tri(g.s, g.sw, bl);
        v[5] = mfd_tri(g.s, g.se, br);
        v[6] = mfd_tri(g.e, g.se, br);
        v[7] = mfd_tri(g.e, g.ne, tr);
    }
    float m = INF;
    int win = -1;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        if (v[k] < m) m = v[k];
        if (v[k] == m && m < INF) win = k;  // min_rhs<1>: `if (rhs == cost) bptr = sp` — the last minimum wins
    }
    if (win < 0) return none;
    const signed char bx[8] = {-1, -1, 0, 1, 1, 1, 0, -1}, by[8] = {0, -1, -1, -1, 0, 1, 1, 1};
    const int ccw = (win + 7) & 7;  // ccw(top) = top-right, ccw(top-left) = top, ... (Graph.cpp:236-237)
    const Parents p = {bx[win], by[win], bx[ccw], by[ccw]};
    return p;
}

// MS-DFM cell. g = the 8 neighbour values, tau = the cell's own cost, cur = its own value.
TASI_HD Parents cell_parents(const Nbr8 &g, float tau, float cur) {
    const Parents none = {0, 0, 0, 0};
    const float INF = inf_f();
    if (!(cur < INF) || cur == 0.0f || tau == INF) return none;
    // best_cell(a, b): a only if strictly lower
    const bool n_best = g.n < g.s, w_best = g.w < g.e, nw_best = g.nw < g.se, sw_best = g.sw < g.ne;
    const float oa = n_best ? g.n : g.s, ob = w_best ? g.w : g.e;
    const float da = nw_best ? g.nw : g.se, db = sw_best ? g.sw : g.ne;
    const float ortho = dfm_st(oa, ob, tau, 1.0f), diag = dfm_st(da, db, tau, TASI_SQRT2F);
    const bool use_diag = diag < ortho;
    float ga = use_diag ? da : oa, gb = use_diag ? db : ob;
    signed char ax, ay, bx, by;
    if (use_diag) {
        ax = nw_best ? -1 : 1; ay = nw_best ? -1 : 1;   // top-left / bottom-right
        bx = sw_best ? 1 : -1; by = sw_best ? -1 : 1;   // bottom-left / top-right
    } else {
        ax = n_best ? -1 : 1; ay = 0;                   // top / bottom
        bx = 0; by = w_best ? -1 : 1;                   // left / right
    }
    if (ga > gb) {  // compute_optimal_cost :325-328: the lower one first
        const float t = ga; ga = gb; gb = t;
        signed char s = ax; ax = bx; bx = s;
        s = ay; ay = by; by = s;
    }
    if (ga == INF && gb == INF) return none;
    const float th = tau * (use_diag ? TASI_SQRT2F : 1.0f);
    Parents p = {ax, ay, 0, 0};
    if (th > gb - ga) {
        p.dx2 = bx;
        p.dy2 = by;
    }
    return p;
}

#ifdef __CUDACC__
// One thread per element of the window (x0, y0, nx, ny) of the field; margins of the padded arrays read as INF /
// obstacle, so no bounds tests are needed inside the grid.
template <int KIND>
__global__ void parents_kernel(const float *g, int64_t gpitch, const uint8_t *cost, int64_t cpitch, int thr, int x0,
                               int y0, int nx, int ny, Parents *out) {
    const int64_t n = (int64_t) nx * ny;
    for (int64_t i = blockIdx.x * (int64_t) blockDim.x + threadIdx.x; i < n; i += (int64_t) gridDim.x * blockDim.x) {
        const int x = x0 + (int) (i / ny), y = y0 + (int) (i % ny);
        const float *c = g + (int64_t) x * gpitch + y;
        const Nbr8 nb = {c[-gpitch], c[gpitch], c[-1], c[1], c[-gpitch - 1], c[-gpitch + 1], c[gpitch - 1], c[gpitch + 1]};
        auto cell = [&](int cx, int cy) {
            const int v = cost[(int64_t) cx * cpitch + cy];
            return v >= thr ? inf_f() : (float) v;
        };
        out[i] = KIND == TASI_KIND_DFM ? cell_parents(nb, cell(x, y), c[0])
                                       : node_parents(KIND, nb, cell(x - 1, y - 1), cell(x - 1, y), cell(x, y - 1),
                                                      cell(x, y), c[0]);
    }
}
#endif

}  // namespace tasi
