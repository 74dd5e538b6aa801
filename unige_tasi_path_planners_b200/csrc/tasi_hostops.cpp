// The local-update operators of tasi_update.cuh compiled for the host (g++ -ffp-contract=off), exported with
// a C ABI so the CPU test-suite can compare them value by value with the oracle / the reference. Not used by
// the product path: the solver evaluates the same header on the device.
#include "tasi_update.cuh"
#include "tasi_extract.cuh"
#include "tasi_parents.cuh"

extern "C" {
float tasi_host_fd_tri(float g1, float g2, float b, float c) { return tasi::fd_tri(g1, g2, b, c); }
float tasi_host_mfd_tri(float g1, float g2, float c) { return tasi::mfd_tri(g1, g2, c); }
float tasi_host_dfm_stencil(float ga, float gb, float tau, float h) { return tasi::dfm_st(ga, gb, tau, h); }

// value of one element from its 8 neighbours (n,s,w,e,nw,ne,sw,se) and the 4 cell costs (tl,tr,bl,br);
// MS-DFM reads only cost[0] (tau)
float tasi_host_update(int kind, const float *nbr, const float *cost) {
    tasi::Nbr8 g = {nbr[0], nbr[1], nbr[2], nbr[3], nbr[4], nbr[5], nbr[6], nbr[7]};
    switch (kind) {
        case TASI_KIND_FD: return tasi::fd_node(g, cost[0], cost[1], cost[2], cost[3]);
        case TASI_KIND_DFM: return tasi::dfm_cell(g, cost[0]);
        default: return tasi::mfd_node(g, cost[0], cost[1], cost[2], cost[3]);
    }
}

// batched forms for the tests
void tasi_host_fd_tri_n(const float *g1, const float *g2, const float *b, const float *c, float *out, long n) {
    for (long i = 0; i < n; ++i) out[i] = tasi::fd_tri(g1[i], g2[i], b[i], c[i]);
}
void tasi_host_mfd_tri_n(const float *g1, const float *g2, const float *c, float *out, long n) {
    for (long i = 0; i < n; ++i) out[i] = tasi::mfd_tri(g1[i], g2[i], c[i]);
}
void tasi_host_dfm_stencil_n(const float *ga, const float *gb, const float *tau, const float *h, float *out, long n) {
    for (long i = 0; i < n; ++i) out[i] = tasi::dfm_st(ga[i], gb[i], tau[i], h[i]);
}

// the path walk of tasi_extract.cuh over host arrays (field: row-major ex x ey floats; cost: L x W bytes)
int tasi_host_extract(int cells, const float *field, const uint8_t *cost, int L, int W, int thr, float sx, float sy,
                      float gx, float gy, int max_steps, int lookahead, int indirect, float *xy, int cap_points,
                      float *costs, int cap_costs, int *n_points, int *n_costs, float *total_cost, float *total_dist) {
    tasi::FieldView v = {field, cells ? W : W + 1, cost, W, L, W, thr, cells};
    tasi::PathOut o = {xy, cap_points, costs, cap_costs, 0, 0, 0.0f, 0.0f};
    const int rc = tasi::walk_path(v, tasi::SerialChooser(), sx, sy, gx, gy, max_steps, lookahead != 0,
                                   indirect != 0, true, o);
    *n_points = o.n_points;
    *n_costs = o.n_costs;
    *total_cost = o.total_cost;
    *total_dist = o.total_dist;
    return rc;
}

// the parent field of tasi_parents.cuh over host arrays (field: ex x ey floats; cost: L x W bytes); out: ex*ey*4 bytes
void tasi_host_parents(int kind, const float *field, const uint8_t *cost, int L, int W, int thr, signed char *out) {
    const bool cells = kind == TASI_KIND_DFM;
    const int ex = cells ? L : L + 1, ey = cells ? W : W + 1;
    auto G = [&](int x, int y) { return (x < 0 || y < 0 || x >= ex || y >= ey) ? INFINITY : field[(size_t) x * ey + y]; };
    auto C = [&](int x, int y) {
        if (x < 0 || y < 0 || x >= L || y >= W) return (float) INFINITY;
        const int v = cost[(size_t) x * W + y];
        return v >= thr ? (float) INFINITY : (float) v;
    };
    for (int x = 0; x < ex; ++x)
        for (int y = 0; y < ey; ++y) {
            const tasi::Nbr8 g = {G(x - 1, y), G(x + 1, y), G(x, y - 1), G(x, y + 1),
                                  G(x - 1, y - 1), G(x - 1, y + 1), G(x + 1, y - 1), G(x + 1, y + 1)};
            const tasi::Parents p = cells ? tasi::cell_parents(g, C(x, y), G(x, y))
                                          : tasi::node_parents(kind, g, C(x - 1, y - 1), C(x - 1, y), C(x, y - 1), C(x, y), G(x, y));
            signed char *o = out + 4 * ((size_t) x * ey + y);
            o[0] = p.dx1; o[1] = p.dy1; o[2] = p.dx2; o[3] = p.dy2;
        }
}

// the 8 triangle values of node (x, y) in the order of Graph::neighbors_8 (triangle k = (b_k, ccw(b_k))), for the tests
void tasi_host_node_triangles(int kind, const float *field, const uint8_t *cost, int L, int W, int thr, int x, int y,
                              float *out) {
    const int ex = L + 1, ey = W + 1;
    auto G = [&](int a, int b) { return (a < 0 || b < 0 || a >= ex || b >= ey) ? INFINITY : field[(size_t) a * ey + b]; };
    auto C = [&](int a, int b) {
        if (a < 0 || b < 0 || a >= L || b >= W) return (float) INFINITY;
        const int v = cost[(size_t) a * W + b];
        return v >= thr ? (float) INFINITY : (float) v;
    };
    const float n = G(x - 1, y), s = G(x + 1, y), w = G(x, y - 1), e = G(x, y + 1), nw = G(x - 1, y - 1),
                ne = G(x - 1, y + 1), sw = G(x + 1, y - 1), se = G(x + 1, y + 1);
    const float tl = C(x - 1, y - 1), tr = C(x - 1, y), bl = C(x, y - 1), br = C(x, y);
    const float g1[8] = {n, n, w, w, s, s, e, e}, g2[8] = {ne, nw, nw, sw, sw, se, se, ne};
    const float b[8] = {tl, tr, bl, tl, br, bl, tr, br}, c[8] = {tr, tl, tl, bl, bl, br, br, tr};
    for (int k = 0; k < 8; ++k)
        out[k] = kind == TASI_KIND_FD ? tasi::fd_tri(g1[k], g2[k], b[k], c[k]) : tasi::mfd_tri(g1[k], g2[k], c[k]);
}
}
