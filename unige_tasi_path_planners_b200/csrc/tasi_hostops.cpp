// The local-update operators of tasi_update.cuh compiled for the host (g++ -ffp-contract=off), exported with
// a C ABI so the CPU test-suite can compare them value by value with the oracle / the reference. Not used by
// the product path: the solver evaluates the same header on the device.
#include "tasi_update.cuh"
#include "tasi_extract.cuh"

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
}
