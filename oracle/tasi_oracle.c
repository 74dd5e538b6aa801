/* TEST INFRASTRUCTURE — not product code. Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference leg may load this library.
 *
 * Plain-C restatement of the cost-propagation core of the three replanners of
 * roncapat/UNIGE-TASI-path-planners. Every function cites the reference file:line it follows
 * (paths relative to the reference root). Parity is PINNED: tests/test_oracle_vs_ref.py checks this
 * file against the unmodified reference planners (oracle/_ref, built by oracle/Makefile from the
 * reference sources) on the four Tests/Tests/ *.bmp maps and on seeded maps, and the outputs of those
 * runs are committed as fixtures under tests/golden/.
 *
 * Build: gcc -std=c11 -O2 -ffp-contract=off (no FMA contraction: the float results must be the
 * reference's).
 *
 * Conventions (ProjectToolkit/Graph.cpp:22-34, 262-268): the map is L rows (x) by W columns (y) of
 * uint8, map[x*W+y]; cost(cell) = INF outside the map or if byte >= thr, else (float)byte. FD* and
 * MFD* values live on the (L+1)x(W+1) nodes, MS-DFM values on the LxW cells.
 */
#include <math.h>
#include <pthread.h>
#include <unistd.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define TASI_FD 0
#define TASI_DFM 1
#define TASI_MFD 2

/* ProjectToolkit/Macros.cpp:2 */
static const float SQRT2F = 1.41421356237309504880168872420969807856967187537694f;

/* ProjectToolkit/include/Macros.h:9,12 : SQUARE(x)=float(x)*float(x); CATH(x,y)=sqrt(SQUARE(x)-SQUARE(y)) */
static inline float sq(float v) { return v * v; }
static inline float cath(float u, float v) { return sqrtf(sq(u) - sq(v)); }

/* FieldDStar/FieldDPlanner_impl.h:292-319 with the Corner::cost bodies of
 * ProjectToolkit/InterpolatedTraversal.cpp:9 (I), 126 (II), 237 (III), 325 (A), 404 (B).
 * g1 = value of the orthogonal neighbour, g2 = of the diagonal one, c = cost of the cell holding the
 * triangle, b = cost of the cell across the edge n-p1. */
float tasi_oracle_fd_tri(float g1, float g2, float b, float c) {
    if (g1 == INFINITY && g2 == INFINITY) return INFINITY; /* :292 */
    if (c == INFINITY) return INFINITY;                    /* :295 */
    float f = g1 - g2;                                     /* :297 */
    if (c > b) {                                           /* :299 */
        if (f <= 0 || sq(f) <= cath(c, b)) return g1 + b;              /* III :300-301 (f^2 vs sqrt: as written) */
        else if (f <= b && c > f * SQRT2F) return g1 + cath(c, f);     /* II  :302-303 */
        else if (f > b && c > b * SQRT2F) return g2 + b + cath(c, b);  /* I   :304-305 */
        else return g2 + c * SQRT2F;                                   /* A   :307 */
    } else {
        if (f <= 0) return g1 + c;                                     /* B   :310-311 */
        else if (f * SQRT2F < c) return g1 + cath(c, f);               /* II  :312-313 */
        else return g2 + c * SQRT2F;                                   /* A   :315 */
    }
}

/* ShiftedGridFastMarching/ShiftedGridPlanner_impl.h:422-436 (note <= where FD* has <). */
float tasi_oracle_mfd_tri(float g1, float g2, float c) {
    if (g1 == INFINITY && g2 == INFINITY) return INFINITY;
    if (c == INFINITY) return INFINITY;
    float f = g1 - g2;
    if (f <= 0) return g1 + c;
    else if (f * SQRT2F <= c) return g1 + cath(c, f);
    else return g2 + c * SQRT2F;
}

/* DynamicFastMarching/DynamicFastMarching_impl.h:322-342 (compute_optimal_cost), value part only. */
float tasi_oracle_dfm_stencil(float ga, float gb, float tau, float h) {
    if (ga > gb) { float t = ga; ga = gb; gb = t; }
    if (ga == INFINITY && gb == INFINITY) return INFINITY;
    if ((tau * h) > (gb - ga))
        return (ga + gb + sqrtf((float) (2 * sq(tau * h) - sq(gb - ga)))) * 0.5f;
    return ga + tau * h;
}

typedef struct {
    const uint8_t *map;
    int L, W, thr;
    int ex, ey; /* element grid extents: nodes (L+1,W+1) or cells (L,W) */
    const float *g;
} ctx_t;

/* ProjectToolkit/Graph.cpp:262-268 */
static inline float cell_cost(const ctx_t *c, int x, int y) {
    if (x < 0 || y < 0 || x >= c->L || y >= c->W) return INFINITY;
    uint8_t v = c->map[(size_t) x * c->W + y];
    return (v >= c->thr) ? INFINITY : (float) v;
}

/* ExpandedMap_impl.h:65-74: absent / off-grid element -> INF */
static inline float val(const ctx_t *c, int x, int y) {
    if (x < 0 || y < 0 || x >= c->ex || y >= c->ey) return INFINITY;
    return c->g[(size_t) x * c->ey + y];
}

static inline float fmin2(float a, float b) { return b < a ? b : a; }

/* FD* node update: min over the <=8 triangles (FieldDPlanner_impl.h:188-194 min_rhs<0>; triangles from
 * Graph::consecutive_neighbors(Node) Graph.cpp:202-230; b/c cells from fill_traversal_costs :322-337).
 * The 4 cells around node (x,y) (Node.cpp:44-50): TL=(x-1,y-1) TR=(x-1,y) BL=(x,y-1) BR=(x,y).
 * Off-grid nodes are skipped by the reference, which equals g=INF padding. */
static float fd_node(const ctx_t *c, int x, int y) {
    float tl = cell_cost(c, x - 1, y - 1), tr = cell_cost(c, x - 1, y);
    float bl = cell_cost(c, x, y - 1), br = cell_cost(c, x, y);
    float gn = val(c, x - 1, y), gs = val(c, x + 1, y), gw = val(c, x, y - 1), ge = val(c, x, y + 1);
    float gnw = val(c, x - 1, y - 1), gne = val(c, x - 1, y + 1);
    float gsw = val(c, x + 1, y - 1), gse = val(c, x + 1, y + 1);
    float m = INFINITY;
    m = fmin2(m, tasi_oracle_fd_tri(gn, gnw, tr, tl));
    m = fmin2(m, tasi_oracle_fd_tri(gn, gne, tl, tr));
    m = fmin2(m, tasi_oracle_fd_tri(gs, gsw, br, bl));
    m = fmin2(m, tasi_oracle_fd_tri(gs, gse, bl, br));
    m = fmin2(m, tasi_oracle_fd_tri(gw, gnw, bl, tl));
    m = fmin2(m, tasi_oracle_fd_tri(gw, gsw, tl, bl));
    m = fmin2(m, tasi_oracle_fd_tri(ge, gne, br, tr));
    m = fmin2(m, tasi_oracle_fd_tri(ge, gse, tr, br));
    return m;
}

/* MFD* node update (ShiftedGridPlanner_impl.h:258-264 min_rhs<0>; only c is read, :440-451). */
static float mfd_node(const ctx_t *c, int x, int y) {
    float tl = cell_cost(c, x - 1, y - 1), tr = cell_cost(c, x - 1, y);
    float bl = cell_cost(c, x, y - 1), br = cell_cost(c, x, y);
    float gn = val(c, x - 1, y), gs = val(c, x + 1, y), gw = val(c, x, y - 1), ge = val(c, x, y + 1);
    float gnw = val(c, x - 1, y - 1), gne = val(c, x - 1, y + 1);
    float gsw = val(c, x + 1, y - 1), gse = val(c, x + 1, y + 1);
    float m = INFINITY;
    m = fmin2(m, tasi_oracle_mfd_tri(gn, gnw, tl));
    m = fmin2(m, tasi_oracle_mfd_tri(gn, gne, tr));
    m = fmin2(m, tasi_oracle_mfd_tri(gs, gsw, bl));
    m = fmin2(m, tasi_oracle_mfd_tri(gs, gse, br));
    m = fmin2(m, tasi_oracle_mfd_tri(gw, gnw, tl));
    m = fmin2(m, tasi_oracle_mfd_tri(gw, gsw, bl));
    m = fmin2(m, tasi_oracle_mfd_tri(ge, gne, tr));
    m = fmin2(m, tasi_oracle_mfd_tri(ge, gse, br));
    return m;
}

/* MS-DFM cell update (DynamicFastMarching_impl.h:157-210 min_rhs<0>; best_cell :344-351). */
static float dfm_cell(const ctx_t *c, int x, int y) {
    float tau = cell_cost(c, x, y);
    if (tau == INFINITY) return INFINITY;
    float a = fmin2(val(c, x - 1, y), val(c, x + 1, y));         /* top / bottom   :169 */
    float b = fmin2(val(c, x, y - 1), val(c, x, y + 1));         /* left / right   :170 */
    float ortho = tasi_oracle_dfm_stencil(a, b, tau, 1.0f);      /* :184 */
    float d1 = fmin2(val(c, x - 1, y - 1), val(c, x + 1, y + 1)); /* TL / BR       :188 */
    float d2 = fmin2(val(c, x + 1, y - 1), val(c, x - 1, y + 1)); /* BL / TR       :189 */
    float diag = tasi_oracle_dfm_stencil(d1, d2, tau, SQRT2F);   /* :203 */
    return (diag < ortho) ? diag : ortho;                        /* :205-209 */
}

static float elem_update(int kind, const ctx_t *c, int x, int y) {
    switch (kind) {
        case TASI_FD: return fd_node(c, x, y);
        case TASI_DFM: return dfm_cell(c, x, y);
        default: return mfd_node(c, x, y);
    }
}

/* one element's local update on a caller-provided field (exported for operator-level tests) */
float tasi_oracle_update(int kind, const uint8_t *map, int W, int L, int thr, const float *g, int x, int y) {
    ctx_t c = {map, L, W, thr, kind == TASI_DFM ? L : L + 1, kind == TASI_DFM ? W : W + 1, g};
    return elem_update(kind, &c, x, y);
}

/* ---------------------------------------------------------------------------------------------
 * Solver 1 (order-free): Gauss-Seidel sweeps in the four raster directions, accepting only
 * decreases, until a full round changes nothing. This is the label-correcting fixed point that
 * SURVEY.md App. C.3 showed to equal the reference's D* Lite result on its closed set.
 * The goal element is pinned to 0 and never updated (FieldDPlanner_impl.h:47,146 `!= goal` guards).
 * Returns the number of sweeps; g must hold ex*ey floats. */
int tasi_oracle_solve_sweep(int kind, const uint8_t *map, int W, int L, int thr, int gx, int gy, float *g,
                            int max_rounds) {
    ctx_t c = {map, L, W, thr, kind == TASI_DFM ? L : L + 1, kind == TASI_DFM ? W : W + 1, g};
    size_t n = (size_t) c.ex * c.ey;
    for (size_t i = 0; i < n; ++i) g[i] = INFINITY;
    if (gx < 0 || gy < 0 || gx >= c.ex || gy >= c.ey) return 0;
    g[(size_t) gx * c.ey + gy] = 0.0f;
    int sweeps = 0;
    for (int round = 0; round < max_rounds; ++round) {
        int changed = 0;
        for (int dir = 0; dir < 4; ++dir) {
            int x0 = (dir & 1) ? c.ex - 1 : 0, xs = (dir & 1) ? -1 : 1;
            int y0 = (dir & 2) ? c.ey - 1 : 0, ys = (dir & 2) ? -1 : 1;
            for (int i = 0, x = x0; i < c.ex; ++i, x += xs)
                for (int j = 0, y = y0; j < c.ey; ++j, y += ys) {
                    if (x == gx && y == gy) continue;
                    float v = elem_update(kind, &c, x, y);
                    float *p = &g[(size_t) x * c.ey + y];
                    if (v < *p) { *p = v; changed = 1; }
                }
            ++sweeps;
        }
        if (!changed) break;
    }
    return sweeps;
}

/* ---------------------------------------------------------------------------------------------
 * Solver 2 (ordered): the reference's expansion order without its (g,rhs) bookkeeping — a binary
 * min-heap keyed by the tentative value (NO_HEURISTIC key = min(g,rhs), FieldDPlanner_impl.h:172-181);
 * popping an element settles it (plan(): G=RHS, :86-88) and re-evaluates its 8 neighbours with the
 * full local update (level-0 behaviour, :89-96); a neighbour that decreased is (re-)pushed, stale
 * heap entries are skipped. It runs to exhaustion (no early stop at the start, so it is a superset
 * of the reference's closed set). Returns the number of settle operations. */
typedef struct { float k; int32_t x, y; } hent_t;
typedef struct { hent_t *a; size_t n, cap; } heap_t;
static void hpush(heap_t *h, hent_t e) {
    if (h->n == h->cap) { h->cap = h->cap ? h->cap * 2 : 1024; h->a = (hent_t *) realloc(h->a, h->cap * sizeof(hent_t)); }
    size_t i = h->n++;
    while (i > 0) {
        size_t p = (i - 1) / 2;
        if (h->a[p].k <= e.k) break;
        h->a[i] = h->a[p]; i = p;
    }
    h->a[i] = e;
}
static hent_t hpop(heap_t *h) {
    hent_t top = h->a[0], e = h->a[--h->n];
    size_t i = 0;
    for (;;) {
        size_t l = 2 * i + 1, r = l + 1, m;
        if (l >= h->n) break;
        m = (r < h->n && h->a[r].k < h->a[l].k) ? r : l;
        if (h->a[m].k >= e.k) break;
        h->a[i] = h->a[m]; i = m;
    }
    if (h->n) h->a[i] = e;
    return top;
}

uint64_t tasi_oracle_solve_heap(int kind, const uint8_t *map, int W, int L, int thr, int gx, int gy, float *g) {
    ctx_t c = {map, L, W, thr, kind == TASI_DFM ? L : L + 1, kind == TASI_DFM ? W : W + 1, g};
    size_t n = (size_t) c.ex * c.ey;
    for (size_t i = 0; i < n; ++i) g[i] = INFINITY;
    if (gx < 0 || gy < 0 || gx >= c.ex || gy >= c.ey) return 0;
    g[(size_t) gx * c.ey + gy] = 0.0f;
    heap_t h = {0, 0, 0};
    hent_t e0 = {0.0f, gx, gy};
    hpush(&h, e0);
    uint64_t settled = 0;
    while (h.n) {
        hent_t e = hpop(&h);
        if (g[(size_t) e.x * c.ey + e.y] != e.k) continue; /* stale entry */
        ++settled;
        for (int dx = -1; dx <= 1; ++dx)
            for (int dy = -1; dy <= 1; ++dy) {
                int x = e.x + dx, y = e.y + dy;
                if ((dx == 0 && dy == 0) || x < 0 || y < 0 || x >= c.ex || y >= c.ey) continue;
                if (x == gx && y == gy) continue;
                float v = elem_update(kind, &c, x, y);
                float *p = &g[(size_t) x * c.ey + y];
                if (v < *p) { *p = v; hent_t ne = {v, x, y}; hpush(&h, ne); }
            }
    }
    free(h.a);
    return settled;
}

/* Fixed-point residual: number of non-goal elements whose stored value differs from the local
 * update of their neighbours (bit compare). 0 for a converged FD* / MFD* field. For MS-DFM
 * `n_above` (value below what the neighbours now give: the ulp-scale non-monotonicity noted in SURVEY.md
 * App. C.3) is reported separately; n_below must be 0 for any converged min-accept field. */
typedef struct { int kind; const ctx_t *c; int gx, gy, x0, x1; uint64_t lo, hi; float mr; } res_job_t;

static void *residual_rows(void *arg) {
    res_job_t *j = (res_job_t *) arg;
    const ctx_t *c = j->c;
    uint64_t lo = 0, hi = 0; float mr = 0.0f;
    for (int x = j->x0; x < j->x1; ++x)
        for (int y = 0; y < c->ey; ++y) {
            if (x == j->gx && y == j->gy) continue;
            float v = elem_update(j->kind, c, x, y), s = c->g[(size_t) x * c->ey + y];
            if (v == s) continue;
            if (v < s) ++lo; else ++hi;
            if (isfinite(v) && isfinite(s)) { float r = fabsf(v - s) / fabsf(s); if (r > mr) mr = r; }
            else mr = INFINITY;
        }
    j->lo = lo; j->hi = hi; j->mr = mr;
    return NULL;
}

/* rows are independent: blocks of rows go to up to 32 threads (the 16384^2 check of the config-size tests then takes
 * seconds); TASI_ORACLE_THREADS overrides, 1 = the plain loop */
void tasi_oracle_residual(int kind, const uint8_t *map, int W, int L, int thr, int gx, int gy, const float *g,
                          uint64_t *n_below, uint64_t *n_above, float *max_rel) {
    ctx_t c = {map, L, W, thr, kind == TASI_DFM ? L : L + 1, kind == TASI_DFM ? W : W + 1, g};
    long nt = sysconf(_SC_NPROCESSORS_ONLN);
    const char *e = getenv("TASI_ORACLE_THREADS");
    if (e) nt = atol(e);
    if (nt > 32) nt = 32;
    if (nt < 1 || (size_t) c.ex * c.ey < (1u << 20)) nt = 1;
    if (nt > c.ex) nt = c.ex;
    res_job_t jobs[32];
    pthread_t th[32];
    for (int t = 0; t < nt; ++t) {
        jobs[t] = (res_job_t){kind, &c, gx, gy, (int) ((int64_t) c.ex * t / nt), (int) ((int64_t) c.ex * (t + 1) / nt), 0, 0, 0.0f};
        if (nt == 1) residual_rows(&jobs[t]);
        else if (pthread_create(&th[t], NULL, residual_rows, &jobs[t]) != 0) { residual_rows(&jobs[t]); th[t] = 0; }
    }
    uint64_t lo = 0, hi = 0; float mr = 0.0f;
    for (int t = 0; t < nt; ++t) {
        if (nt > 1 && th[t]) pthread_join(th[t], NULL);
        lo += jobs[t].lo; hi += jobs[t].hi;
        if (jobs[t].mr > mr) mr = jobs[t].mr;
    }
    *n_below = lo; *n_above = hi; *max_rel = mr;
}

/* Graph::update (ProjectToolkit/Graph.cpp:36-51): write the patch (rows x..x+h-1, cols y..y+w-1) into
 * the map and return how many bytes changed. */
uint64_t tasi_oracle_patch(uint8_t *map, int W, int L, const uint8_t *patch, int x, int y, int w, int h) {
    (void) L;
    uint64_t changed = 0;
    for (int i = 0; i < h; ++i)
        for (int j = 0; j < w; ++j) {
            uint8_t *mv = &map[(size_t) (x + i) * W + (y + j)];
            uint8_t pv = patch[(size_t) i * w + j];
            if (*mv != pv) ++changed;
            *mv = pv;
        }
    return changed;
}
