"""Parity at the sizes BASELINE.json names ("What's weak" #1).

* 1024^2 (the cfg5 map size), all three planners, LIVE against the unmodified reference planners (oracle/_ref) on the
  reference's closed set {g == rhs < INF and g <= top_key}: FD* / MFD* bit-exact, MS-DFM 1e-4 relative.
* cfg2 — 4096^2, seed 1 — all three planners against the digest of the reference's own plan committed under
  tests/golden/cfg2_4096_*.npz (scripts/make_golden_cfg_sizes.py): FD* / MFD* row CRCs of the closed set, MS-DFM sampled
  rows and row sums within 1e-4.
* cfg3 — MFD* 16384^2 — after the plan and after the 1 % replan: one application of the oracle's operator to the whole
  field changes nothing (lo == hi == 0: a fixed point of the reference's operator, which is unique), and the repaired
  field equals a from-scratch plan.
* cfg5 — the query generator draws starts and goals from free cells, every MS-DFM query reaches its start.
"""
import os
import zlib

import numpy as np
import pytest

from unige_tasi_path_planners_b200 import synth

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
NAMES = {0: "fd", 1: "dfm", 2: "mfd"}
REL_TOL = 1e-4  # north_star: converged value field within 1e-4 relative (MS-DFM; FD* / MFD* are compared bit for bit)


def make(kind, level=None):
    from unige_tasi_path_planners_b200 import PLANNERS
    cls = PLANNERS[kind]
    p = cls(cls.max_level if level is None else level)
    p.set_occupancy_threshold(1.0)
    return p


@pytest.mark.gpu
@pytest.mark.parametrize("kind", [0, 1, 2])
def test_1024_live_against_the_reference(oracle, kind):
    if not oracle.ref_available():
        pytest.skip("oracle/_ref not built (needs /root/reference at build time)")
    side = 1024
    start, goal = synth.default_endpoints(side, side)
    cmap = synth.synth_map(1000 + kind, side, side, 10, start, goal)
    r = oracle.RefPlanner(kind, 2 if kind == 2 else 1)
    r.set_map(cmap, 1.0)
    r.set_start(*start)
    r.set_goal(*goal)
    assert r.step() == 0
    closed, g_ref = r.closed_mask()
    r.close()
    assert closed.sum() > 0.85 * closed.size  # the far-corner start makes the reference expand (nearly) everything
    p = make(kind)
    p.set_map(cmap)
    p.set_start(*start)
    p.set_goal(*goal)
    assert p.step() == 0
    got = p.read_field()
    p.close()
    assert np.all(np.isfinite(got[closed])), "an element the reference reached is INF on the GPU"
    if kind == 1:
        rel = np.abs(got[closed] - g_ref[closed]) / np.maximum(g_ref[closed], 1e-30)
        assert rel.max() <= REL_TOL, rel.max()
    else:
        bad = got[closed] != g_ref[closed]
        assert not bad.any(), "%d of %d closed elements differ from the reference" % (bad.sum(), closed.sum())
    # obstacle-enclosed / unreachable elements: INF in the reference's full-grid view must be INF here too
    lo, hi, _ = oracle.residual(kind, cmap, goal, got)
    assert lo == 0 and (kind == 1 or hi == 0)


def check_against_digest(kind, got, d):
    """`got`: a converged field of the cfg2 grid; `d`: the digest of the reference's plan (make_golden_cfg_sizes.py)."""
    top = np.float32(d["top_key"])
    rows, want = d["rows"], d["row_values"]
    closed_rows = ~np.isnan(want)
    g_rows = got[rows]
    assert np.all(np.isfinite(g_rows[closed_rows]))
    mask = got < top  # the reference's closed set, rebuilt from our own field
    if kind == 1:
        rel = np.abs(g_rows[closed_rows] - want[closed_rows]) / np.maximum(want[closed_rows], 1e-30)
        assert rel.max() <= REL_TOL, rel.max()
        # an ulp of difference may move an element across top_key: allow a handful per row
        assert np.abs(mask.sum(axis=1) - d["closed_per_row"]).max() <= 4
        rs = np.where(mask, got, 0).astype(np.float64).sum(axis=1)
        ref = d["row_sum"]
        assert np.all(np.abs(rs - ref) <= REL_TOL * np.maximum(ref, 1.0) + 4 * float(top))
    else:
        assert np.array_equal(g_rows[closed_rows], want[closed_rows])
        assert np.array_equal(mask.sum(axis=1).astype(np.int32), d["closed_per_row"])
        crc = np.array([zlib.crc32(np.ascontiguousarray(got[i][mask[i]]).tobytes()) for i in range(got.shape[0])],
                       np.uint32)
        bad = np.nonzero(crc != d["row_crc"])[0]
        assert bad.size == 0, "rows whose closed set differs from the reference's: %s" % bad[:8]


def load_digest(kind):
    d = np.load(os.path.join(GOLDEN, "cfg2_4096_%s.npz" % NAMES[kind]))
    side, seed = int(d["side"]), int(d["seed"])
    start, goal = synth.default_endpoints(side, side)
    return d, side, seed, start, goal


@pytest.mark.gpu
@pytest.mark.parametrize("kind", [0, 1, 2])
def test_cfg2_4096_against_the_reference_digest(kind):
    d, side, seed, start, goal = load_digest(kind)
    p = make(kind)
    p.synth_map(seed, side, side, 0, 10, start, goal)
    p.set_start(*start)
    p.set_goal(*goal)
    assert p.step() == 0
    got = p.read_field()
    p.close()
    check_against_digest(kind, got, d)


def test_oracle_cfg2_4096_against_the_reference_digest(oracle):
    """The oracle itself at config size: its heap solver on the cfg2 grid against the reference's digest (the three
    planners side by side in threads: the C calls drop the GIL)."""
    from concurrent.futures import ThreadPoolExecutor
    d0, side, seed, start, goal = load_digest(0)
    cmap = synth.synth_map(seed, side, side, 10, start, goal)
    with ThreadPoolExecutor(3) as ex:
        fields = list(ex.map(lambda k: oracle.solve(k, cmap, goal), [0, 1, 2]))
    for kind in (0, 1, 2):
        check_against_digest(kind, fields[kind], load_digest(kind)[0])


@pytest.mark.gpu
def test_cfg3_16384_plan_and_replan_are_fixed_points_of_the_reference_operator(oracle):
    import torch
    if torch.cuda.mem_get_info()[0] < 8 << 30:
        pytest.skip("needs 8 GB of device memory")
    side, kind = 16384, 2
    start, goal = synth.default_endpoints(side, side)
    p = make(kind, 2)
    p.synth_map(2, side, side, 0, 10, start, goal)
    p.set_start(*start)
    p.set_goal(*goal)
    assert p.step() == 0
    cmap = p.read_map()
    got = p.read_field()
    assert np.isfinite(got[5, 5])
    lo, hi, _ = oracle.residual(kind, cmap, goal, got)
    assert lo == 0 and hi == 0, (lo, hi)
    # the cfg3 replan: 1 % of the cells around the rover re-drawn (raises and lowers mixed)
    ps = int(0.1 * side)
    rover = (float(start[0] + 6), float(start[1] + 6))
    x0 = int(min(max(rover[0] - ps // 2, 0), side - ps))
    y0 = int(min(max(rover[1] - ps // 2, 0), side - ps))
    patch = synth.synth_rows(2 + 7777, 0, ps, ps, 5)
    patch[:12, :12] = np.minimum(patch[:12, :12], 64)
    changed = p.patch_map(patch, x0, y0)
    assert changed > 0.9 * ps * ps
    p.set_start(*rover)
    assert p.step() == 0
    assert p.stats()["elem_invalidated"] > 0
    cmap[x0:x0 + ps, y0:y0 + ps] = patch
    got = p.read_field()
    p.close()
    lo, hi, _ = oracle.residual(kind, cmap, goal, got)
    assert lo == 0 and hi == 0, (lo, hi)


def test_cfg5_queries_are_drawn_from_free_cells():
    from unige_tasi_path_planners_b200 import batch
    side = 256
    maps = {m: synth.synth_map(1000 + m, side, side, 10, None, None) for m in range(3)}
    qs = batch.make_queries(3, 16, side, maps=maps)
    assert len(qs) == 48
    for m, s, g in qs:
        assert maps[m][int(s[0]), int(s[1])] != 255 and maps[m][int(g[0]), int(g[1])] != 255
    assert qs == batch.make_queries(3, 16, side, maps=maps)  # deterministic (seed 42)


@pytest.mark.gpu
def test_cfg5_every_query_reaches_its_start():
    from unige_tasi_path_planners_b200 import batch, planner
    side = 256
    maps = {m: synth.synth_map(1000 + m, side, side, 10, None, None) for m in range(4)}
    qs = batch.make_queries(4, 8, side, maps=maps)
    for kind in (planner.DFM, planner.FD):
        sb = batch.StackedBatch(kind, 1, 0)
        sb.plan([maps[q[0]] for q in qs], [q[2] for q in qs])
        reached = 0
        for i, q in enumerate(qs):
            w = sb.window(i, int(q[1][0]), int(q[1][1]), 1, 1)
            reached += bool(np.isfinite(w[0, 0]))
        sb.close()
        # free start and goal cells on a 10 %-obstacle map: connected with overwhelming probability
        assert reached >= len(qs) - 1, (kind, reached)
