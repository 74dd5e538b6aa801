"""GPU parity: the CUDA solver (through the C ABI) against the CPU oracle on the same seeded inputs.

Bars (BASELINE.json north_star): reachability masks identical; FD*/MFD* fields bit-exact with the oracle
(their fixed point is order-independent); MS-DFM within 1e-4 relative (observed ~1e-7).
"""
import numpy as np
import pytest

import hostops

from unige_tasi_path_planners_b200 import synth

pytestmark = pytest.mark.gpu

KINDS = [0, 1, 2]
REL_TOL = 1e-4  # north_star: converged value field within 1e-4 relative


def make(kind, level=None):
    from unige_tasi_path_planners_b200 import PLANNERS
    cls = PLANNERS[kind]
    return cls(cls.max_level if level is None else level)


def compare(kind, got, want, exact=None):
    assert got.shape == want.shape
    fin_g, fin_w = np.isfinite(got), np.isfinite(want)
    assert np.array_equal(fin_g, fin_w), "reachability masks differ at %d elements" % (fin_g != fin_w).sum()
    if exact is None:
        exact = kind != 1
    if exact:
        bad = got[fin_w] != want[fin_w]
        assert not bad.any(), "%d of %d elements not bit-exact, max rel %g" % (
            bad.sum(), fin_w.sum(), np.max(np.abs(got[fin_w] - want[fin_w]) / np.maximum(want[fin_w], 1e-30)))
    else:
        rel = np.abs(got[fin_w] - want[fin_w]) / np.maximum(np.abs(want[fin_w]), 1e-30)
        assert rel.max(initial=0.0) <= REL_TOL, "max rel err %g" % rel.max()


def plan(kind, cmap, start, goal, occ=1.0):
    p = make(kind)
    p.set_occupancy_threshold(occ)
    p.set_map(cmap)
    p.set_start(*start)
    p.set_goal(*goal)
    assert p.step() == 0
    return p


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("shape", [(64, 64), (100, 37), (33, 200), (31, 31), (256, 256), (1, 1), (5, 70)])
def test_initial_plan_matches_oracle(oracle, kind, shape):
    L, W = shape
    start, goal = (min(5.0, L - 1), min(5.0, W - 1)), (float(max(L - 6, 0)), float(max(W - 6, 0)))
    cmap = synth.synth_map(11 + kind, L, W, 10, start, goal)
    p = plan(kind, cmap, start, goal)
    got = p.read_field()
    want = oracle.solve(kind, cmap, goal)
    compare(kind, got, want)
    lo, hi, _ = oracle.residual(kind, cmap, goal, got)
    assert lo == 0  # nothing can still be lowered: a true fixed point of the reference's operator
    if kind != 1:
        assert hi == 0


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("p_obs", [0, 30, 45])
def test_obstacle_density(oracle, kind, p_obs):
    L = W = 160
    start, goal = (5.0, 5.0), (80.0, 80.0)
    cmap = synth.synth_map(21, L, W, p_obs, start, goal)
    p = plan(kind, cmap, start, goal)
    compare(kind, p.read_field(), oracle.solve(kind, cmap, goal))


@pytest.mark.parametrize("kind", KINDS)
def test_occupancy_threshold_and_uniform(oracle, kind):
    L, W = 96, 130
    cmap = np.full((L, W), 7, np.uint8)
    cmap[40:44, 10:120] = 200  # a wall once the threshold is 0.5 (>= 127)
    start, goal = (3.0, 3.0), (90.0, 64.0)
    p = plan(kind, cmap, start, goal, occ=0.5)
    thr = int(0.5 * 255.0)
    compare(kind, p.read_field(), oracle.solve(kind, cmap, goal, thr=thr))


@pytest.mark.parametrize("kind", KINDS)
def test_unreachable_goal_pocket(oracle, kind):
    L = W = 80
    cmap = np.full((L, W), 3, np.uint8)
    cmap[30:51, 30] = 255; cmap[30:51, 50] = 255; cmap[30, 30:51] = 255; cmap[50, 30:51] = 255
    start, goal = (5.0, 5.0), (40.0, 40.0)
    p = plan(kind, cmap, start, goal)
    got = p.read_field()
    compare(kind, got, oracle.solve(kind, cmap, goal))
    assert not np.isfinite(got[5, 5])


@pytest.mark.parametrize("kind", KINDS)
def test_step_return_codes(kind):
    p = make(kind)
    assert p.step() == -1  # LOOP_FAILURE_NO_GRAPH
    p.set_map(np.ones((8, 8), np.uint8))
    assert p.step() == -2  # LOOP_FAILURE_NO_GOAL
    p.set_goal(4.0, 4.0)
    p.set_start(1.0, 1.0)
    assert p.step() == 0


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("mode", ["mixed", "raise", "lower"])
def test_replan_equals_from_scratch(oracle, kind, mode):
    """Two-phase repair after a patch == from-scratch solve of the patched map."""
    L = W = 200
    start, goal = (5.0, 5.0), (float(L - 6), float(W - 6))
    cmap = synth.synth_map(31, L, W, 10, start, goal)
    p = plan(kind, cmap, start, goal)
    rng = np.random.default_rng(5)
    x, y, h, w = 60, 70, 48, 40
    old = cmap[x:x + h, y:y + w]
    if mode == "mixed":
        patch = synth.synth_rows(977, 0, h, w, 15)
    elif mode == "raise":
        patch = np.where(rng.random((h, w)) < 0.3, 255, np.minimum(old.astype(int) + 9, 255)).astype(np.uint8)
    else:
        patch = np.maximum(old.astype(int) // 2, 1).astype(np.uint8)
    n_changed = p.patch_map(patch, x, y)
    cm2 = cmap.copy()
    assert n_changed == oracle.patch(cm2, patch, x, y)
    assert np.array_equal(p.read_map(), cm2)
    p.set_start(8.0, 8.0)
    assert p.step() == 0
    assert p.num_nodes_updated > 0
    compare(kind, p.read_field(), oracle.solve(kind, cm2, goal))
    # a second, disjoint patch on top
    patch2 = synth.synth_rows(1234, 0, 20, 33, 25)
    p.patch_map(patch2, 150, 20)
    oracle.patch(cm2, patch2, 150, 20)
    p.set_start(9.0, 9.0)
    assert p.step() == 0
    compare(kind, p.read_field(), oracle.solve(kind, cm2, goal))


@pytest.mark.parametrize("kind", KINDS)
def test_new_goal_reinitialises(oracle, kind):
    L = W = 120
    cmap = synth.synth_map(41, L, W, 10, (5, 5), (100, 100))
    p = plan(kind, cmap, (5.0, 5.0), (100.0, 100.0))
    p.set_goal(20.0, 90.0)
    assert p.step() == 0
    compare(kind, p.read_field(), oracle.solve(kind, cmap, (20.0, 90.0)))
    f1 = p.read_field()
    p.set_goal(20.2, 89.9)  # same element: no re-init, nothing to do
    assert p.step() == 0
    assert np.array_equal(f1, p.read_field())


@pytest.mark.parametrize("kind", KINDS)
def test_window_read(kind):
    L, W = 70, 90
    cmap = synth.synth_map(51, L, W, 5, (5, 5), (60, 80))
    p = plan(kind, cmap, (5.0, 5.0), (60.0, 80.0))
    f = p.read_field()
    w = p.read_window(-3, 80, 10, 20)
    ex, ey = f.shape
    ref = np.full((10, 20), np.inf, np.float32)
    ref[3:, :ey - 80] = f[0:7, 80:ey]
    assert np.array_equal(w, ref)


def test_device_synth_map_matches_host():
    from unige_tasi_path_planners_b200 import DFMPlanner
    p = DFMPlanner(1)
    p.synth_map(77, 300, 130, row0=0, p_obs=10, start=(5, 5), goal=(120, 290))
    host = synth.synth_map(77, 130, 300, 10, (5, 5), (120, 290))
    assert np.array_equal(p.read_map(), host)


@pytest.mark.parametrize("kind", KINDS)
def test_concurrent_planners_with_capped_ctas(oracle, kind):
    """Independent handles on their own streams, each with a share of the persistent CTAs (the query-batch mode):
    four host threads planning different maps at once give the oracle's fields."""
    import threading
    L = W = 192
    maps = [synth.synth_map(300 + i, L, W, 10, (5, 5), (150 + i, 160 - i)) for i in range(4)]
    out = [None] * 4

    def work(i):
        p = make(kind)
        p.set_solver_blocks(16)
        p.set_occupancy_threshold(1.0)
        for rep in range(3):
            p.set_map(maps[i])
            p.set_start(5.0, 5.0)
            p.set_goal(150.0 + i, 160.0 - i)
            assert p.step() == 0
        out[i] = p.read_field()
        p.close()

    th = [threading.Thread(target=work, args=(i,)) for i in range(4)]
    [t.start() for t in th]
    [t.join() for t in th]
    for i in range(4):
        compare(kind, out[i], oracle.solve(kind, maps[i], (150.0 + i, 160.0 - i)))


@pytest.mark.parametrize("kind", KINDS)
def test_stacked_batch_equals_single_plans(oracle, kind):
    """A batch of independent queries solved as one multi-goal plan on the stacked map (obstacle row between maps):
    every block is the field a planner holding only that map returns."""
    from unige_tasi_path_planners_b200 import batch
    L, W = 70, 100
    maps = [synth.synth_map(400 + i, L, W, 10, None, None) for i in range(3)]
    maps.append(maps[1])  # a second query on an already used map
    goals = [(60.0, 90.0), (5.0, 50.0), (35.0, 3.0), (66.0, 8.0)]
    for m, g in zip(maps, goals):
        synth.clear_block(m, int(g[0]), int(g[1]))
    sb = batch.StackedBatch(kind, device=0)
    sb.plan(maps, goals)
    for i, (m, g) in enumerate(zip(maps, goals)):
        compare(kind, sb.field(i), oracle.solve(kind, m, g))
    w = sb.window(2, -5, 10, 20, 30)
    f = sb.field(2)
    assert np.all(np.isinf(w[:5])) and np.array_equal(w[5:], f[0:15, 10:40])
    # one walk per query in one launch == the walk over each block's own field (rows at the seams included)
    from unige_tasi_path_planners_b200 import _lib
    starts = [(0.0, 0.0), (69.0, 99.0), (70.0, 50.0), (2.5, 60.0)]
    for m, s in zip(maps, starts):
        synth.clear_block(m, int(s[0]), int(s[1]))
    sb.plan(maps, goals)
    walks, _ = sb.paths(starts, 2000, with_costs=True)
    for i, (m, s, g) in enumerate(zip(maps, starts, goals)):
        hw = hostops.host_extract(kind == 1, sb.field(i), m, s, g, 255, 2000)
        assert np.array_equal(walks[i]["path"], hw[0]) and np.array_equal(walks[i]["costs"], hw[1]), i
        assert walks[i]["status"] == hw[4] and (kind == 1 or hw[4] == 0)
        if oracle.ref_available() and kind != 1:  # the reference's own extractor over the same block (oracle/_ref)
            r = oracle.RefPlanner(kind, 1)
            r.set_map(m)
            r.set_start(*s)
            r.set_goal(*g)
            xy, cost, _, _ = r.extract(field=sb.field(i), max_steps=2000)
            r.close()
            assert np.array_equal(xy, walks[i]["path"]), i
    # a replan on the stack: patch map 1's block, every block still equals its own from-scratch plan
    patch = synth.synth_rows(77, 0, 20, 30, 20)
    sb.p.patch_map(patch, 1 * (L + 1) + 10, 40)
    sb.p.set_start(1.0, 1.0)
    assert sb.p.step() == 0
    m1 = maps[1].copy()
    oracle.patch(m1, patch, 10, 40)
    compare(kind, sb.field(1), oracle.solve(kind, m1, goals[1]))
    compare(kind, sb.field(0), oracle.solve(kind, maps[0], goals[0]))
    sb.close()


@pytest.mark.parametrize("kind", KINDS)
def test_full_size_cfg2_field_is_a_fixed_point(oracle, kind):
    """BASELINE.json's 4096 x 4096 workload: the oracle cannot solve it in seconds, but one pass of its operator over
    the GPU field can: no element may be lowerable (residual lo == 0), for FD*/MFD* none may be raisable either
    (a true fixed point of the reference's operator), and the reachability mask must match a flood fill."""
    from scipy import ndimage
    N = 4096
    start, goal = synth.default_endpoints(N, N)
    p = make(kind)
    p.set_occupancy_threshold(1.0)
    p.synth_map(1, N, N, 0, 10, start, goal)
    p.set_start(*start)
    p.set_goal(*goal)
    assert p.step() == 0
    got = p.read_field()
    cmap = p.read_map()
    assert np.array_equal(cmap, synth.synth_map(1, N, N, 10, start, goal))
    lo, hi, maxrel = oracle.residual(kind, cmap, goal, got)
    assert lo == 0
    if kind != 1:
        assert hi == 0
    else:
        assert maxrel <= 1e-6
    # reachability: free cells 8-connected to the goal cell (MS-DFM values live on cells)
    if kind == 1:
        lab, _ = ndimage.label(cmap < 255, structure=np.ones((3, 3), int))
        reach = lab == lab[int(goal[0]), int(goal[1])]
        assert np.array_equal(np.isfinite(got), reach)
    assert np.isfinite(got[int(start[0]), int(start[1])])


def compare_settled(kind, got, want, settled_below):
    """Focused plans (tasi_set_focus): values below settled_below are final, the others upper bounds."""
    m = want < settled_below
    assert np.array_equal(np.isfinite(got[m]), np.isfinite(want[m]))
    if kind == 1:
        assert (np.abs(got[m] - want[m]) / np.maximum(want[m], 1e-30)).max(initial=0.0) <= REL_TOL
    else:
        assert np.array_equal(got[m], want[m])
    fin = np.isfinite(want)
    assert np.all(got[fin] >= want[fin] * (1 - 1e-5 if kind == 1 else 1.0)) and np.all(np.isinf(got[~fin]))


@pytest.mark.parametrize("focus", [False, True])
@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("seed", [0, 1, 2])
def test_random_patch_sequences(oracle, kind, seed, focus):
    """A drive of random events — patches anywhere (borders, overlapping, two before one step, the whole map, walls
    that cut the goal off and patches that reopen it), moved starts, a changed goal — after every step the field is
    the from-scratch field of the map as it then is."""
    import random
    rng = random.Random(1000 * kind + seed)
    L, W = rng.choice([(97, 130), (160, 75), (129, 129)])
    start, goal = (3.0, 4.0), (float(L - 5), float(W - 7))
    cmap = synth.synth_map(50 + seed, L, W, rng.choice([0, 10, 20]), start, goal)
    p = plan(kind, cmap, start, goal)
    compare(kind, p.read_field(), oracle.solve(kind, cmap, goal))
    p.set_focus(focus, 0.0)  # focused: every later step stops once its start is settled and parks the rest
    for step in range(9):
        for _ in range(rng.choice([1, 1, 2])):
            what = rng.randrange(6)
            if what == 0:  # the whole map re-drawn
                x = y = 0
                patch = synth.synth_rows(900 + step, 0, L, W, rng.choice([5, 15]))
            elif what == 1:  # a wall across the map (may cut the goal off), with or without a door
                x, y = rng.randrange(10, L - 10), 0
                patch = np.full((2, W), 255, np.uint8)
                if rng.random() < 0.5:
                    patch[:, rng.randrange(W - 3):][:, :3] = 7
            elif what == 2:  # everything cheap again in a big block (reopens, lowers)
                h, w = rng.randrange(20, L), rng.randrange(20, W)
                x, y = rng.randrange(L - h + 1), rng.randrange(W - w + 1)
                patch = np.full((h, w), rng.randrange(1, 9), np.uint8)
            else:  # a random block, often touching a border
                h, w = rng.randrange(1, 50), rng.randrange(1, 50)
                x = rng.choice([0, L - h, rng.randrange(L - h + 1)])
                y = rng.choice([0, W - w, rng.randrange(W - w + 1)])
                patch = synth.synth_rows(7000 + 31 * step + what, 0, h, w, rng.choice([0, 20, 60]))
            p.patch_map(patch, x, y)
            oracle.patch(cmap, patch, x, y)
        if step == 5:
            goal = (float(rng.randrange(5, L - 5)), float(rng.randrange(5, W - 5)))
            p.set_goal(*goal)
        p.set_start(float(rng.randrange(1, L - 1)), float(rng.randrange(1, W - 1)))
        assert p.step() == 0
        assert np.array_equal(p.read_map(), cmap)
        if focus:
            compare_settled(kind, p.read_field(), oracle.solve(kind, cmap, goal), p.stats()["settled_below"])
        else:
            compare(kind, p.read_field(), oracle.solve(kind, cmap, goal))
    p.set_focus(False)
    p.set_start(1.0, 1.0)
    assert p.step() == 0  # whatever was parked is finished now
    compare(kind, p.read_field(), oracle.solve(kind, cmap, goal))
    p.close()
