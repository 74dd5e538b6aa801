"""Path extraction: csrc/tasi_extract.cuh against the reference's LinearInterpolationPathExtractor.

CPU tests run the header compiled for the host (libtasi_hostops.so) against the committed golden walks
(tests/golden/paths.npz, scripts/make_golden_paths.py) and, when oracle/_ref is present, against the reference
extractor on fresh seeded fields. GPU tests run the warp-per-walk kernel through the C ABI against the same two
checkers. Bar: identical positions, piece costs, total cost and length (float32, bit for bit).
"""
import os
import random

import numpy as np
import pytest

import hostops

from unige_tasi_path_planners_b200 import _lib, synth

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "paths.npz")


def golden_cases():
    d = np.load(GOLDEN)
    for i in range(int(d["n"][0])):
        k = "c%02d_" % i
        kind, ms, la, ind = (int(v) for v in d[k + "meta"])
        e = d[k + "ends"]
        yield dict(i=i, kind=kind, max_steps=ms, lookahead=bool(la), indirect=bool(ind), cmap=d[k + "map"],
                   field=d[k + "field"], start=(float(e[0]), float(e[1])), goal=(float(e[2]), float(e[3])),
                   path=d[k + "path"], costs=d[k + "costs"], totals=d[k + "totals"])


def same_walk(c, xy, sc, cost, dist, what):
    assert xy.shape == c["path"].shape, "%s: %d points, reference %d" % (what, len(xy), len(c["path"]))
    assert np.array_equal(xy, c["path"]), what
    assert np.array_equal(sc, c["costs"]), what
    assert np.float32(cost) == c["totals"][0] and np.float32(dist) == c["totals"][1], what


def test_golden_file_covers_all_planners_and_options():
    cs = list(golden_cases())
    assert {c["kind"] for c in cs} == {0, 1, 2}
    assert {c["lookahead"] for c in cs} == {True, False} and {c["indirect"] for c in cs} == {True, False}
    assert any(c["start"][0] != int(c["start"][0]) or c["start"][1] != int(c["start"][1]) for c in cs)
    assert any(len(c["path"]) > 1 and tuple(c["path"][-1]) == c["goal"] for c in cs)


def test_host_walk_equals_golden():
    for c in golden_cases():
        xy, sc, cost, dist, rc = hostops.host_extract(c["kind"] == 1, c["field"], c["cmap"], c["start"], c["goal"], 255,
                                                   c["max_steps"], c["lookahead"], c["indirect"])
        same_walk(c, xy, sc, cost, dist, "golden case %d" % c["i"])
        reached = tuple(xy[-1]) == c["goal"]
        assert rc == (0 if reached else 1)


def fresh_cases(n_seeds, base=0):
    for kind in (0, 1, 2):
        for seed in range(base, base + n_seeds):
            rng = random.Random(77 * seed + kind)
            L, W = rng.choice([24, 40, 64]), rng.choice([24, 33, 64])
            start = (float(rng.randrange(2, L - 2)), float(rng.randrange(2, W - 2)))
            goal = (float(rng.randrange(2, L - 2)), float(rng.randrange(2, W - 2)))
            cmap = synth.synth_map(seed, L, W, rng.choice([0, 10, 25, 40]), start, goal)
            m = rng.randrange(3)
            if m == 1:
                start = (start[0] + rng.choice([0.5, 0.25, 0.7313]), start[1])
            if m == 2:
                start = (start[0], start[1] + rng.choice([0.5, 0.125, 0.3331]))
            yield kind, seed, cmap, start, goal


def test_host_walk_equals_reference_extractor(oracle):
    if not oracle.ref_available():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    n = 0
    for kind, seed, cmap, start, goal in fresh_cases(12):
        field = oracle.solve(kind, cmap, goal)
        r = oracle.RefPlanner(kind, 1)
        r.set_map(cmap)
        r.set_start(*start)
        r.set_goal(*goal)
        for la in (True, False):
            for ind in (True, False):
                ms = 20 if (seed + la + ind) % 2 else 500
                xy, cost, dist, _ = r.extract(field=field, max_steps=ms, lookahead=la, allow_indirect=ind)
                c = dict(path=xy, costs=r.step_costs(), totals=np.array([cost, dist], np.float32))
                got = hostops.host_extract(kind == 1, field, cmap, start, goal, 255, ms, la, ind)
                same_walk(c, *got[:4], "kind %d seed %d lookahead %d indirect %d" % (kind, seed, la, ind))
                n += 1
        r.close()
    assert n == 144


def test_walk_respects_the_occupancy_threshold_and_capacity(oracle):
    cmap = np.full((16, 16), 10, np.uint8)
    cmap[:, 8] = 200  # a wall once the threshold drops below 200
    field = np.full((17, 17), np.inf, np.float32)
    field[:, 9:] = 10.0 * np.abs(np.arange(9, 17) - 12)[None, :] + 10.0 * np.abs(np.arange(17) - 8)[:, None]
    xy, _, _, _, rc = hostops.host_extract(False, field, cmap, (8.0, 4.0), (8.0, 12.0), thr=128)
    assert rc == 2 and len(xy) == 1  # every node around the start is unreached: stuck
    field = oracle.solve(0, cmap, (8.0, 12.0))
    xy, _, _, _, rc = hostops.host_extract(False, field, cmap, (8.0, 2.0), (8.0, 12.0), thr=255, cap=4)
    assert rc == 3 and len(xy) <= 4
    xy, _, _, _, rc = hostops.host_extract(False, field, cmap, (8.0, 2.0), (8.0, 12.0), thr=255)
    assert rc == 0 and tuple(xy[-1]) == (8.0, 12.0)


# ---------------------------------------------------------------- GPU: the warp-per-walk kernel through the C ABI

def gpu_planner(kind, cmap, start, goal):
    from unige_tasi_path_planners_b200 import PLANNERS
    p = PLANNERS[kind](1)
    p.set_occupancy_threshold(1.0)
    p.set_map(cmap)
    p.set_start(*start)
    p.set_goal(*goal)
    assert p.step() == 0
    return p


@pytest.mark.gpu
def test_gpu_walk_equals_host_walk_and_reference(oracle):
    have_ref = oracle.ref_available()
    for kind, seed, cmap, start, goal in fresh_cases(6, base=100):
        p = gpu_planner(kind, cmap, start, goal)
        field = p.read_field()
        r = None
        if have_ref:
            r = oracle.RefPlanner(kind, 1)
            r.set_map(cmap)
            r.set_start(*start)
            r.set_goal(*goal)
        for la in (True, False):
            for ind in (True, False):
                ms = 20 if (seed + la + ind) % 2 else 500
                res, _ = p.extract_paths(None, None, ms, la, ind)
                g = res[0]
                what = "kind %d seed %d lookahead %d indirect %d" % (kind, seed, la, ind)
                h = hostops.host_extract(kind == 1, field, cmap, start, goal, 255, ms, la, ind)
                c = dict(path=h[0], costs=h[1], totals=np.array(h[2:4], np.float32))
                same_walk(c, g["path"], g["costs"], g["total_cost"], g["total_dist"], what + " (host walk)")
                assert g["status"] == h[4]
                if r is not None:
                    xy, cost, dist, _ = r.extract(field=field, max_steps=ms, lookahead=la, allow_indirect=ind)
                    c = dict(path=xy, costs=r.step_costs(), totals=np.array([cost, dist], np.float32))
                    same_walk(c, g["path"], g["costs"], g["total_cost"], g["total_dist"], what + " (reference)")
        if r is not None:
            r.close()
        p.close()


@pytest.mark.gpu
def test_gpu_extractor_class_mirrors_the_reference_members():
    from unige_tasi_path_planners_b200 import planner as P
    start, goal = (50.0, 50.0), (6.0, 9.0)
    cmap = synth.synth_map(5, 64, 64, 10, start, goal)
    p = gpu_planner(0, cmap, start, goal)
    ex = P.PathExtractor(p)
    assert (ex.max_steps, ex.lookahead, ex.allow_indirect_traversals) == (20, True, True)
    ex.extract_path()
    assert ex.status == 1 and len(ex.path_) >= 21 and tuple(ex.path_[0]) == start
    assert ex.total_cost == pytest.approx(float(np.sum(ex.cost_)), rel=1e-5) and ex.e_time > 0
    ex.max_steps = 10000
    ex.extract_path()
    assert ex.status == 0 and tuple(ex.path_[-1]) == goal
    # the walk re-integrates the cost along the path: close to the field value at the start, not equal to it
    assert ex.total_cost == pytest.approx(float(p.read_window(50, 50, 1, 1)[0, 0]), rel=1e-2)
    p.close()


@pytest.mark.gpu
def test_gpu_batch_of_walks_equals_single_walks(oracle):
    start, goal = (100.0, 90.0), (10.0, 12.0)
    cmap = synth.synth_map(9, 128, 128, 10, start, goal)
    p = gpu_planner(2, cmap, start, goal)
    rng = random.Random(3)
    starts = [(float(rng.randrange(4, 124)), float(rng.randrange(4, 124))) for _ in range(37)]
    goals = [goal] * len(starts)
    many, _ = p.extract_paths(starts, goals, 4000)
    field = p.read_field()
    for s, m in zip(starts, many):
        one = hostops.host_extract(False, field, cmap, s, goal, 255, 4000)
        assert np.array_equal(m["path"], one[0]) and np.array_equal(m["costs"], one[1])
        assert m["status"] == one[4] and np.float32(m["total_cost"]) == np.float32(one[2])
    assert sum(m["status"] == 0 for m in many) >= 30
    # ... and against the reference's own extractor (oracle/_ref, where it was built) for a few of them: the host build
    # of the product's walk above is a stand-in for it, test_extract pins the two to each other
    if oracle.ref_available():
        r = oracle.RefPlanner(2, 1)
        r.set_map(cmap)
        r.set_goal(*goal)
        for s, m in list(zip(starts, many))[:6]:
            r.set_start(*s)
            xy, cost, dist, _ = r.extract(field=field, max_steps=4000)
            assert np.array_equal(xy, m["path"]) and np.float32(cost) == np.float32(m["total_cost"])
        r.close()
    p.close()


@pytest.mark.gpu
def test_gpu_walk_argument_errors():
    from unige_tasi_path_planners_b200 import PLANNERS
    p = PLANNERS[0](1)
    p.set_occupancy_threshold(1.0)
    p.set_map(np.full((16, 16), 5, np.uint8))
    p.set_start(2.0, 2.0)
    p.set_goal(12.0, 12.0)
    with pytest.raises(Exception):
        p.extract_paths()  # no plan yet
    assert p.step() == 0
    with pytest.raises(Exception):
        p.extract_paths([(99.0, 2.0)], [(12.0, 12.0)])  # start outside the map
    res, _ = p.extract_paths([(2.0, 2.0)], [(12.0, 12.0)], 100, cap_points=3)
    assert res[0]["status"] == 3
    p.close()


@pytest.mark.gpu
@pytest.mark.parametrize("kind", [0, 2])
def test_gpu_full_size_walk_descends_to_the_goal(kind):
    """BASELINE.json's 4096 x 4096 grid: no oracle at this size, but a walk has size-independent properties — it ends
    on the goal, the field never rises along its vertices, and its cost re-integrates the field value at the start."""
    from unige_tasi_path_planners_b200 import PLANNERS
    N = 4096
    start, goal = synth.default_endpoints(N, N)
    p = PLANNERS[kind](1)
    p.set_occupancy_threshold(1.0)
    p.synth_map(1, N, N, 0, 10, start, goal)
    p.set_start(*start)
    p.set_goal(*goal)
    assert p.step() == 0
    res, ms = p.extract_paths(None, None, 40000, cap_points=120001)
    w = res[0]
    assert w["status"] == 0 and tuple(w["path"][-1]) == goal and tuple(w["path"][0]) == start
    assert len(w["path"]) > N  # at least one point per row crossed
    g0 = float(p.read_window(int(start[0]), int(start[1]), 1, 1)[0, 0])
    assert w["total_cost"] == pytest.approx(g0, rel=2e-2)
    assert w["total_dist"] >= np.hypot(goal[0] - start[0], goal[1] - start[1]) * 0.999
    # field values at the path's vertices (positions with integer coordinates) never rise
    pts = w["path"]
    vert = (pts == np.floor(pts)).all(axis=1)
    vs = pts[vert].astype(np.int64)
    step = max(1, len(vs) // 400)
    vals = [float(p.read_window(int(x), int(y), 1, 1)[0, 0]) for x, y in vs[::step]]
    assert all(b <= a for a, b in zip(vals, vals[1:]))
    p.close()
