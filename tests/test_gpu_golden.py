"""GPU parity against the committed reference vectors (tests/golden, made by scripts/make_golden.py from the
unmodified reference): value field on the reference's closed set, reachability, replan, and the extracted path.

Bars (BASELINE.json north_star): reachability identical on the closed set; FD*/MFD* values bit-exact, MS-DFM
within 1e-4 relative (observed <= 1e-7); path cost within 1e-3 relative and the path's cell sequence identical.
"""
import glob
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")
KIND = {"fd": 0, "dfm": 1, "mfd": 2}
FIELD_TOL = 1e-4
PATH_COST_TOL = 1e-3
CASES = sorted(os.path.basename(p) for p in glob.glob(os.path.join(GOLD, "field_*.npz")))


def case_kind(fname):
    return KIND[fname[:-4].rsplit("_", 1)[1]]


def check(kind, got, want, mask):
    assert np.isfinite(got[mask]).all(), "reference-closed elements unreached on the GPU"
    if kind == 1:
        rel = np.abs(got[mask] - want[mask]) / np.maximum(want[mask], 1e-30)
        assert rel.max(initial=0.0) <= FIELD_TOL
    else:
        assert np.array_equal(got[mask], want[mask])


def gpu_plan(kind, d):
    from unige_tasi_path_planners_b200 import PLANNERS
    p = PLANNERS[kind](PLANNERS[kind].max_level)
    p.set_occupancy_threshold(1.0)
    p.set_map(d["cmap"])
    p.set_start(*[float(v) for v in d["start"]])
    p.set_goal(*[float(v) for v in d["goal"]])
    assert p.step() == 0
    return p


@pytest.mark.parametrize("fname", CASES)
def test_field_and_replan_match_reference(fname):
    d = np.load(os.path.join(GOLD, fname))
    kind = case_kind(fname)
    p = gpu_plan(kind, d)
    got = p.read_field()
    check(kind, got, d["g"], d["mask"])
    # elements the reference left open must not be cheaper than its top key
    open_fin = np.isfinite(got) & ~d["mask"]
    if open_fin.any() and np.isfinite(d["top_key"]):
        assert got[open_fin].min() >= d["top_key"] * (1 - 1e-6)
    px, py = [int(v) for v in d["patch_xy"]]
    p.patch_map(d["patch"], px, py)
    p.set_start(*[float(v) for v in d["start2"]])
    assert p.step() == 0
    assert p.num_nodes_updated == int(d["updated2"])  # same seed set as the reference's update()
    check(kind, p.read_field(), d["g2"], d["mask2"])
    p.close()


def cells_of(path):
    """Cell sequence of a path: floor of consecutive midpoints, repeats collapsed."""
    mid = (path[:-1] + path[1:]) * 0.5
    c = np.floor(mid).astype(np.int64)
    keep = np.ones(len(c), bool)
    keep[1:] = (c[1:] != c[:-1]).any(axis=1)
    return c[keep]


@pytest.mark.parametrize("fname", CASES)
def test_path_through_reference_extractor(fname, oracle):
    """The unmodified LinearInterpolationPathExtractor run on the GPU field gives the reference's path."""
    if not oracle.ref_available():
        pytest.skip("oracle/_ref not present on this machine")
    d = np.load(os.path.join(GOLD, fname))
    kind = case_kind(fname)
    p = gpu_plan(kind, d)
    got = p.read_field()
    p.close()
    r = oracle.RefPlanner(kind, 1)
    r.set_map(d["cmap"], 1.0)
    r.set_start(*[float(v) for v in d["start"]])
    r.set_goal(*[float(v) for v in d["goal"]])
    xy, cost, dist, _ = r.extract(got, 20 if kind == 1 else 2000, True, kind != 2)
    r.close()
    want = d["path"]
    assert len(xy) == len(want)
    if len(want) > 1:
        assert np.array_equal(cells_of(xy), cells_of(want))
    wc = float(d["path_cost"])
    if np.isfinite(wc) and wc > 0:
        assert abs(cost - wc) / wc <= PATH_COST_TOL
        assert abs(dist - float(d["path_dist"])) / max(float(d["path_dist"]), 1e-30) <= PATH_COST_TOL
    if kind != 1:
        assert np.array_equal(xy, want)  # bit-identical field => identical extraction
