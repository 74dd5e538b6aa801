"""BASELINE.json configs[0] — Field D* level 1, initial plan + replans on the Tests/Tests maps through the demo process
(headless restatement of Tests/run_test.py, see tests/cfg1_drive.py). The reference's side of the comparison is its own
process, oracle/_ref/field_d_planner_1_no_heur (the unmodified Tests/Planners/FDSTAR/main.cpp), whose drive is committed
as tests/golden/cfg1_trace.npz and re-run live where the binary exists.

CPU: the fixture loads; the reference process reproduces the committed trace; the ORACLE's converged field of the map
known at every step carries the reference's closed set bit for bit (pins the oracle at cfg1).
GPU: host/fifo_demo (the GPU planners behind the same wire protocol) drives the same maps: reported positions, paths,
piece costs, totals and every value of the reference's closed set are bit-identical at every step.
"""
import os

import numpy as np
import pytest

import cfg1_drive as D
from unige_tasi_path_planners_b200 import build

MAPS = ["noise-trap", "square", "wall-a", "wall-b"]


def test_fixture_has_the_four_maps():
    m = D.load_maps()
    assert sorted(m) == MAPS
    assert m["noise-trap"][0].shape == (100, 100) and m["noise-trap"][1] == (90.0, 90.0) and m["noise-trap"][2] == (25.0, 25.0)
    t = D.load_trace()
    assert sorted(t) == MAPS and all(len(t[k]) >= 6 for k in MAPS)


@pytest.mark.parametrize("name", MAPS)
def test_oracle_field_carries_the_reference_closed_set(oracle, name):
    img, start, goal = D.load_maps()[name]
    trace = D.load_trace()[name]
    beliefs = D.replay_beliefs(img, [s["pos"] for s in trace])
    rep = D.compare_with_trace(trace, [None] * len(trace), lambda i: oracle.solve(oracle.FD, beliefs[i], goal))
    assert rep["closed_values_compared"] > 0


@pytest.mark.skipif(not os.path.exists(D.REF_DEMO), reason="oracle/_ref/field_d_planner_1_no_heur not built")
def test_reference_process_reproduces_the_committed_trace(tmp_path):
    name = "wall-b"
    img, start, goal = D.load_maps()[name]
    dr = D.drive(D.ref_cmd(name, start, goal), img, str(tmp_path), "ref")
    assert dr.ended and dr.returncode == 0
    shape = (img.shape[0] + 1, img.shape[1] + 1)

    def ref_field(i):  # dense field of the reference's own records (closed or not: only the closed set is read)
        f = np.full(shape, np.inf, np.float32)
        r = dr.steps[i]["recs"]
        f[r["x"], r["y"]] = r["g"]
        return f
    D.compare_with_trace(D.load_trace()[name], dr.steps, ref_field)


@pytest.fixture(scope="module")
def exe():
    build.build()
    return build.build_fifo_demo()


@pytest.mark.gpu
@pytest.mark.parametrize("name", MAPS)
def test_gpu_process_drives_like_the_reference_process(exe, name, tmp_path):
    img, start, goal = D.load_maps()[name]
    shape = (img.shape[0] + 1, img.shape[1] + 1)
    g = D.drive(D.gpu_cmd(exe, start, goal), img, str(tmp_path), "gpu")
    assert g.ended and g.returncode == 0, g.stderr
    rep = D.compare_with_trace(D.load_trace()[name], g.steps, lambda i: D.field_of(g.steps[i], shape))
    assert rep["closed_values_compared"] > 0
    if os.path.exists(D.REF_DEMO):  # live, against the reference's process on this machine
        r = D.drive(D.ref_cmd(name, start, goal), img, str(tmp_path), "ref")
        assert r.ended and r.returncode == 0
        D.compare(r.steps, g.steps, shape)
