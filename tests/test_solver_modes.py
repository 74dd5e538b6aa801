"""Every instantiation of the tile solver reaches the oracle's field: strips / CTA list / warp lists (TASI_COMPACT),
2- and 4-element strips (TASI_STRIP), TMA staging (TASI_TMA=1), pile-order release (TASI_REL_PASSES=1), the FIFO near
queue (TASI_BUCKETS=0) and other ring counts of the bucketed one, the upwind deferral (TASI_DEFER) and chaining
(TASI_CHAIN). The modes are
read once per process, so each runs in a worker process (tests/_mode_worker.py)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from unige_tasi_path_planners_b200 import synth

HERE = os.path.dirname(os.path.abspath(__file__))
KNOBS = ("TASI_TMA", "TASI_COMPACT", "TASI_STRIP", "TASI_REL_PASSES", "TASI_DELTA_FACTOR", "TASI_BUCKETS", "TASI_DEFER", "TASI_CHAIN", "TASI_SLIDE")
MODES = [{"TASI_TMA": "1"}, {"TASI_TMA": "1", "TASI_STRIP": "2"}, {"TASI_COMPACT": "0"}, {"TASI_COMPACT": "1"},
         {"TASI_COMPACT": "2"}, {"TASI_COMPACT": "1", "TASI_STRIP": "4"}, {"TASI_COMPACT": "3"},
         {"TASI_COMPACT": "3", "TASI_STRIP": "4"}, {"TASI_STRIP": "4"}, {"TASI_REL_PASSES": "1"},
         {"TASI_REL_PASSES": "32", "TASI_DELTA_FACTOR": "6"}, {"TASI_BUCKETS": "0"}, {"TASI_BUCKETS": "4"},
         {"TASI_BUCKETS": "16", "TASI_DELTA_FACTOR": "3"}, {"TASI_BUCKETS": "60"}, {"TASI_DEFER": "0.5"},
         {"TASI_DEFER": "0.1", "TASI_DELTA_FACTOR": "8"}, {"TASI_CHAIN": "1"}, {"TASI_SLIDE": "1"},
         {"TASI_SLIDE": "1", "TASI_DELTA_FACTOR": "2"}]


@pytest.fixture(scope="module")
def want(oracle):
    L, W = 200, 170
    start, goal = (5.0, 5.0), (float(L - 6), float(W - 6))
    cmap = synth.synth_map(777, L, W, 12, start, goal)
    cm2 = cmap.copy()
    oracle.patch(cm2, synth.synth_rows(31, 0, 60, 70, 25), 80, 50)
    return {k: (oracle.solve(k, cmap, goal), oracle.solve(k, cm2, goal)) for k in (0, 1, 2)}


@pytest.mark.gpu
@pytest.mark.parametrize("mode", MODES, ids=lambda m: ",".join("%s=%s" % kv for kv in m.items()))
def test_solver_mode_matches_oracle(mode, want, gpu_lib, tmp_path):
    out = str(tmp_path / "fields.npz")
    env = dict(os.environ)
    for k in KNOBS:
        env.pop(k, None)
    env.update(mode)
    r = subprocess.run([sys.executable, os.path.join(HERE, "_mode_worker.py"), out], env=env, capture_output=True,
                       text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    got = np.load(out)
    for k in (0, 1, 2):
        for name, w in zip(("plan", "replan"), want[k]):
            g = got["%s%d" % (name, k)]
            assert np.array_equal(np.isfinite(g), np.isfinite(w)), (k, name)
            m = np.isfinite(w)
            if k == 1:  # MS-DFM: 1e-4 relative (north star); the node planners are bit-exact
                assert (np.abs(g[m] - w[m]) / np.maximum(w[m], 1e-30)).max(initial=0.0) <= 1e-4
            else:
                assert np.array_equal(g[m], w[m]), (k, name)
