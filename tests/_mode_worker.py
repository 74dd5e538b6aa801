"""Worker of tests/test_solver_modes.py: plans + replans one map with the three planners in THIS process's solver mode
(TASI_TMA / TASI_COMPACT / TASI_STRIP / TASI_REL_PASSES are read once per process) and writes the fields to argv[1]."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from unige_tasi_path_planners_b200 import PLANNERS, synth  # noqa: E402

L, W = 200, 170
start, goal = (5.0, 5.0), (float(L - 6), float(W - 6))
cmap = synth.synth_map(777, L, W, 12, start, goal)
patch = synth.synth_rows(31, 0, 60, 70, 25)
out = {}
for kind in (0, 1, 2):
    p = PLANNERS[kind](PLANNERS[kind].max_level, device=0)
    p.set_occupancy_threshold(1.0)
    p.set_map(cmap)
    p.set_start(*start)
    p.set_goal(*goal)
    assert p.step() == 0
    out["plan%d" % kind] = p.read_field()
    p.patch_map(patch, 80, 50)
    p.set_start(9.0, 9.0)
    assert p.step() == 0
    out["replan%d" % kind] = p.read_field()
    p.close()
np.savez(sys.argv[1], **out)
