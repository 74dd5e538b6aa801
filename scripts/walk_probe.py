"""Latency of one device walk as a function of its length: fixed cost of the call vs cost per step."""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from unige_tasi_path_planners_b200 import PLANNERS, synth  # noqa: E402

side = 1024
start, goal = (900.0, 880.0), (20.0, 30.0)
for kind, name in ((0, "FD*"), (1, "MS-DFM"), (2, "MFD*")):
    p = PLANNERS[kind](1)
    p.set_occupancy_threshold(1.0)
    p.synth_map(7, side, side, 0, 10 if kind != 1 else 0, start, goal)
    p.set_start(*start)
    p.set_goal(*goal)
    assert p.step() == 0
    out = []
    for steps in (1, 5, 20, 80, 320):
        for _ in range(5):
            p.extract_paths(None, None, steps)
        t0 = time.perf_counter()
        for _ in range(50):
            res, ms = p.extract_paths(None, None, steps)
        dt = (time.perf_counter() - t0) / 50 * 1e6
        out.append("%d steps: %.0f us (%d pts)" % (steps, dt, len(res[0]["path"])))
    print(name, " | ".join(out), flush=True)
    p.close()
