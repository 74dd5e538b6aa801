"""Focused planning (tasi_set_focus) against full convergence: initial plan and a replan after a patch around a
rover that stands part-way to the goal. One JSON line per planner."""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from unige_tasi_path_planners_b200 import PLANNERS, synth  # noqa: E402

side = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
frac = float(sys.argv[2]) if len(sys.argv) > 2 else 0.6   # rover at this fraction of the way from the far corner
goal = (float(side - 6), float(side - 6))
rover = (float(int(side * frac)), float(int(side * frac)))
ps = side // 10
for kind, name in ((0, "FD*"), (1, "MS-DFM"), (2, "MFD*")):
    out = {"planner": name, "grid": side, "rover": rover}
    for focus in (False, True):
        p = PLANNERS[kind](1)
        p.set_occupancy_threshold(1.0)
        p.synth_map(3, side, side, 0, 10, rover, goal)
        p.set_focus(focus)
        p.set_start(*rover)
        p.set_goal(*goal)
        for _ in range(2):
            p.reset()
            t0 = time.perf_counter()
            assert p.step() == 0
            plan_ms = (time.perf_counter() - t0) * 1e3
        st = p.stats()
        base = p.read_map()
        x0 = y0 = int(rover[0]) - ps // 2
        patch = synth.synth_rows(991, 0, ps, ps, 5)
        patch[ps // 2 - 4:ps // 2 + 5, ps // 2 - 4:ps // 2 + 5] = 1
        t0 = time.perf_counter()
        p.patch_map(patch, x0, y0)
        p.set_start(rover[0] + 3.0, rover[1] + 2.0)
        assert p.step() == 0
        replan_ms = (time.perf_counter() - t0) * 1e3
        st2 = p.stats()
        out["focused" if focus else "full"] = {
            "plan_ms": plan_ms, "plan_visits": st["tile_visits"], "settled_below": st["settled_below"],
            "replan_ms": replan_ms, "replan_device_ms": st2["step_ms_device"],
            "replan_visits": st2["tile_visits"] + st2["inval_visits"]}
        p.close()
    print(json.dumps(out), flush=True)
