"""Development probe (not the bench contract): time an initial plan per planner and size, print the work counters."""
import json
import sys
import time

sys.path.insert(0, __file__.rsplit("/", 2)[0])
from unige_tasi_path_planners_b200 import PLANNERS, synth  # noqa: E402

import os  # noqa: E402
sizes = [int(a) for a in sys.argv[1:]] or [1024, 4096]
kinds = [int(k) for k in os.environ.get("QUICK_KINDS", "1,2,0").split(",")]
for N in sizes:
    for kind in kinds:
        p = PLANNERS[kind](PLANNERS[kind].max_level)
        p.set_occupancy_threshold(1.0)
        start, goal = synth.default_endpoints(N, N)
        p.synth_map(1, N, N, 0, 10, start, goal)
        p.set_start(*start)
        p.set_goal(*goal)
        if os.environ.get("QUICK_BLOCKS"):
            p.set_solver_blocks(int(os.environ["QUICK_BLOCKS"]))
        best = None
        for it in range(3):
            p.reset()
            t = time.time()
            rc = p.step()
            dt = time.time() - t
            s = p.stats()
            if best is None or s["kernel_ms_solve"] < best["kernel_ms_solve"]:
                best = dict(s, wall_ms=dt * 1e3, rc=rc)
        best.update(kind=kind, N=N, mcells_s=N * N / best["kernel_ms_solve"] / 1e3,
                    visits_per_tile=best["tile_visits"] / best["tiles"],
                    rounds_per_visit=best["tile_rounds"] / max(best["tile_visits"], 1),
                    evals_per_elem=best["elem_evals"] / (N * N), lowered_per_elem=best["elem_lowered"] / (N * N))
        print(json.dumps(best), flush=True)
        p.close()
