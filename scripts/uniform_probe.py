"""Development probe: how much of the visit count is due to heterogeneous costs (uniform map vs random map)."""
import sys, json
import numpy as np
sys.path.insert(0, __file__.rsplit("/", 2)[0])
from unige_tasi_path_planners_b200 import PLANNERS, synth
N = 4096
for name, cmap in (("uniform", np.full((N, N), 10, np.uint8)), ("random-nocost-obst10", None), ("random", synth.synth_map(1, N, N, 10, (5, 5), (N - 6, N - 6)))):
    if cmap is None:
        cmap = np.where(synth.synth_map(1, N, N, 10, (5, 5), (N - 6, N - 6)) == 255, 255, 10).astype(np.uint8)
    for kind in (1, 2):
        p = PLANNERS[kind](PLANNERS[kind].max_level)
        p.set_occupancy_threshold(1.0); p.set_map(cmap); p.set_start(5.0, 5.0); p.set_goal(N - 6.0, N - 6.0)
        for _ in range(2):
            p.reset(); p.step()
        s = p.stats()
        print(name, kind, "ms %.2f" % s["kernel_ms_solve"], "vis/tile %.2f" % (s["tile_visits"] / s["tiles"]), "rnd/vis %.2f" % (s["tile_rounds"] / s["tile_visits"]), "phases", s["phases"], "low/el %.2f" % (s["elem_lowered"] / N / N), flush=True)
        p.close()
