"""Generate tests/golden/paths.npz by running the UNMODIFIED reference path extractor
(PathExtraction/LinearInterpolationPathExtractor.h, built into oracle/_ref by oracle/Makefile) over converged fields.
Run in the build container only (needs /root/reference):

    python scripts/make_golden_paths.py

Per case: the map, the field handed to the extractor (an ExpandedMap filled with g = rhs = field), start, goal, the
options (max_steps, lookahead, allow_indirect_traversals) and the reference's path_, cost_, total_cost, total_dist.
"""
import os
import random
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle_py as O  # noqa: E402
from unige_tasi_path_planners_b200 import synth  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "paths.npz")


def cases():
    """(kind, seed, L, W, p_obs, start, goal, max_steps, lookahead, indirect)"""
    out = []
    for kind in (0, 1, 2):
        for seed in range(8):
            rng = random.Random(1000 * kind + seed)
            L, W = rng.choice([24, 40, 56]), rng.choice([24, 33, 48])
            start = [float(rng.randrange(2, L - 2)), float(rng.randrange(2, W - 2))]
            goal = (float(rng.randrange(2, L - 2)), float(rng.randrange(2, W - 2)))
            p_obs = rng.choice([0, 10, 25, 40])
            frac = rng.randrange(3)  # the rover may stand on an edge (it advances along path points)
            off = rng.choice([0.5, 0.25, 0.7313])
            out.append((kind, seed, L, W, p_obs, tuple(start), goal, frac, off,
                        rng.choice([20, 1000]), seed % 4 != 3, seed % 3 != 2))
    return out


def main():
    if not O.ref_available():
        raise SystemExit("oracle/_ref missing: run `make -C oracle ref` in the build container")
    data = {}
    n = 0
    for kind, seed, L, W, p_obs, start, goal, frac, off, ms, la, ind in cases():
        cmap = synth.synth_map(seed, L, W, p_obs, start, goal)
        if frac == 1:
            start = (start[0] + off, start[1])
        elif frac == 2:
            start = (start[0], start[1] + off)
        field = O.solve(kind, cmap, goal)
        r = O.RefPlanner(kind, 1)
        r.set_map(cmap)
        r.set_start(*start)
        r.set_goal(*goal)
        xy, cost, dist, _ = r.extract(field=field, max_steps=ms, lookahead=la, allow_indirect=ind)
        sc = r.step_costs()
        r.close()
        k = "c%02d_" % n
        data[k + "map"] = cmap
        data[k + "field"] = field
        data[k + "meta"] = np.array([kind, ms, int(la), int(ind)], np.int32)
        data[k + "ends"] = np.array([start[0], start[1], goal[0], goal[1]], np.float32)
        data[k + "path"] = xy
        data[k + "costs"] = sc
        data[k + "totals"] = np.array([cost, dist], np.float32)
        n += 1
    data["n"] = np.array([n], np.int32)
    np.savez_compressed(OUT, **data)
    print("wrote %s: %d cases, %d bytes" % (OUT, n, os.path.getsize(OUT)))


if __name__ == "__main__":
    main()
