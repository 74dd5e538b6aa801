"""Generate tests/golden/parents.npz: the INFO (level-1 back-pointers) the UNMODIFIED reference planners hold after
step() (oracle/_ref, ref_dump_info), with the map, the goal, their closed set and their g values. Run in the build
container only (needs /root/reference):

    python scripts/make_golden_parents.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle_py as O  # noqa: E402
from unige_tasi_path_planners_b200 import synth  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "parents.npz")


def main():
    if not O.ref_available():
        raise SystemExit("oracle/_ref missing: run `make -C oracle ref` in the build container")
    data, n = {}, 0
    for kind in (0, 1, 2):
        for seed in (0, 1):
            L, W = 40 + 7 * seed, 50 - 3 * seed
            start, goal = (3.0, 3.0), (float(L - 9), float(W - 8))
            cmap = synth.synth_map(20 + seed, L, W, [10, 25][seed], start, goal)
            r = O.RefPlanner(kind, 1)
            r.set_map(cmap)
            r.set_start(*start)
            r.set_goal(*goal)
            r.step()
            closed, g = r.closed_mask()
            info = r.dump_info()
            r.close()
            k = "c%d_" % n
            data[k + "kind"] = np.array([kind], np.int32)
            data[k + "map"] = cmap
            data[k + "goal"] = np.array(goal, np.float32)
            data[k + "closed"] = closed
            data[k + "g"] = g
            data[k + "info"] = info
            n += 1
    data["n"] = np.array([n], np.int32)
    np.savez_compressed(OUT, **data)
    print("wrote %s: %d cases, %d bytes" % (OUT, n, os.path.getsize(OUT)))


if __name__ == "__main__":
    main()
