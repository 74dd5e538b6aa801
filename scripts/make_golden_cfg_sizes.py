"""Golden fixtures at BASELINE config size: the UNMODIFIED reference planners (oracle/_ref, built from /root/reference by
oracle/Makefile) plan the 4096 x 4096 grid of cfg2 (generator of SURVEY.md 8d, seed 1, 10 % obstacles, start (5,5),
goal (4090,4090), occupancy threshold 1.0) once per planner, and a compact digest of the reference's closed set
{g == rhs < INF and g < top_key} is committed:

    tests/golden/cfg2_4096_<fd|dfm|mfd>.npz
        top_key         the reference queue's top key when plan() returned (float32; +INF = queue empty)
        rows            indices of the sampled rows (every 64th element row + the last)
        row_values      reference g of the sampled rows, NaN where the element is not in the closed set
        closed_per_row  number of closed elements in every element row
        row_sum         float64 sum of the closed values of every element row
        row_crc         CRC32 of the closed values of every element row, in column order (float32 bytes)
        p_ms, expanded  the reference's own p_time and num_nodes_expanded of that plan (context for the CPU baseline)

The GPU test (tests/test_config_sizes.py) rebuilds the closed set from its own field (value < top_key) and compares
FD* / MFD* bit for bit (row CRCs) and MS-DFM within 1e-4 (sampled rows and row sums).

    python scripts/make_golden_cfg_sizes.py [fd|dfm|mfd ...]       (about 1-2 minutes and 1 GB per planner)
"""
import os
import sys
import time
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import oracle_py as O  # noqa: E402
from unige_tasi_path_planners_b200 import synth  # noqa: E402

SIDE, SEED, P_OBS = 4096, 1, 10
NAMES = {"fd": 0, "dfm": 1, "mfd": 2}


def digest(kind, side=SIDE, seed=SEED):
    start, goal = synth.default_endpoints(side, side)
    cmap = synth.synth_map(seed, side, side, P_OBS, start, goal)
    r = O.RefPlanner(kind, 2 if kind == 2 else 1)
    r.set_map(cmap, 1.0)
    r.set_start(*start)
    r.set_goal(*goal)
    t0 = time.time()
    assert r.step() == 0
    wall = time.time() - t0
    _, p_ms = r.times()
    _, expanded = r.counters()
    g, rhs = r.dump()
    top = np.float32(r.top_key())
    r.close()
    closed = (g == rhs) & np.isfinite(g) & (g < top)
    ex = g.shape[0]
    rows = np.unique(np.concatenate([np.arange(0, ex, 64), [ex - 1]])).astype(np.int32)
    row_values = np.where(closed[rows], g[rows], np.float32(np.nan)).astype(np.float32)
    cnt = closed.sum(axis=1).astype(np.int32)
    rsum = np.array([g[i][closed[i]].astype(np.float64).sum() for i in range(ex)])
    crc = np.array([zlib.crc32(np.ascontiguousarray(g[i][closed[i]]).tobytes()) for i in range(ex)], np.uint32)
    print("kind %d: p_time %.0f ms (wall %.1f s), expanded %d, closed %d of %d, top_key %r" % (
        kind, p_ms, wall, expanded, int(cnt.sum()), g.size, float(top)), flush=True)
    return dict(top_key=top, rows=rows, row_values=row_values, closed_per_row=cnt, row_sum=rsum, row_crc=crc,
                p_ms=np.float64(p_ms), expanded=np.int64(expanded), side=np.int32(side), seed=np.int32(seed))


def main():
    assert O.ref_available(), "oracle/_ref is not built (needs /root/reference): make -C oracle ref"
    which = [a for a in sys.argv[1:] if a in NAMES] or list(NAMES)
    for name in which:
        d = digest(NAMES[name])
        out = os.path.join(ROOT, "tests", "golden", "cfg2_%d_%s.npz" % (SIDE, name))
        np.savez_compressed(out, **d)
        print(out, os.path.getsize(out), "bytes", flush=True)


if __name__ == "__main__":
    main()
