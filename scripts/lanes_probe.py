"""Development probe: throughput of the query batch against lanes and CTAs per lane."""
import sys, time, threading
sys.path.insert(0, __file__.rsplit("/", 2)[0])
from unige_tasi_path_planners_b200 import batch, planner
side = 1024
qs = batch.make_queries(8, 8, side)
for lanes, blocks in ((1, 444), (1, 148), (1, 55), (2, 148), (4, 111), (8, 55), (8, 148)):
    lane_q = [[q for q in qs if q[0] % lanes == l] for l in range(lanes)]
    out = [None] * lanes
    t0 = time.perf_counter()
    th = [threading.Thread(target=batch.run_lane, args=(planner.DFM, 1, 0, blocks, side, lane_q[l], out, l)) for l in range(lanes)]
    [t.start() for t in th]; [t.join() for t in th]
    dt = time.perf_counter() - t0
    print("lanes", lanes, "blocks", blocks, "queries/s %.1f" % (len(qs) / dt), flush=True)
