"""Development probe (1 GPU): cost of band_solve_until launches that find little or no work."""
import sys, time
import numpy as np
sys.path.insert(0, __file__.rsplit("/", 2)[0])
from unige_tasi_path_planners_b200 import bands, planner, synth
side, parts = 8192, 8
start, goal = synth.default_endpoints(side, side)
bs = [bands.GpuBand(planner.FD, 1, side, side, parts, r) for r in range(parts)]
for b in bs:
    b.synth_map(3, 10, start, goal); b.set_endpoints(start, goal); b.begin()
step = 2 * max(b.delta() for b in bs); limit = step
rows = []
outer = 0
while True:
    parked = []
    for r, b in enumerate(bs):
        v0 = b.stats()["tile_visits"]; k0 = b.stats()["kernel_ms_solve"]
        t0 = time.perf_counter(); pk = b.solve_until(limit); dt = (time.perf_counter() - t0) * 1e3
        st = b.stats()
        rows.append((st["tile_visits"] - v0, dt, st["kernel_ms_solve"] - k0, st["phases"]))
        parked.append(pk)
    for r in range(parts - 1):
        bs[r + 1].recv_stage(0).copy_(bs[r].send_row(1)); bs[r].recv_stage(1).copy_(bs[r + 1].send_row(0))
    bands._sync(bs)
    low = sum((b.import_side(0) if b.open_top else 0) + (b.import_side(1) if b.open_bottom else 0) for b in bs)
    outer += 1
    g = min(parked)
    if low == 0 and g == float("inf"): break
    if g < float("inf"): limit = max(limit, g + step)
rows = np.array(rows)
idle = rows[rows[:, 0] == 0]; small = rows[(rows[:, 0] > 0) & (rows[:, 0] < 50)]; big = rows[rows[:, 0] >= 50]
for name, a in (("idle", idle), ("small", small), ("big", big)):
    if len(a): print(name, "calls", len(a), "visits/call %.1f" % a[:, 0].mean(), "wall ms %.3f" % a[:, 1].mean(), "kernel ms %.3f" % a[:, 2].mean(), "total kernel ms %.1f" % a[:, 2].sum())
print("outer", outer)
