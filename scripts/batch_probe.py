"""Where a cfg5 chunk's time goes (stacked multi-goal batch): host stacking, H2D, plan, windows, walks."""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from unige_tasi_path_planners_b200 import batch, synth  # noqa: E402

side, chunk = 1024, 256
qs = batch.make_queries(16, 16, side)
maps = {m: synth.synth_map(1000 + m, side, side, 10, None, None) for m in range(16)}
for kind, name in ((0, "FD*"), (1, "MS-DFM")):
    sb = batch.StackedBatch(kind, 1, 0)
    for rep in range(2):
        t = [time.perf_counter()]
        stacked = batch.stack_maps([maps[q[0]] for q in qs[:chunk]]); t.append(time.perf_counter())
        sb.L, sb.W, sb.n = side, side, chunk
        sb.p.set_map(stacked); t.append(time.perf_counter())
        sb.goals = [q[2] for q in qs[:chunk]]
        sb.p.set_start(0.0, 0.0)
        sb.p.set_goals([(g[0] + i * (side + 1), g[1]) for i, g in enumerate(sb.goals)]); t.append(time.perf_counter())
        assert sb.p.step() == 0; t.append(time.perf_counter())
        buf = np.empty((64, 64), np.float32)
        for i, q in enumerate(qs[:chunk]):
            sb.window(i, int(q[1][0]) - 32, int(q[1][1]) - 32, 64, 64, out=buf)
        t.append(time.perf_counter())
        res, ms = sb.paths([q[1] for q in qs[:chunk]], 20); t.append(time.perf_counter())
    d = np.diff(t) * 1e3
    print(name, "stack %.1f  set_map %.1f  goals %.1f  step %.1f (device %.1f)  windows %.1f  walks %.1f  total %.1f ms" % (
        d[0], d[1], d[2], d[3], sb.p.stats()["step_ms_device"], d[4], d[5], d.sum()), flush=True)
    sb.close()
