"""Development probe: the fused sharded planner (ONE process over the visible GPUs, tasi_create_sharded) on one FD* grid,
swept over the band factor of the value-ordered schedule and the bands per GPU; every setting is checked against the
checksum of the unsharded plan.   python scripts/shard_sweep.py [side] [n_dev] [settings "factor:bands,..."; factor 0 = rule]"""
import json
import os
import sys
import time

sys.path.insert(0, __file__.rsplit("/", 2)[0])
import torch  # noqa: E402
from unige_tasi_path_planners_b200 import PLANNERS, ShardedPlanner, synth  # noqa: E402

side = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
ndev = int(sys.argv[2]) if len(sys.argv) > 2 else torch.cuda.device_count()
settings = sys.argv[3] if len(sys.argv) > 3 else "0:4,0:8,2:4,3:4,4:4,3:8,4:8"
kind = 0
start, goal = synth.default_endpoints(side, side)
os.environ.pop("TASI_DELTA_FACTOR", None)
q = PLANNERS[kind](1)
q.set_occupancy_threshold(1.0)
q.synth_map(3, side, side, 0, 10, start, goal)
q.set_start(*start); q.set_goal(*goal)
for _ in range(2):
    q.reset(); assert q.step() == 0
want = q.field_checksum()
base = q.stats()["step_ms_device"]
print(json.dumps({"side": side, "unsharded_ms": base, "visits_per_tile": q.stats()["tile_visits"] / q.stats()["tiles"]}), flush=True)
q.close()
torch.cuda.empty_cache()
for st in settings.split(","):
    f, per = st.split(":")
    if float(f) > 0:
        os.environ["TASI_DELTA_FACTOR"] = f
    else:
        os.environ.pop("TASI_DELTA_FACTOR", None)
    p = ShardedPlanner(kind, 1, devices=list(range(ndev)), bands_per_device=int(per))
    p.set_occupancy_threshold(1.0)
    p.synth_map(3, side, side, 0, 10, start, goal)
    p.set_start(*start); p.set_goal(*goal)
    best = None
    for _ in range(3):
        p.reset()
        t = time.perf_counter()
        assert p.step() == 0
        dt = (time.perf_counter() - t) * 1e3
        if best is None or dt < best[0]:
            best = (dt, p.shard_stats(), p.stats())
    ok = p.field_checksum() == want
    ss, s2 = best[1], best[2]
    print(json.dumps({"n_dev": ndev, "factor": float(f), "bands_per_gpu": ss["bands_per_rank"], "ms": best[0], "speedup": base / best[0],
                      "fused": ss["fused"], "remote_act": ss["remote_activations"], "busy": [ss["busy_frac_min_rank"], ss["busy_frac_max_rank"]],
                      "visits_per_tile": s2["tile_visits"] / s2["tiles"], "phases": s2["phases"], "verified": ok}), flush=True)
    p.close()
