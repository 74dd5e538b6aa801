"""Development probe: the sharded planner driven by ONE process over the visible GPUs (tasi_create_sharded), FD* by
default: time per plan and exchanges for several bands-per-GPU / phases-per-exchange settings, checksum against the
unsharded plan.  python scripts/shard_probe.py [side] [kind] [n_dev]"""
import json
import os
import sys
import time

sys.path.insert(0, __file__.rsplit("/", 2)[0])
import torch  # noqa: E402
from unige_tasi_path_planners_b200 import PLANNERS, ShardedPlanner, synth  # noqa: E402

side = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
kind = int(sys.argv[2]) if len(sys.argv) > 2 else 0
ndev = int(sys.argv[3]) if len(sys.argv) > 3 else torch.cuda.device_count()
start, goal = synth.default_endpoints(side, side)
q = PLANNERS[kind](1)
q.set_occupancy_threshold(1.0)
q.synth_map(3, side, side, 0, 10, start, goal)
q.set_start(*start); q.set_goal(*goal)
for _ in range(2):
    q.reset(); assert q.step() == 0
want = q.field_checksum()
print(json.dumps({"unsharded_ms": q.stats()["step_ms_device"], "visits_per_tile": q.stats()["tile_visits"] / q.stats()["tiles"]}), flush=True)
q.close()
settings = [(b, k) for b in (1, 2, 4, 8, 16) for k in (2,)] + [(4, 1), (4, 4), (8, 4)]
if os.environ.get("PROBE_SETTINGS"):
    settings = [tuple(int(v) for v in s.split(":")) for s in os.environ["PROBE_SETTINGS"].split(",")]
for per, k in settings:
    p = ShardedPlanner(kind, 1, devices=list(range(ndev)), bands_per_device=per, phases_per_exchange=k)
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
    ss, st = best[1], best[2]
    print(json.dumps({"n_dev": ndev, "bands_per_gpu": ss["bands_per_rank"], "phases": k, "ms": best[0], "exchanges": ss["exchanges"],
                      "solver_ms_max": ss["solver_ms_max_rank"], "solver_ms_min": ss["solver_ms_min_rank"], "nccl": ss["nccl"], "fused": ss["fused"], "remote_act": ss["remote_activations"], "busy": [ss["busy_frac_min_rank"], ss["busy_frac_max_rank"]],
                      "visits_per_tile": st["tile_visits"] / st["tiles"], "verified": ok}), flush=True)
    p.close()
