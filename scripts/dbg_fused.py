import sys
sys.path.insert(0, __file__.rsplit("/", 2)[0])
import numpy as np
from unige_tasi_path_planners_b200 import PLANNERS, ShardedPlanner, synth
kind = int(sys.argv[1]) if len(sys.argv) > 1 else 0
L, W = 420, 260
start, goal = (5.0, 5.0), (float(L - 6), float(W - 6))
cmap = synth.synth_map(61, L, W, 10, start, goal)
p = ShardedPlanner(kind, devices=[0], bands_per_device=4)
p.set_occupancy_threshold(1.0)
p.set_map(cmap); p.set_start(*start); p.set_goal(*goal)
assert p.step() == 0
print("plan", p.shard_stats(), {k: p.stats()[k] for k in ("tile_visits", "phases", "tiles")})
x, y, h, w = 90, 60, 180, 120
old = cmap[x:x + h, y:y + w]
mode = sys.argv[2] if len(sys.argv) > 2 else "lower"
rng = np.random.default_rng(7)
if mode == "raise":
    patch = np.where(rng.random(old.shape) < 0.1, 255, np.minimum(old.astype(np.int32) + 9, 255)).astype(np.uint8)
else:
    patch = np.maximum(old.astype(np.int32) - 9, 1).astype(np.uint8)
print("changed", p.patch_map(patch, x, y))
patched = cmap.copy(); patched[x:x + h, y:y + w] = patch
p.set_start(8.0, 8.0)
assert p.step() == 0
print("ALLSTATS", p.stats())
print("replan", p.shard_stats(), {k: p.stats()[k] for k in ("tile_visits", "phases", "tiles", "seed_elems", "elem_invalidated", "elem_lowered")})
got = p.read_field()
q = PLANNERS[kind](1); q.set_occupancy_threshold(1.0); q.set_map(patched); q.set_start(8.0, 8.0); q.set_goal(*goal); q.step()
want = q.read_field()
bad = np.argwhere((got != want) & ~(np.isnan(got) | np.isnan(want)))
print("inf mismatch", int((np.isfinite(got) != np.isfinite(want)).sum()), "got<want", int((got < want).sum()), "got>want", int((got > want).sum()))
print("bad", len(bad), "rows", np.unique(bad[:, 0])[:40], "cols range", bad[:, 1].min() if len(bad) else None, bad[:, 1].max() if len(bad) else None)
if len(bad):
    for r, c in bad[:10]:
        print(r, c, got[r, c], want[r, c])
    # per tile summary
    t = {(r // 32, c // 32) for r, c in bad}
    print("tiles", sorted(t)[:60])
