import sys
sys.path.insert(0, __file__.rsplit("/", 2)[0])
from unige_tasi_path_planners_b200 import PLANNERS, synth
kind = int(sys.argv[1]) if len(sys.argv) > 1 else 0
L = W = 160
start, goal = (5.0, 5.0), (80.0, 80.0)
cmap = synth.synth_map(21, L, W, 0, start, goal)
p = PLANNERS[kind](1)
p.set_occupancy_threshold(1.0)
p.set_map(cmap); p.set_start(*start); p.set_goal(*goal)
import ctypes as C
try:
    print("rc", p.step(), p.stats()["tile_visits"])
except Exception as e:
    print("ERR", e)
