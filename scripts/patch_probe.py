import sys, time
import numpy as np
sys.path.insert(0, ".")
from unige_tasi_path_planners_b200 import PLANNERS, synth
import torch
side=16384
start, goal = synth.default_endpoints(side, side)
p = PLANNERS[2](2); p.set_occupancy_threshold(1.0)
p.synth_map(2, side, side, 0, 10, start, goal); p.set_start(*start); p.set_goal(*goal); assert p.step()==0
ps=1638
patch = synth.synth_rows(7779, 0, ps, ps, 5)
pin = torch.from_numpy(patch.copy()).pin_memory().numpy()
for name, arr in (("pageable", patch), ("pinned", pin)):
    for rep in range(3):
        torch.cuda.synchronize(); t0=time.perf_counter()
        n = p.patch_map(arr, 0, 0)
        t1=time.perf_counter()
        p.set_start(11.0, 11.0); assert p.step()==0
        t2=time.perf_counter()
        st=p.stats()
        print(name, rep, "patch_map %.2f ms  step %.2f ms (device %.2f: inval %.2f solve %.2f)  changed %d" % ((t1-t0)*1e3, (t2-t1)*1e3, st["step_ms_device"], st["kernel_ms_inval"], st["kernel_ms_solve"], n))
        # alternate patches so that each call changes cells
        arr = (arr ^ 1) | 1
