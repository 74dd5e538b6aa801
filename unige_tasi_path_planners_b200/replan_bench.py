"""bench.py --workload cfg3: MFD* (ShiftedGridPlanner level 2) on a side x side synthetic grid (16384 by default):
initial plan, then the rover advances and a square patch covering ~1 % of the cells around it changes cost
(SURVEY.md 8d): mixed (re-drawn with another seed, 5 % obstacles), raise-only and lower-only variants.

Reports the replan latency (patch_map + set_start + step, wall clock with host buffers, and the device time of
the step) and checks, outside the timed region, that the repaired field equals a from-scratch plan of the
patched map on the same device (bit-exact for MFD*): the size-independent parity property of a replan.
"""
import json
import time

import numpy as np

from . import planner as _planner
from . import synth


def _patch_variant(kind, old, seed):
    if kind == "mixed":
        return synth.synth_rows(seed, 0, old.shape[0], old.shape[1], 5)
    rng = np.random.default_rng(seed)
    if kind == "raise":
        return np.where(rng.random(old.shape) < 0.05, 255, np.minimum(old.astype(np.int32) + 7, 255)).astype(np.uint8)
    return np.maximum(old.astype(np.int32) - 7, 1).astype(np.uint8)  # lower (obstacles open up too)


def bench_main(args, rank, world):
    import torch
    if rank != 0:
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    side = args.side if args.side != 4096 else 16384
    start, goal = synth.default_endpoints(side, side)
    p = _planner.ShiftedGridPlanner(2, device=0)
    p.set_occupancy_threshold(1.0)
    p.synth_map(2, side, side, 0, 10, start, goal)
    p.set_start(*start)
    p.set_goal(*goal)
    plan_ms = []
    for _ in range(max(args.warmup, 1) + args.steps):
        p.reset()
        assert p.step() == 0
        plan_ms.append(p.stats()["step_ms_device"])
    plan_ms = plan_ms[max(args.warmup, 1):]
    base_map = p.read_map()

    ps = int(0.1 * side)                       # 0.1 N x 0.1 N = 1 % of the cells
    rover = (float(start[0] + 6), float(start[1] + 6))  # the rover moved on along its path
    x0 = int(min(max(rover[0] - ps // 2, 0), side - ps))
    y0 = int(min(max(rover[1] - ps // 2, 0), side - ps))
    results = {}
    ref = _planner.ShiftedGridPlanner(2, device=0)  # from-scratch plans of the patched maps (the check)
    ref.set_occupancy_threshold(1.0)
    for variant in ("mixed", "raise", "lower", "mixed_focused"):
        focused = variant.endswith("_focused")
        lat, dev, inval, solve, changed, visits = [], [], [], [], 0, 0
        old = base_map[x0:x0 + ps, y0:y0 + ps]
        patch = _patch_variant(variant.split("_")[0], old, 2 + 7777)
        patch[:12, :12] = np.minimum(patch[:12, :12], 64)  # keep the rover's own cells free
        for it in range(args.steps):
            # restore the unpatched, converged state (untimed)
            p.set_map(base_map)
            p.set_start(*start)
            p.set_goal(*goal)
            p.reset()
            assert p.step() == 0
            p.set_focus(focused)  # focused: the repair stops once the rover's cell is settled (tasi_set_focus)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            changed = p.patch_map(patch, x0, y0)
            p.set_start(*rover)
            rc = p.step()
            lat.append((time.perf_counter() - t0) * 1e3)
            assert rc == 0
            st = p.stats()
            dev.append(st["step_ms_device"]); inval.append(st["kernel_ms_inval"]); solve.append(st["kernel_ms_solve"])
            visits = st["tile_visits"] + st["inval_visits"]
            p.set_focus(False)
        got = p.read_field()
        patched = base_map.copy()
        patched[x0:x0 + ps, y0:y0 + ps] = patch
        ref.set_map(patched)
        ref.set_start(*rover)
        ref.set_goal(*goal)
        ref.reset()
        assert ref.step() == 0
        want = ref.read_field()
        if focused:  # only values below settled_below are final; the rest must be upper bounds
            m = want < st["settled_below"]
            same = bool(np.array_equal(got[m], want[m]) and np.all(got >= want))
        else:
            same = bool(np.array_equal(got, want))
        results[variant] = {"latency_ms": float(np.mean(lat)), "device_ms": float(np.mean(dev)),
                            "invalidate_ms": float(np.mean(inval)), "repropagate_ms": float(np.mean(solve)),
                            "changed_cells": int(changed), "seed_nodes": int(p.num_nodes_updated),
                            "tile_visits": int(visits), "tiles": int(st["tiles"]),
                            "settled_below": float(st["settled_below"]),
                            "equals_from_scratch_plan": same}
        del got, want
    cells = side * side
    ms = float(np.mean(plan_ms))
    print(json.dumps({
        "metric": "replan latency ms after 1%% of cells change near the rover, MFD* %dx%d" % (side, side),
        "value": results["mixed"]["latency_ms"], "unit": "ms", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": results["mixed"]["latency_ms"], "higher_is_better": False, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "cfg3", "planner": "MFD* (ShiftedGridPlanner level 2)", "grid": "%dx%d" % (side, side),
                   "patch": "%dx%d at (%d,%d)" % (ps, ps, x0, y0), "rover": list(rover),
                   "timing": "wall clock of patch_map(host patch) + set_start + step; device_ms = CUDA events"},
        "initial_plan": {"ms": ms, "mcells_per_s": cells / ms / 1e3,
                         "roofline_frac": 5.0 * cells / (ms / 1e3) / 1e9 / 6543.4},
        "replan": results}), flush=True)
    p.close()
    ref.close()
