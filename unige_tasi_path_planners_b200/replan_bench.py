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


def measure(side=16384, steps=5, warmup=1, device=0, peak_gbs=None, variants=("mixed", "raise", "lower", "mixed_focused")):
    """The cfg3 record: MFD* initial plan on side x side (device time), then per patch variant the replan latency
    (wall clock of patch_map(host patch) + set_start + step) and whether the repaired field equals a from-scratch
    plan of the patched map bit for bit. Returns a dict (bench.py puts it into its JSON line as "replan")."""
    import torch
    start, goal = synth.default_endpoints(side, side)
    p = _planner.ShiftedGridPlanner(2, device=device)
    p.set_occupancy_threshold(1.0)
    p.synth_map(2, side, side, 0, 10, start, goal)
    p.set_start(*start)
    p.set_goal(*goal)
    plan_ms, solve_ms = [], []
    for _ in range(max(warmup, 1) + steps):
        p.reset()
        assert p.step() == 0
        st = p.stats()
        plan_ms.append(st["step_ms_device"])
        solve_ms.append(st["kernel_ms_solve"])
    plan_ms, solve_ms = plan_ms[max(warmup, 1):], solve_ms[max(warmup, 1):]
    plan_stats = p.stats()
    base_map = p.read_map()

    ps = int(0.1 * side)                       # 0.1 N x 0.1 N = 1 % of the cells
    rover = (float(start[0] + 6), float(start[1] + 6))  # the rover moved on along its path
    x0 = int(min(max(rover[0] - ps // 2, 0), side - ps))
    y0 = int(min(max(rover[1] - ps // 2, 0), side - ps))
    results = {}
    ref = _planner.ShiftedGridPlanner(2, device=device)  # from-scratch plans of the patched maps (the check)
    ref.set_occupancy_threshold(1.0)
    launches = 0
    for variant in variants:
        focused = variant.endswith("_focused")
        lat, dev, inval, solve, changed, visits = [], [], [], [], 0, 0
        old = base_map[x0:x0 + ps, y0:y0 + ps]
        patch = _patch_variant(variant.split("_")[0], old, 2 + 7777)
        patch[:12, :12] = np.minimum(patch[:12, :12], 64)  # keep the rover's own cells free
        for it in range(steps):
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
            launches += st["solver_launches"] + 2
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
                            "settled_below": float(st["settled_below"]) if np.isfinite(st["settled_below"]) else None,
                            "equals_from_scratch_plan": same}
        del got, want
    p.close()
    ref.close()
    cells = side * side
    ms, k_ms = float(np.mean(plan_ms)), float(np.mean(solve_ms))
    if peak_gbs is None:
        peak_gbs = 6450.6
    return {"workload": "cfg3", "planner": "MFD* (ShiftedGridPlanner level 2)", "grid": "%dx%d" % (side, side),
            "plan_ms": ms, "mcells_per_s": cells / ms / 1e3,
            "roofline_frac": 5.0 * cells / (k_ms / 1e3) / 1e9 / peak_gbs, "solver_kernel_ms": k_ms,
            "tile_visits_per_tile": plan_stats["tile_visits"] / max(plan_stats["tiles"], 1),
            "elem_evals_per_cell": plan_stats["elem_evals"] / cells,
            "patch": "%dx%d at (%d,%d), ~1%% of the cells, re-drawn around the rover" % (ps, ps, x0, y0),
            "rover": list(rover), "reps": steps,
            "replan_ms": {v: results[v]["latency_ms"] for v in results},
            "replan_device_ms": {v: results[v]["device_ms"] for v in results},
            "timing": "replan_ms = wall clock of patch_map(host patch) + set_start + step; plan_ms = CUDA events",
            "verified": all(r["equals_from_scratch_plan"] for r in results.values()),
            "verified_how": "repaired field == from-scratch plan of the patched map on the same device, bit for bit "
                            "(tests/test_config_sizes.py also applies the oracle's operator to the 16384^2 fields)",
            "gpu_launches": launches, "detail": results}


def bench_main(args, rank, world):
    import torch
    if rank != 0:
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    side = args.side if args.side != 4096 else 16384
    rec = measure(side, args.steps, args.warmup, 0)
    print(json.dumps({
        "metric": "replan latency ms after 1%% of cells change near the rover, MFD* %dx%d" % (side, side),
        "value": rec["replan_ms"]["mixed"], "unit": "ms", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": rec["replan_ms"]["mixed"], "higher_is_better": False, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "cfg3", "planner": rec["planner"], "grid": rec["grid"], "patch": rec["patch"],
                   "rover": rec["rover"], "timing": rec["timing"]},
        "initial_plan": {"ms": rec["plan_ms"], "mcells_per_s": rec["mcells_per_s"],
                         "roofline_frac": rec["roofline_frac"]},
        "replan": rec["detail"], "verified": rec["verified"]}), flush=True)
