"""Batches of independent start/goal queries (BASELINE.json cfg5; SURVEY.md 8e: "replicas only", no collective).

A query = (map, start, goal): set_goal on a planner that already holds the map forces a fresh plan
(ReplannerBase.h:48-55), then the hand-off window around the start is read back. Queries are sharded across
ranks (one slice per GPU) and, inside a rank, across `lanes` planners that run side by side — each on its own
CUDA stream with a share of the persistent CTAs (tasi_set_solver_blocks), driven by one host thread each
(ctypes drops the GIL during the calls). A 1024^2 plan has at most a few dozen active tiles at a time, so lanes
fill SMs that a single query would leave idle.
"""
import json
import os
import random
import threading
import time

import numpy as np

from . import planner as _planner
from . import synth


def make_queries(n_maps, per_map, side, seed=42):
    """(map index, start, goal) triples: endpoints drawn uniformly, 9x9 blocks around them are cleared per map."""
    rng = random.Random(seed)
    out = []
    for m in range(n_maps):
        for _ in range(per_map):
            s = (float(rng.randrange(8, side - 8)), float(rng.randrange(8, side - 8)))
            g = (float(rng.randrange(8, side - 8)), float(rng.randrange(8, side - 8)))
            out.append((m, s, g))
    return out


def run_lane(kind, level, device, blocks, side, queries, out, idx, win=64):
    """One planner working through its queries; consecutive queries on the same map keep the map on the device."""
    p = _planner.PLANNERS[kind](level, device=device)
    p.set_occupancy_threshold(1.0)
    p.set_solver_blocks(blocks)
    buf = np.empty((win, win), np.float32)
    cur_map = None
    reached = 0
    visits = 0
    launches = 0
    for m, s, g in queries:
        if m != cur_map:
            p.synth_map(1000 + m, side, side, 0, 10, None, None)
            cur_map = m
        p.set_start(*s)
        p.set_goal(*g)
        p.reset()
        rc = p.step()
        assert rc == 0, rc
        p.read_window(int(s[0]) - win // 2, int(s[1]) - win // 2, win, win, out=buf)
        reached += bool(np.isfinite(buf[win // 2, win // 2]))
        st = p.stats()
        visits += st["tile_visits"]
        launches += st["solver_launches"]
    p.close()
    out[idx] = (reached, visits, launches)


def bench_main(args, rank, world):
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    side = 1024 if args.side == 4096 else args.side
    lanes = int(os.environ.get("TASI_BATCH_LANES", "4"))
    n_maps, per_map = (64, 64) if args.full else (8 * world, 16)
    queries = make_queries(n_maps, per_map, side)
    mine = [q for q in queries if q[0] % world == rank]  # whole maps per rank: a map is generated once per lane
    results = {}
    for kind, name in ((_planner.FD, "FD*"), (_planner.DFM, "MS-DFM")):
        maps = sorted({q[0] for q in mine})
        lane_q = [[q for q in mine if maps.index(q[0]) % lanes == l] for l in range(lanes)]
        blocks = max(1, 444 // lanes)

        def run_all():
            out = [None] * lanes
            th = [threading.Thread(target=run_lane, args=(kind, 1, local, blocks, side, lane_q[l], out, l))
                  for l in range(lanes) if lane_q[l]]
            for t in th:
                t.start()
            for t in th:
                t.join()
            return [o for o in out if o]

        run_lane(kind, 1, local, blocks, side, mine[:2], [None], 0)  # warm-up: library, streams, allocator
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        out = run_all()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        tm = torch.tensor([dt], dtype=torch.float64, device="cuda")
        cnt = torch.tensor([len(mine), sum(o[0] for o in out), sum(o[2] for o in out)], dtype=torch.int64, device="cuda")
        if world > 1:
            dist.all_reduce(tm, op=dist.ReduceOp.MAX)
            dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
        results[name] = {"queries": int(cnt[0]), "reached": int(cnt[1]), "seconds": float(tm[0]),
                         "queries_per_s": int(cnt[0]) / float(tm[0]),
                         "mcells_per_s": int(cnt[0]) * side * side / float(tm[0]) / 1e6,
                         "gpu_launches": int(cnt[2])}
    if rank == 0:
        print(json.dumps({
            "metric": "Mcells/s over a batch of independent start/goal queries on %dx%d maps (FD* and MS-DFM)" % (side, side),
            "value": results["MS-DFM"]["mcells_per_s"], "unit": "Mcells/s", "n_gpus": world, "steps": 1, "warmup": 1,
            "ms_per_step": results["MS-DFM"]["seconds"] * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "cfg5", "maps": n_maps, "queries_per_map": per_map, "grid": "%dx%d" % (side, side),
                       "lanes_per_gpu": lanes, "solver_ctas_per_lane": max(1, 444 // lanes),
                       "sharding": "whole maps per rank, no collective on the data path",
                       "timing": "wall clock of the whole batch incl. map generation on the device, goal upload, "
                                 "plan and D2H of the 64x64 hand-off window per query; max over ranks"},
            "per_planner": results}), flush=True)
    if world > 1:
        dist.destroy_process_group()
