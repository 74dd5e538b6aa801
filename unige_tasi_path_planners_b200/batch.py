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


def make_queries(n_maps, per_map, side, seed=42, maps=None):
    """(map index, start, goal) triples, endpoints drawn uniformly from the FREE cells of their map (SURVEY.md 8d;
    rejection sampling with a seeded generator, so the batch is reproducible). `maps`: {index: uint8 array} of the maps
    the queries run on; by default the generator's maps 1000 + index are built here."""
    rng = random.Random(seed)
    out = []
    for m in range(n_maps):
        cmap = maps[m] if maps is not None and m in maps else synth.synth_map(1000 + m, side, side, 10, None, None)

        def free_cell():
            while True:
                x, y = rng.randrange(8, side - 8), rng.randrange(8, side - 8)
                if cmap[x, y] != 255:
                    return (float(x), float(y))

        for _ in range(per_map):
            s = free_cell()
            g = free_cell()
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


def stack_maps(maps):
    """Independent maps of equal shape -> one tall map with an obstacle row (255) between consecutive maps.
    Element rows of map i start at i * (L + 1), for cell fields (L rows + the obstacle row) and for node fields
    (the L + 1 node rows of a map are contiguous and the next map's first node row follows directly)."""
    L, W = maps[0].shape
    out = np.full((len(maps) * (L + 1) - 1, W), 255, np.uint8)
    for i, m in enumerate(maps):
        out[i * (L + 1):i * (L + 1) + L] = m
    return out


class StackedBatch:
    """A batch of independent (map, start, goal) queries planned as ONE multi-goal plan on the stacked map: all the
    wavefronts share one launch of the persistent solver and keep the whole GPU busy (tasi_set_goals)."""

    def __init__(self, kind, level=1, device=0):
        self.kind = kind
        self.p = _planner.PLANNERS[kind](level, device=device)
        self.p.set_occupancy_threshold(1.0)
        self.L = self.W = self.n = 0

    def plan(self, maps, goals, occ=1.0):
        """maps: list of (L, W) uint8 arrays (repeat an array to ask several goals on one map); goals: (x, y) per map."""
        self.L, self.W = maps[0].shape
        self.n = len(maps)
        self.p.set_occupancy_threshold(occ)
        distinct, index, seen = [], [], {}
        for m in maps:  # the same array object asked several times crosses the bus once
            if id(m) not in seen:
                seen[id(m)] = len(distinct)
                distinct.append(m)
            index.append(seen[id(m)])
        self.p.set_map_stack(distinct, index)
        off = self.L + 1
        self.p.set_start(0.0, 0.0)
        self.goals = [tuple(g) for g in goals]
        self.p.set_goals([(g[0] + i * off, g[1]) for i, g in enumerate(goals)])
        rc = self.p.step()
        assert rc == 0, rc

    def field(self, i):
        """Converged field of query i (what a planner holding only that map would return)."""
        ex = self.L if self.kind == _planner.DFM else self.L + 1
        ey = self.W if self.kind == _planner.DFM else self.W + 1
        return self.p.read_window(i * (self.L + 1), 0, ex, ey)

    def window(self, i, x0, y0, nx, ny, out=None):
        """Hand-off window of query i in that map's own coordinates (INF outside the map)."""
        ex = self.L if self.kind == _planner.DFM else self.L + 1
        xa, xb = max(x0, 0), min(x0 + nx, ex)
        w = out if out is not None else np.empty((nx, ny), np.float32)
        w[:] = np.inf
        if xa < xb:
            w[xa - x0:xb - x0] = self.p.read_window(i * (self.L + 1) + xa, y0, xb - xa, ny)
        return w

    def paths(self, starts, max_steps=20, lookahead=True, allow_indirect=True, with_costs=False):
        """One walk per query from its start to its goal (map coordinates), all in one kernel launch; each
        walk sees only its own block of rows (row_ranges), so it is the walk a planner holding only that map gives."""
        n = len(self.goals)
        rr = [(i * (self.L + 1), self.L) for i in range(n)]
        res, ms = self.p.extract_paths(starts, self.goals, max_steps, lookahead, allow_indirect,
                                       with_costs=with_costs, row_ranges=rr)
        return res, ms

    def close(self):
        self.p.close()


def measure(rank, world, local, side=1024, full=False, n_maps=None, per_map=None):
    """The cfg5 record: a batch of independent (map, start, goal) queries, whole maps per rank, planned as stacked
    multi-goal plans; wall clock of the whole batch, max over ranks. Collective: every rank of an initialised
    torch.distributed group must call it. Returns (per-planner results, config dict)."""
    import torch
    import torch.distributed as dist
    lanes = int(os.environ.get("TASI_BATCH_LANES", "4"))
    if n_maps is None:
        n_maps, per_map = (64, 64) if full else (8 * world, 16)
    # every rank builds every map of the batch (the draw of the endpoints needs them and has to be the same everywhere)
    all_maps = {m: synth.synth_map(1000 + m, side, side, 10, None, None) for m in range(n_maps)}
    queries = make_queries(n_maps, per_map, side, maps=all_maps)
    mine = [q for q in queries if q[0] % world == rank]  # whole maps per rank: a map is generated once per lane
    results = {}
    stacked = os.environ.get("TASI_BATCH_MODE", "stacked") == "stacked"
    chunk = int(os.environ.get("TASI_BATCH_CHUNK", "256"))
    host_maps = {m: all_maps[m] for m in sorted({q[0] for q in mine})} if stacked else {}
    for kind, name in ((_planner.FD, "FD*"), (_planner.DFM, "MS-DFM")):
        if stacked:
            sb = StackedBatch(kind, 1, local)
            buf = np.empty((64, 64), np.float32)

            walk_steps = int(os.environ.get("TASI_BATCH_WALK_STEPS", "20"))  # 0 = fields and windows only
            walk = {"ms": 0.0, "points": 0, "at_goal": 0}

            def run_chunks(qs):
                reached = launches = 0
                for c0 in range(0, len(qs), chunk):
                    part = qs[c0:c0 + chunk]
                    sb.plan([host_maps[q[0]] for q in part], [q[2] for q in part])
                    launches += sb.p.stats()["solver_launches"]
                    if walk_steps:
                        res, ms = sb.paths([q[1] for q in part], walk_steps)
                        walk["ms"] += ms
                        walk["points"] += sum(len(r["path"]) for r in res)
                        walk["at_goal"] += sum(r["status"] == 0 for r in res)
                        launches += 1
                    for i, q in enumerate(part):
                        sb.window(i, int(q[1][0]) - 32, int(q[1][1]) - 32, 64, 64, out=buf)
                        reached += bool(np.isfinite(buf[32, 32]))
                return reached, launches

            run_chunks(mine[:8])  # warm-up
            walk.update(ms=0.0, points=0, at_goal=0)
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
            # the sample is short (tens of ms): the batch is run three times and the median is reported
            dts = []
            for _ in range(3):
                walk.update(ms=0.0, points=0, at_goal=0)
                torch.cuda.synchronize()
                if world > 1:
                    dist.barrier()
                t0 = time.perf_counter()
                reached, launches = run_chunks(mine)
                torch.cuda.synchronize()
                dts.append(time.perf_counter() - t0)
            dt = sorted(dts)[1]
            sb.close()
            tm = torch.tensor([dt], dtype=torch.float64, device="cuda")
            cnt = torch.tensor([len(mine), reached, launches], dtype=torch.int64, device="cuda")
            if world > 1:
                dist.all_reduce(tm, op=dist.ReduceOp.MAX)
                dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
            results[name] = {"queries": int(cnt[0]), "reached": int(cnt[1]), "seconds": float(tm[0]),
                             "queries_per_s": int(cnt[0]) / float(tm[0]),
                             "mcells_per_s": int(cnt[0]) * side * side / float(tm[0]) / 1e6,
                             "gpu_launches": int(cnt[2]),
                             "walks": {"max_steps": walk_steps, "ms_total_rank0": walk["ms"],
                                       "points_rank0": walk["points"], "at_goal_rank0": walk["at_goal"]}}
            continue
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
    config = {"workload": "cfg5", "maps": n_maps, "queries_per_map": per_map, "grid": "%dx%d" % (side, side),
              "endpoints": "drawn from free cells (seed 42)",
              "mode": "stacked: %d queries per multi-goal plan" % chunk if stacked else "lanes",
              "lanes_per_gpu": lanes, "solver_ctas_per_lane": max(1, 444 // lanes),
              "sharding": "whole maps per rank, no collective on the data path",
              "timing": "wall clock of the whole batch: H2D of the (stacked) maps from host memory, goal "
                        "upload, plan, one device walk per query (TASI_BATCH_WALK_STEPS, default 20 steps as the "
                        "reference extractor) and D2H of the 64x64 hand-off window per query; median of 3 runs, max over ranks"}
    return results, config


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
    results, config = measure(rank, world, local, side, args.full)
    if rank == 0:
        print(json.dumps({
            "metric": "Mcells/s over a batch of independent start/goal queries on %dx%d maps (FD* and MS-DFM)" % (side, side),
            "value": results["MS-DFM"]["mcells_per_s"], "unit": "Mcells/s", "n_gpus": world, "steps": 1, "warmup": 1,
            "ms_per_step": results["MS-DFM"]["seconds"] * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config,
            "per_planner": results}), flush=True)
    if world > 1:
        dist.destroy_process_group()
