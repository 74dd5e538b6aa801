"""bench.py's `bands` record and `--workload bands`: ONE FD* grid in row bands over the GPUs of the job (BASELINE.json
cfg4 shape), strong scaling, through the sharded planner of the C ABI (tasi_create_sharded_rank, csrc/tasi_shard.cuh):
one process per GPU, the 128-byte NCCL id is broadcast with torch.distributed, everything else — halo rows by
ncclSend/ncclRecv, convergence quantities by ncclAllReduce, the common value limit — runs inside libtasi_gpu.so on the
planner's stream.

Every timed configuration is VERIFIED: the checksum of the sharded field (tasi_field_checksum, summed over the ranks by
NCCL) must equal the checksum of an unsharded plan of the same grid computed on rank 0's GPU — bit-identity, the contract
of SURVEY.md 8e. At N = 1 the record's `ms` is the unsharded single-device plan (the honest baseline of the strong-scaling
curve); the sharded path with the same number of bands per GPU is run beside it and verified too.
"""
import json
import os
import time

from . import planner as _planner
from . import synth


def broadcast_unique_id(dist, rank, world, device=None):
    """Rank 0 draws the NCCL id (tasi_nccl_unique_id), torch.distributed hands it to every rank (any backend)."""
    import torch
    if world == 1:
        return None
    t = torch.zeros(128, dtype=torch.uint8, device=device if device is not None else "cpu")
    if rank == 0:
        uid = _planner.ShardedPlanner.nccl_unique_id()
        t.copy_(torch.frombuffer(bytearray(uid), dtype=torch.uint8))
    dist.broadcast(t, 0)
    return bytes(t.cpu().numpy().tobytes())


def measure_side(rank, world, local, side, steps, warmup, kind=_planner.FD, level=1, seed=3):
    """One grid size. Collective over the ranks. Returns the record (valid on every rank)."""
    import torch
    import torch.distributed as dist
    start, goal = synth.default_endpoints(side, side)
    dev = torch.device("cuda", local)
    uid = broadcast_unique_id(dist, rank, world, dev)  # an NCCL id makes ONE communicator: a fresh one per planner

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    p = _planner.ShardedPlanner(kind, level, rank=rank, n_ranks=world, unique_id=uid, device=local)
    p.set_occupancy_threshold(1.0)
    p.synth_map(seed, side, side, 0, 10, start, goal)
    if world > 1 and os.environ.get("TASI_SHARD_FUSED", "1") != "0":
        p.connect_peers(dist, dev)  # CUDA IPC handles of every rank's arrays: the fused mode (else NCCL iterations)
    p.set_start(*start)
    p.set_goal(*goal)
    for _ in range(max(warmup, 1)):
        p.reset()
        assert p.step() == 0
    barrier()
    t0 = time.perf_counter()
    solver_ms = dev_ms = 0.0
    exchanges = visits = launches = 0
    for _ in range(steps):
        p.reset()
        rc = p.step()
        assert rc == 0, rc
        ss, st = p.shard_stats(), p.stats()
        solver_ms += ss["solver_ms_max_rank"]
        dev_ms += st["step_ms_device"]
        exchanges = ss["exchanges"]
        visits += st["tile_visits"]
        launches += st["solver_launches"]
    barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3 / steps
    tm = torch.tensor([wall_ms, dev_ms / steps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    wall_ms, dev_ms = float(tm[0]), float(tm[1])
    tiles = p.stats()["tiles"]
    cnt = torch.tensor([visits, tiles, launches], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    sharded_sum = p.field_checksum()  # collective: the whole field's checksum on every rank
    ss = p.shard_stats()
    p.close()
    torch.cuda.empty_cache()

    # the unsharded plan of the same grid on rank 0's GPU: the check, and at N = 1 the baseline time
    flag = torch.zeros(2, dtype=torch.float64, device=dev)
    if rank == 0:
        q = _planner.PLANNERS[kind](level, device=local)
        q.set_occupancy_threshold(1.0)
        q.synth_map(seed, side, side, 0, 10, start, goal)
        q.set_start(*start)
        q.set_goal(*goal)
        ms1 = []
        for _ in range((max(warmup, 1) + steps) if world == 1 else 1):
            q.reset()
            assert q.step() == 0
            ms1.append(q.stats()["step_ms_device"])
        flag[0] = 1.0 if q.field_checksum() == sharded_sum else 0.0
        flag[1] = sum(ms1[-steps:]) / len(ms1[-steps:])
        q.close()
        torch.cuda.empty_cache()
    if world > 1:
        dist.broadcast(flag, 0)
    verified, single_ms = bool(flag[0] > 0), float(flag[1])
    cells = side * side
    ms = single_ms if world == 1 else wall_ms
    rec = {"grid": "%dx%d" % (side, side), "planner": "FD* (FieldDPlanner level 1)", "n_gpus": world,
           "ms": ms, "mcells_per_s": cells / ms / 1e3, "scaling": "strong",
           "exchanges": int(exchanges), "solver_ms_max_rank": solver_ms / steps,
           "solver_share_of_wall": (solver_ms / steps) / wall_ms if wall_ms > 0 else None,
           "bands": ss["bands"], "bands_per_gpu": ss["bands_per_rank"], "rows_per_band": ss["rows_per_band"],
           "mode": "fused" if ss["fused"] else ("nccl iterations" if ss["nccl"] else "one rank"),
           "collective": ("none on the data path: a border tile's visit stores its boundary row into the neighbour rank's halo "
                          "row (peer-mapped memory over NVLink, CUDA IPC) and activates that rank's tiles with system-scope "
                          "atomics; thresholds advance globally through command words (rank 0 coordinates); NCCL only "
                          "sums the checksum") if ss["fused"] and world > 1 else
                         ("ncclSend/ncclRecv of the packed boundary rows to the two ring neighbours + ncclAllReduce(max) of "
                          "{halo lowered, -lowest parked value} per outer iteration, on the planner stream" if ss["nccl"]
                          else "none (one rank)"),
           "remote_activations": ss["remote_activations"],
           "busy_frac_min_max_rank": [ss["busy_frac_min_rank"], ss["busy_frac_max_rank"]],
           "tile_visits_per_tile": int(cnt[0]) / steps / max(int(cnt[1]), 1),
           "gpu_launches_per_plan": int(cnt[2]) / steps,
           "steps": steps, "verified": verified,
           "verified_how": "tasi_field_checksum of the sharded field (summed over ranks) == checksum of an unsharded plan "
                           "of the same grid on rank 0's GPU: bit-identical fields",
           "timing": "wall clock of reset + step between barriers, max over ranks; device time %.2f ms" % dev_ms}
    if world == 1:
        rec["ms_is"] = "the unsharded single-device plan (CUDA events): the baseline of the strong-scaling curve"
        rec["sharded_path_on_one_gpu_ms"] = wall_ms
    else:
        rec["single_gpu_unsharded_ms_rank0"] = single_ms
    return rec


def measure(rank, world, local, steps=3, warmup=1, sides=(16384, 65536)):
    """The `bands` record of the bench line: {side: record}. Collective."""
    import torch
    import torch.distributed as dist
    out = {}
    for side in sides:
        need = 5.3 * side * side * (1.0 + 1.2 / world)  # unsharded check on rank 0 + this rank's share, bytes
        free = torch.cuda.mem_get_info(local)[0]
        ok = torch.tensor([1.0 if free > need else 0.0], device=torch.device("cuda", local))
        if world > 1:
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if float(ok[0]) == 0.0:
            out[str(side)] = {"skipped": "needs %.0f GiB of free device memory" % (need / 2 ** 30)}
            continue
        k = steps if side <= 16384 else min(steps, 2)
        out[str(side)] = measure_side(rank, world, local, side, k, warmup)
    return out


def bench_main(args, rank, world):
    """`bench.py --workload bands [--side S]`: the bands record alone, as its own JSON line."""
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    side = args.side if args.side != 4096 else 16384
    rec = measure_side(rank, world, local, side, max(1, args.steps), args.warmup)
    if rank == 0:
        print(json.dumps({
            "metric": "Mcells/s to convergence (initial plan), FD* %dx%d in row bands" % (side, side),
            "value": rec["mcells_per_s"], "unit": "Mcells/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": rec["ms"], "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "bands (cfg4 shape)", "planner": rec["planner"], "grid": rec["grid"]},
            "bands": rec}), flush=True)
    if world > 1:
        dist.destroy_process_group()
