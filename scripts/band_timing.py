"""Development probe (torchrun, 2+ GPUs): where an outer iteration of the ordered band solve spends its time."""
import os, sys, time
sys.path.insert(0, __file__.rsplit("/", 2)[0])
import torch, torch.distributed as dist
from unige_tasi_path_planners_b200 import bands, synth, planner
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
side = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
start, goal = synth.default_endpoints(side, side)
band = bands.GpuBand(planner.FD, 1, side, side, world, rank, device=local)
band.synth_map(3, 10, start, goal); band.set_endpoints(start, goal)
comm = bands.DistComm(dist, rank, world, device=torch.device("cuda", local))
for rep in range(2):
    T = dict(solve=0.0, exch=0.0, imp=0.0, red=0.0); n = 0
    band.begin(); step = 2 * comm.reduce_float(band.delta(), "max"); limit = step
    torch.cuda.synchronize(); dist.barrier(); t00 = time.perf_counter()
    while True:
        t0 = time.perf_counter(); parked = band.solve_until(limit); t1 = time.perf_counter()
        comm.exchange(band); t2 = time.perf_counter()
        low = (band.import_side(0) if band.open_top else 0) + (band.import_side(1) if band.open_bottom else 0); t3 = time.perf_counter()
        anyl, gmin = comm.lowered_and_min(low, parked); t4 = time.perf_counter()
        T["solve"] += t1 - t0; T["exch"] += t2 - t1; T["imp"] += t3 - t2; T["red"] += t4 - t3; n += 1
        if not anyl and gmin == float("inf"): break
        if gmin < float("inf"): limit = max(limit, gmin + step)
    tot = time.perf_counter() - t00
    print("rank", rank, "iters", n, "total ms %.1f" % (tot * 1e3), {k: round(v * 1e3, 1) for k, v in T.items()}, "kernel_ms %.1f" % band.stats()["kernel_ms_solve"], flush=True)
dist.destroy_process_group()
