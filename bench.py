"""bench.py — the headline measurement (BASELINE.json): Mcells/s to convergence of an initial plan.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload NAME]

Workload at every N (config.workload = "cfg2"): MS-DFM (DFMPlanner level 1) initial plan on the 4096 x 4096
synthetic traversability grid with 10 % random obstacles of SURVEY.md 8(d) (seed 1 + rank, start (5,5), goal
(4090,4090), occupancy threshold 1.0). One step = one plan to convergence (reset + step). With N > 1 every
rank plans its own grid of the same shape — the "independent queries, one slice per GPU" sharding of the
north star, no data-path collective — so scaling is weak and value = N * cells / max-over-ranks time.
The same JSON line carries the rest of BASELINE.json's metric, each measured by its own module outside cfg2's timed
region: `replan` (cfg3: MFD* 16384^2 plan + replan latency after a 1 % patch, N = 1), `batch` (cfg5: independent queries
on 1024^2 maps, sharded over the ranks) and `bands` (cfg4 shape: ONE FD* grid, 16384^2 and 65536^2, in row bands over
the N GPUs through the sharded planner — NCCL halo exchange, strong scaling, verified by checksum against an unsharded
plan). `--workload bands|cfg3|cfg5` print one of them as a line of its own; `--no-extras` leaves them out.

Timing: W warm-up steps, then K steps; each step's device time is taken with CUDA events recorded on the
planner's own stream inside tasi_step (stats.step_ms_device: field fill + seeding + persistent solver), L2 is
flushed between steps by writing a 256 MiB buffer (the 80 MiB working set would otherwise stay in the 126 MB
L2), ranks are barrier'ed + synchronised on both sides, and the max over ranks of the summed step time is used.
The e2e leg times the public API with HOST buffers by wall clock: set_map (H2D of the 16 MiB grid from pinned
memory) + set_start + set_goal + step + read_window around the start (the hand-off the path extractor
consumes), every step.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

# rank 0 prints ONE JSON line on stdout. NCCL writes its version banner and its INFO lines to file descriptor 1 as well
# (NCCL_DEBUG from the environment or from /etc/nccl.conf), so descriptor 1 is pointed at stderr for the whole run and
# Python's sys.stdout (the JSON line) writes to a copy of the original stdout. An NCCL_DEBUG the caller has set is
# left as it is (the driver's INFO setting is its evidence of the ranks); only an unset one becomes WARN.
if "NCCL_DEBUG" not in os.environ:
    os.environ["NCCL_DEBUG"] = "WARN"
sys.stdout.flush()
_JSON_FD = os.dup(1)
os.dup2(2, 1)
sys.stdout = os.fdopen(_JSON_FD, "w")  # Python-level prints (the JSON line of every workload) keep the real stdout


def emit(line):
    print(json.dumps(line), flush=True)

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "Mcells/s to convergence (initial plan), MS-DFM 4096x4096, 10% obstacles"
UNIT = "Mcells/s"
N_SIDE = 4096
P_OBS = 10
KIND_DFM = 1
ALGO_BYTES_PER_CELL = 5.0  # 1 B traversability read + 4 B converged value written
REF_SAMPLE_SIDE = 1536     # bounded CPU sample: the reference plans ~0.36 Mcells/s on one core


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=["cfg2", "cfg3", "cfg5", "bands"])
    ap.add_argument("--side", type=int, default=N_SIDE, help="grid side (development only; the contract is 4096)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true",
                    help="cfg2 only: skip the replan (cfg3), batch (cfg5) and bands (cfg4) records of the line")
    ap.add_argument("--full", action="store_true", help="cfg5: the full 64 maps x 64 queries instead of a sample")
    return ap.parse_args()


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def recorded_traffic(side):
    """dram bytes per solver launch from the committed ncu capture (profiles/traffic.json), or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            t = json.load(f)
        e = t.get("dfm_%d" % side)
        return float(e["dram_bytes_per_launch"]) if e else None
    except Exception:
        return None


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md, clocks line)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except Exception:
                continue
            for n, v in zip(names, r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---- the CPU side: the unmodified reference (oracle/_ref) or, without it, the oracle port ------------------
def cpu_plan_once(side, seed, fast=True):
    """One MS-DFM initial plan on the CPU on a side x side grid of the bench's generator.
    Returns (seconds, kind, build): kind "reference" = oracle/_ref (the reference's own DFMPlanner<1>, its p_time),
    "port" = oracle/tasi_oracle.c heap solver (wall clock). build = which of the two reference builds ran: the timing
    build (the reference's Release flags, oracle/Makefile REF_FAST_FLAGS) when this CPU can run it, else the parity
    build (-ffp-contract=off)."""
    from oracle import oracle_py as O  # CPU baseline leg: the one place bench.py executes oracle/
    from unige_tasi_path_planners_b200 import synth
    start, goal = synth.default_endpoints(side, side)
    cmap = synth.synth_map(seed, side, side, P_OBS, start, goal)
    if O.ref_available():
        use_fast = bool(fast and O.ref_fast_available(KIND_DFM))
        r = O.RefPlanner(KIND_DFM, 1, fast=use_fast)
        r.set_map(cmap, 1.0)
        r.set_start(*start)
        r.set_goal(*goal)
        rc = r.step()
        u_ms, p_ms = r.times()
        r.close()
        assert rc == 0
        return (u_ms + p_ms) / 1e3, "reference", ("-O2 -march=x86-64-v3 -DNDEBUG" if use_fast else
                                                  "-O2 -ffp-contract=off -DNDEBUG")
    t0 = time.perf_counter()
    O.solve(KIND_DFM, cmap, goal, method="heap")
    return time.perf_counter() - t0, "port", "gcc -O2 -ffp-contract=off"


def cpu_sample_desc(side, kind, build):
    what = ("the reference's DFMPlanner<1> (oracle/_ref, built from the unmodified sources with %s, NO_HEURISTIC), "
            "its own u_time+p_time" % build) if kind == "reference" else \
           "oracle/tasi_oracle.c heap solver (CPU restatement, %s), wall clock" % build
    return "MS-DFM initial plan on a %dx%d grid of the same generator (seed 1, 10%% obstacles), 1 thread: %s" % (
        side, side, what)


def run_reference(args, rank):
    """--impl reference: the reference's CPU planner on the host cores on the SAME workload as the B200 arm: the full
    4096 x 4096 grid of cfg2. The reference is single-threaded, so `cores` is 1; rank 0 alone runs.
    A plan takes 45-75 s on one core, so min(steps, 3) full-size plans are timed (fewer if 150 s have passed) and the
    line says how many; the rate of a bounded 1536^2 sample is reported beside it."""
    if rank != 0:
        return
    side = args.side
    kind, build = "reference", ""
    for _ in range(max(args.warmup, 0) and 1):  # one warm-up plan is enough to page the libraries in
        _, kind, build = cpu_plan_once(256, 1)
    dt_s, _, _ = cpu_plan_once(REF_SAMPLE_SIDE, 1)
    tot, n = 0.0, 0
    t_begin = time.perf_counter()
    for _ in range(max(1, min(args.steps, 3))):
        dt, kind, build = cpu_plan_once(side, 1)
        tot += dt
        n += 1
        if time.perf_counter() - t_begin > 150.0:
            break
    ms = tot / n * 1e3
    val = side * side / (ms / 1e3) / 1e6
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "cfg2", "planner": "MS-DFM (DFMPlanner level 1)", "grid": "%dx%d" % (side, side),
                       "p_obs": P_OBS, "seed": 1, "plans_timed": n,
                       "note": "the full cfg2 grid, %d plan(s) timed out of the %d steps asked for (a plan takes about a "
                               "minute on one core)" % (n, args.steps),
                       "sample_1536": {"mcells_per_s": REF_SAMPLE_SIDE ** 2 / dt_s / 1e6, "seconds": dt_s}},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": 1, "kind": kind,
                             "sample": cpu_sample_desc(side, kind, build), "host_cpus": os.cpu_count()},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


def cpu_baseline():
    """Bounded CPU sample timed beside the GPU number (rank 0, N = 1)."""
    cpu_plan_once(128, 1)
    dt, kind, build = cpu_plan_once(REF_SAMPLE_SIDE, 1)
    return {"value": REF_SAMPLE_SIDE * REF_SAMPLE_SIDE / dt / 1e6, "unit": UNIT, "cores": 1, "kind": kind,
            "sample": cpu_sample_desc(REF_SAMPLE_SIDE, kind, build), "seconds": dt, "host_cpus": os.cpu_count()}


# ---- the B200 arm --------------------------------------------------------------------------------------------
def run_b200(args, rank, world):
    import numpy as np
    import torch
    import torch.distributed as dist
    from unige_tasi_path_planners_b200 import DFMPlanner, synth

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    side = args.side
    cells = side * side
    start, goal = synth.default_endpoints(side, side)
    seed = 1 + rank

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    p = DFMPlanner(1, device=local)
    p.set_occupancy_threshold(1.0)
    p.synth_map(seed, side, side, 0, P_OBS, start, goal)  # generated on the device: resident in HBM
    p.set_start(*start)
    p.set_goal(*goal)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def one_step():
        flush.fill_(1)           # evict the working set from L2
        torch.cuda.synchronize()
        p.reset()
        rc = p.step()
        assert rc == 0, rc
        return p.stats()

    for _ in range(args.warmup):
        one_step()
    sampler = ClockSampler(local) if rank == 0 else None
    barrier()
    if sampler:
        sampler.start()
    dev_ms = solve_ms = 0.0
    launches = 0
    visits = evals = 0
    t_wall0 = time.perf_counter()
    for _ in range(args.steps):
        st = one_step()
        dev_ms += st["step_ms_device"]
        solve_ms += st["kernel_ms_solve"]
        launches += st["solver_launches"]
        visits += st["tile_visits"]
        evals += st["elem_evals"]
    barrier()
    wall_ms = (time.perf_counter() - t_wall0) * 1e3
    clocks = sampler.stop() if sampler else None
    tiles = st["tiles"]

    # ---- e2e: the public API with host buffers, wall clock, every step H2D map + D2H hand-off window ----
    host_map = torch.from_numpy(synth.synth_map(seed, side, side, P_OBS, start, goal)).pin_memory()
    host_np = host_map.numpy()
    win = 64
    out = torch.empty((win, win), dtype=torch.float32).pin_memory()
    out_np = out.numpy()

    def e2e_step():
        flush.fill_(1)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        p.set_map(host_np)                      # H2D, 1 B/cell (set_map forces a re-plan of the new grid)
        p.set_start(*start)
        p.set_goal(*goal)
        p.reset()
        rc = p.step()
        p.read_window(int(start[0]) - win // 2, int(start[1]) - win // 2, win, win, out=out_np)  # D2H hand-off
        dt = time.perf_counter() - t0
        assert rc == 0
        return dt

    e2e_step()
    barrier()
    e2e_s = 0.0
    for _ in range(args.steps):
        e2e_s += e2e_step()
    barrier()
    assert np.isfinite(out_np[win // 2, win // 2]), "start not reached"
    t0 = time.perf_counter()
    p.read_field()
    full_read_ms = (time.perf_counter() - t0) * 1e3

    tm = torch.tensor([dev_ms, e2e_s * 1e3, wall_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
    dev_ms_max, e2e_ms_max, wall_ms_max = [float(v) for v in tm.tolist()]
    p.close()
    del flush, host_map, out
    torch.cuda.empty_cache()

    # ---- the rest of BASELINE.json's metric in the same line: replan latency (cfg3, N = 1), the query batch (cfg5)
    # and the row-band sharded grid (cfg4 shape, strong scaling) — each measured by its own module, outside cfg2's
    # timed region
    extras = {}
    if not args.no_extras and side == N_SIDE:
        peak_gbs, _ = measured_peak_gbs()
        if world == 1:
            from unige_tasi_path_planners_b200 import replan_bench
            extras["replan"] = replan_bench.measure(16384, max(1, min(args.steps, 5)), 1, local, peak_gbs,
                                                    variants=("mixed", "raise", "lower"))
            torch.cuda.empty_cache()
        from unige_tasi_path_planners_b200 import batch
        res, cfg = batch.measure(rank, world, local, 1024, False)
        extras["batch"] = {"config": cfg, "queries_per_s": {k: v["queries_per_s"] for k, v in res.items()},
                           "mcells_per_s": {k: v["mcells_per_s"] for k, v in res.items()},
                           "verified": all(v["reached"] == v["queries"] for v in res.values()),
                           "verified_how": "every query's start has a finite value (endpoints are free cells)",
                           "per_planner": res}
        torch.cuda.empty_cache()
        from unige_tasi_path_planners_b200 import sharded_bench
        sides = tuple(int(v) for v in os.environ.get("TASI_BENCH_BAND_SIDES", "16384,65536").split(",") if v)
        extras["bands"] = sharded_bench.measure(rank, world, local, steps=max(1, min(args.steps, 3)), warmup=1, sides=sides)
        torch.cuda.empty_cache()
    if rank == 0:
        ms_per_step = dev_ms_max / args.steps
        value = world * cells / (ms_per_step / 1e3) / 1e6
        e2e_val = world * cells / (e2e_ms_max / args.steps / 1e3) / 1e6
        peak, peak_src = measured_peak_gbs()
        k_ms = solve_ms / args.steps  # the persistent solver launch (plus its 2 tiny queue-arming launches)
        achieved = ALGO_BYTES_PER_CELL * cells / (k_ms / 1e3) / 1e9
        traffic = recorded_traffic(side)
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": "cfg2", "planner": "MS-DFM (DFMPlanner level 1)",
                           "grid": "%dx%d" % (side, side), "p_obs": P_OBS, "seed": "1+rank",
                           "start": list(start), "goal": list(goal), "sharding": "one independent grid per GPU",
                           "l2": "flushed between steps (256 MiB write)", "timing": "CUDA events on the planner "
                           "stream inside tasi_step; wall %.2f ms/step" % (wall_ms_max / args.steps)},
                "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": cells,
                        "d2h_bytes_per_step": win * win * 4, "ms_per_step": e2e_ms_max / args.steps,
                        "what": "set_map(host) + set_start + set_goal + step + read_window(64x64 at start): the hand-off "
                                "the path extractor consumes; GpuPathExtractor (the default hand-off of the C++ classes) "
                                "walks on the device and only the path crosses the bus. Reading the WHOLE field back "
                                "costs full_field_d2h_ms on top",
                        "full_field_d2h_ms": full_read_ms},
                "gpu_launches": launches,
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                             "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                             "kernel": "solver_kernel<DFM,0>", "kernel_ms": k_ms,
                             "algorithmic_bytes_per_launch": ALGO_BYTES_PER_CELL * cells},
                "work": {"tiles": tiles, "tile_visits_per_tile": visits / args.steps / tiles,
                         "elem_evals_per_cell": evals / args.steps / cells},
                "clocks": clocks}
        line.update(extras)
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline()
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if args.workload == "bands":
        from unige_tasi_path_planners_b200 import sharded_bench
        sharded_bench.bench_main(args, rank, world)
        return
    if args.workload == "cfg5":
        from unige_tasi_path_planners_b200 import batch
        batch.bench_main(args, rank, world)
        return
    if args.workload == "cfg3":
        from unige_tasi_path_planners_b200 import replan_bench
        replan_bench.bench_main(args, rank, world)
        return
    run_b200(args, rank, world)


if __name__ == "__main__":
    main()
