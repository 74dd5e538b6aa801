#!/bin/bash
# Records of one version for profiles/: parity suite (with durations), bench lines of cfg2 / cfg3 / cfg5, quick sizes,
# ncu launch list and one full capture of the solver (each after the same command exited 0 without ncu).
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
timeout 600 python bench.py --workload cfg3 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_cfg3.json 2> gpurun_out/bench_cfg3.err; echo "cfg3 rc=$?"
timeout 600 python bench.py --workload cfg5 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_cfg5.json 2> gpurun_out/bench_cfg5.err; echo "cfg5 rc=$?"
timeout 600 python scripts/quick_bench.py 1024 4096 8192 16384 > gpurun_out/quick.log 2>&1; echo "quick rc=$?"
timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
bash scripts/gpu_ncu_full.sh
ncu -i gpurun_out/prof_solver.ncu-rep --page raw --csv > gpurun_out/prof_solver_raw.csv 2>/dev/null
tail -12 gpurun_out/pytest_gpu.log; cat gpurun_out/bench.json; cut -c1-600 gpurun_out/bench_cfg3.json gpurun_out/bench_cfg5.json
