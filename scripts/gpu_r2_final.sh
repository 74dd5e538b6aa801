#!/bin/bash
# round 2: evidence at HEAD (1 GPU): all GPU tests, smoke, the default bench line, the reference arm (3 full-size plans),
# quick sizes, ncu launch list of the cfg2 bench command + full captures (MS-DFM 4096^2, FD* 16384^2)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_full.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/pytest_gpu_full.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -n 1 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
timeout 600 python scripts/quick_bench.py 1024 4096 8192 16384 > gpurun_out/quick.log 2>&1; echo "quick rc=$?"
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-extras"
timeout 600 $CMD > gpurun_out/plain.json 2> gpurun_out/plain.err; echo "plain rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_cfg2.csv $CMD > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
CMD1="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-extras"
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:solver_kernel -s 1 -c 1 -f -o gpurun_out/r02_prof_solver_dfm4096 $CMD1 > gpurun_out/ncu_full.log 2>&1; echo "ncu full dfm rc=$?"
QUICK_KINDS=0 timeout 300 python scripts/quick_bench.py 16384 > gpurun_out/plain_fd.json 2>&1; echo "plain fd rc=$?"
QUICK_KINDS=0 timeout 1500 ncu --set full --clock-control none --import-source on -k regex:solver_kernel -s 1 -c 1 -f -o gpurun_out/r02_prof_solver_fd16384 python scripts/quick_bench.py 16384 > gpurun_out/ncu_full_fd.log 2>&1; echo "ncu full fd rc=$?"
cut -c1-300 gpurun_out/bench.json
