#!/bin/bash
# Full ncu capture of the persistent solver kernel (one launch), after the same command exited 0 without ncu.
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/plain_full.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:solver_kernel -s 1 -c 1 \
    -f -o gpurun_out/prof_solver $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"; tail -5 gpurun_out/ncu_full.log
