#!/bin/bash
# round 2: ncu evidence at HEAD (1 GPU). (1) launch list of the cfg2 bench command (extras left out: the list is capped at 400
# launches) after the same command exited 0 without ncu; (2) full capture of one solver launch of cfg2 (MS-DFM 4096^2);
# (3) full capture of one FD* 16384^2 launch (the CTA-list instance).
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-extras"
timeout 600 $CMD > gpurun_out/plain.json 2> gpurun_out/plain.err; echo "plain rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches_cfg2.csv $CMD > gpurun_out/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
CMD1="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-extras"
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:solver_kernel -s 1 -c 1 -f -o gpurun_out/r02_prof_solver_dfm4096 $CMD1 > gpurun_out/ncu_full.log 2>&1; echo "ncu full dfm rc=$?"
QUICK_KINDS=0 timeout 300 python scripts/quick_bench.py 16384 > gpurun_out/plain_fd.json 2>&1; echo "plain fd rc=$?"
QUICK_KINDS=0 timeout 1500 ncu --set full --clock-control none --import-source on -k regex:solver_kernel -s 1 -c 1 -f -o gpurun_out/r02_prof_solver_fd16384 python scripts/quick_bench.py 16384 > gpurun_out/ncu_full_fd.log 2>&1; echo "ncu full fd rc=$?"
ls -la gpurun_out/*.ncu-rep; cut -c1-600 gpurun_out/plain.json
