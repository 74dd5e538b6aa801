#!/bin/bash
# round 2, visit H (1 GPU): state of HEAD after the session restart: GPU tests, smoke, the default bench line (timed by wall clock), quick sizes
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
t0=$(date +%s)
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$? wall $(( $(date +%s) - t0 )) s"
timeout 600 python scripts/quick_bench.py 1024 4096 8192 16384 > gpurun_out/quick.log 2>&1; echo "quick rc=$?"
tail -3 gpurun_out/pytest_gpu.log; tail -2 gpurun_out/smoke.log; cut -c1-1500 gpurun_out/bench.json; tail -3 gpurun_out/bench.err
