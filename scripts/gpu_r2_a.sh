#!/bin/bash
# round 2, visit A: the whole GPU suite (with the new config-size tests), the bench line with its new records.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q --durations=15 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
timeout 600 python scripts/quick_bench.py 1024 4096 16384 > gpurun_out/quick.log 2>&1; echo "quick rc=$?"
tail -25 gpurun_out/pytest_gpu.log; cut -c1-300 gpurun_out/quick.log; cut -c1-3000 gpurun_out/bench.json; tail -5 gpurun_out/bench.err
