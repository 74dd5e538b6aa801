#!/bin/bash
# round 2, visit O (2 GPUs): sharded tests, the bands record at N = 2 with the new defaults (band factor by tiles per GPU / world, bands by grid size), MFD* cfg3 with the CTA list for the initial plan
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_sharded.py tests/test_bands.py tests/test_config_sizes.py -m gpu -x -q > gpurun_out/pytest_sharded.log 2>&1; echo "pytest sharded rc=$?"; tail -n 3 gpurun_out/pytest_sharded.log
timeout 600 python bench.py --workload cfg3 --steps 3 --warmup 1 > gpurun_out/cfg3.json 2> gpurun_out/cfg3.err; echo "cfg3 rc=$?"; cut -c1-900 gpurun_out/cfg3.json
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --workload bands --side 16384 --steps 3 --warmup 1 > gpurun_out/bands_16384_n2.json 2> gpurun_out/bands_n2.err; echo "bands n2 rc=$?"; cut -c1-1800 gpurun_out/bands_16384_n2.json
