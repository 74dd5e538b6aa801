#!/bin/bash
# round 2, visit D (2 GPUs): the FUSED sharded planner — one process over both GPUs, then one process per GPU (CUDA IPC).
mkdir -p gpurun_out
PROBE_SETTINGS=1:1,2:1,4:1,8:1,16:1 timeout 600 python scripts/shard_probe.py 16384 0 2 > gpurun_out/shard_probe_fused_n2.jsonl 2> gpurun_out/shard_probe_fused_n2.err; echo "probe fused rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --workload bands --steps 2 --warmup 1 > gpurun_out/bands_fused_n2.json 2> gpurun_out/bands_fused_n2.err; echo "bands n2 rc=$?"
cat gpurun_out/shard_probe_fused_n2.jsonl; tail -n 3 gpurun_out/shard_probe_fused_n2.err; cut -c1-3000 gpurun_out/bands_fused_n2.json; grep -v "^\[W\|^W1\|^\*\*\*\|OMP_NUM" gpurun_out/bands_fused_n2.err | tail -n 8
