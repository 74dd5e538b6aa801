#!/bin/bash
# round 2, visit E (8 GPUs): the fused sharded planner on 65536^2, bands per GPU swept.
mkdir -p gpurun_out
: > gpurun_out/bands_n8_fused.jsonl
for b in 4 8 16; do
  TASI_BANDS_PER_GPU=$b timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519 \
    bench.py --gpus 8 --workload bands --side 65536 --steps 2 --warmup 1 >> gpurun_out/bands_n8_fused.jsonl 2> gpurun_out/bands_n8_fused.err
  echo "bands=$b rc=$?"
done
cut -c1-2200 gpurun_out/bands_n8_fused.jsonl; grep -v "^\[W\|^W1\|^\*\*\*\|OMP_NUM" gpurun_out/bands_n8_fused.err | tail -n 5
