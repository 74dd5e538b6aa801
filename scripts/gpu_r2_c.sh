#!/bin/bash
# round 2, visit C (8 GPUs): the iteration-based sharded planner (NCCL, one process per GPU) on 65536^2, a few settings.
mkdir -p gpurun_out
: > gpurun_out/bands_n8_iter.jsonl
for cfg in "1 2" "1 1" "2 1" "4 1"; do
  set -- $cfg
  TASI_BANDS_PER_GPU=$1 TASI_BAND_PHASES=$2 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519 \
    bench.py --gpus 8 --workload bands --side 65536 --steps 2 --warmup 1 >> gpurun_out/bands_n8_iter.jsonl 2> gpurun_out/bands_n8_iter.err
  echo "bands=$1 phases=$2 rc=$?"
done
cut -c1-1500 gpurun_out/bands_n8_iter.jsonl; grep -v "^\[W\|^W1\|^\*\*\*\|OMP_NUM" gpurun_out/bands_n8_iter.err | tail -5
