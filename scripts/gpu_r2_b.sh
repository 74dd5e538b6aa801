#!/bin/bash
# round 2, visit B (2 GPUs): the sharded planner on real devices — one process over both (peer reads, then NCCL), then
# the bench line under torchrun with its bands record.
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv > gpurun_out/gpu2.txt 2>&1
timeout 600 python scripts/shard_probe.py 16384 0 2 > gpurun_out/shard_probe_p2p_n2.jsonl 2> gpurun_out/shard_probe_p2p_n2.err; echo "probe p2p rc=$?"
TASI_SHARD_NCCL=1 PROBE_SETTINGS=4:2 timeout 600 python scripts/shard_probe.py 16384 0 2 > gpurun_out/shard_probe_nccl_n2.jsonl 2> gpurun_out/shard_probe_nccl_n2.err; echo "probe nccl rc=$?"
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench n2 rc=$?"
cat gpurun_out/shard_probe_p2p_n2.jsonl gpurun_out/shard_probe_nccl_n2.jsonl; tail -3 gpurun_out/shard_probe_p2p_n2.err gpurun_out/shard_probe_nccl_n2.err; cut -c1-6000 gpurun_out/bench_n2.json; tail -5 gpurun_out/bench_n2.err
