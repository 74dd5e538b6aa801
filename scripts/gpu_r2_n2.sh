#!/bin/bash
# round 2 (2 GPUs): sharded tests + the whole default bench line under torchrun at N = 2 (what the driver's SCALE run does)
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_sharded.py tests/test_bands.py -m gpu -x -q 2>&1 | tail -2
t0=$(date +%s)
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench n2 rc=$? wall $(( $(date +%s) - t0 )) s"
python - <<'PY'
import json
d=json.loads(open("gpurun_out/bench_n2.json").read().strip().splitlines()[-1])
print("value", d["value"], "e2e", d["e2e"]["value"])
print("batch", d["batch"]["queries_per_s"], d["batch"]["verified"])
for k,v in d["bands"].items(): print("bands", k, {x: v.get(x) for x in ("ms","mcells_per_s","verified","bands_per_gpu","busy_frac_min_max_rank","single_gpu_unsharded_ms_rank0","tile_visits_per_tile","skipped")})
PY
tail -3 gpurun_out/bench_n2.err
