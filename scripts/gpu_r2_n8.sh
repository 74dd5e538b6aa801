#!/bin/bash
# round 2 (N GPUs): the default bench line under torchrun, as the driver's SCALE run launches it
N=${1:-8}
mkdir -p gpurun_out
t0=$(date +%s)
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench n$N rc=$? wall $(( $(date +%s) - t0 )) s"
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_n$N.json").read().strip().splitlines()[-1])
print("value", d["value"], "e2e", d["e2e"]["value"])
print("batch", d["batch"]["queries_per_s"], d["batch"]["verified"])
for k,v in d["bands"].items(): print("bands", k, {x: v.get(x) for x in ("ms","mcells_per_s","verified","bands_per_gpu","busy_frac_min_max_rank","single_gpu_unsharded_ms_rank0","tile_visits_per_tile","skipped")})
PY
tail -2 gpurun_out/bench_n$N.err
