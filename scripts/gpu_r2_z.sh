#!/bin/bash
# round 2, visit Z (1 GPU): upwind filter of the relaxation: parity (all GPU tests), then timings
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_full.log 2>&1; echo "full rc=$?"; tail -n 6 gpurun_out/pytest_gpu_full.log
python scripts/quick_bench.py 1024 4096 8192 16384 > gpurun_out/quick.log 2>&1
python - <<'PY'
import json
for l in open("gpurun_out/quick.log"):
    try: d=json.loads(l)
    except Exception: print(l.strip()[:300]); continue
    tot=d['cyc_pop']+d['cyc_stage']+d['cyc_relax']+d['cyc_post']
    print('kind %d N %5d ms %7.2f v/t %5.2f r/v %4.1f ev/el %5.1f cyc/round %5.0f cyc/visit %6.0f pop %.2f' % (d['kind'], d['N'], d['kernel_ms_solve'], d['visits_per_tile'], d['rounds_per_visit'], d['evals_per_elem'], d['cyc_relax']/d['tile_rounds'], tot/d['tile_visits'], d['cyc_pop']/tot))
PY
