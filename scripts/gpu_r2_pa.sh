#!/bin/bash
# round 2 (1 GPU): pop-ahead: parity + timings
mkdir -p gpurun_out
export TASI_TIMEOUT_S=10
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_full.log 2>&1; echo "full rc=$?"; tail -n 3 gpurun_out/pytest_gpu_full.log
python scripts/quick_bench.py 1024 4096 8192 16384 > gpurun_out/quick.log 2>&1
python - <<'PY'
import json
for l in open("gpurun_out/quick.log"):
    try: d=json.loads(l)
    except Exception: print(l.strip()[:300]); continue
    tot=d['cyc_pop']+d['cyc_stage']+d['cyc_relax']+d['cyc_post']
    print('kind %d N %5d ms %7.2f v/t %5.2f r/v %4.1f ev/el %5.1f cyc/visit %6.0f pop %.2f post %.2f' % (d['kind'], d['N'], d['kernel_ms_solve'], d['visits_per_tile'], d['rounds_per_visit'], d['evals_per_elem'], tot/d['tile_visits'], d['cyc_pop']/tot, d['cyc_post']/tot))
PY
