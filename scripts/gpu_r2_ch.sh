#!/bin/bash
# round 2 (1 GPU): chaining (TASI_CHAIN=1): parity + timings against off
mkdir -p gpurun_out
export TASI_TIMEOUT_S=10
TASI_CHAIN=1 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_focus.py tests/test_config_sizes.py -m gpu -x -q > gpurun_out/pytest_chain.log 2>&1; echo "chain pytest rc=$?"; tail -n 3 gpurun_out/pytest_chain.log
for c in 0 1; do
TASI_CHAIN=$c python scripts/quick_bench.py 1024 4096 8192 16384 > gpurun_out/quick_chain$c.log 2>&1
python - <<PY
import json
for l in open("gpurun_out/quick_chain$c.log"):
    try: d=json.loads(l)
    except Exception: print(l.strip()[:300]); continue
    tot=d['cyc_pop']+d['cyc_stage']+d['cyc_relax']+d['cyc_post']
    print('chain $c kind %d N %5d ms %7.2f v/t %5.2f r/v %4.1f ev/el %5.1f cyc/visit %6.0f pop %.2f post %.2f pushes/tile %.1f' % (d['kind'], d['N'], d['kernel_ms_solve'], d['visits_per_tile'], d['rounds_per_visit'], d['evals_per_elem'], tot/d['tile_visits'], d['cyc_pop']/tot, d['cyc_post']/tot, d['queue_pushes']/d['tiles']))
PY
done
