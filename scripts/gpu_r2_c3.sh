#!/bin/bash
# round 2 (1 GPU): push-style list relaxation (TASI_COMPACT=3) against the scanned list (1) and strips (0)
mkdir -p gpurun_out
export TASI_TIMEOUT_S=10
TASI_COMPACT=3 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_config_sizes.py tests/test_gpu_golden.py tests/test_focus.py -m gpu -x -q 2>&1 | tail -n 2
for c in 1 3; do
TASI_COMPACT=$c python scripts/quick_bench.py 1024 4096 8192 16384 > gpurun_out/quick_c$c.log 2>&1
python - <<PY
import json
for l in open("gpurun_out/quick_c$c.log"):
    try: d=json.loads(l)
    except Exception: print(l.strip()[:300]); continue
    tot=d['cyc_pop']+d['cyc_stage']+d['cyc_relax']+d['cyc_post']
    print('cmp $c kind %d N %5d ms %7.2f v/t %5.2f r/v %4.1f ev/el %5.1f cyc/round %5.0f cyc/visit %6.0f' % (d['kind'], d['N'], d['kernel_ms_solve'], d['visits_per_tile'], d['rounds_per_visit'], d['evals_per_elem'], d['cyc_relax']/d['tile_rounds'], tot/d['tile_visits']))
PY
done
