#!/bin/bash
# round 2 (1 GPU): upwind deferral (TASI_DEFER = margin in tile widths): parity, then timings
mkdir -p gpurun_out
export TASI_TIMEOUT_S=5
TASI_DEFER=0.5 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_focus.py -m gpu -x -q > gpurun_out/pytest_defer.log 2>&1; echo "pytest defer rc=$?"; tail -n 3 gpurun_out/pytest_defer.log
out=gpurun_out/defer_sweep.txt; : > $out
run() { echo "== $* sizes $SIZES" >> $out; env "$@" timeout 600 python scripts/quick_bench.py $SIZES 2>&1 | python -c "
import json,sys
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    tot=d['cyc_pop']+d['cyc_stage']+d['cyc_relax']+d['cyc_post']
    print('kind %d N %5d ms %7.2f v/t %5.2f r/v %4.1f ev/el %5.1f cyc/visit %6.0f pop %.2f phases %d' % (d['kind'], d['N'], d['kernel_ms_solve'], d['visits_per_tile'], d['rounds_per_visit'], d['evals_per_elem'], tot/d['tile_visits'], d['cyc_pop']/tot, d['phases']))
" >> $out; }
SIZES="1024 4096 16384"
run TASI_DEFER=-1
for m in 0.25 0.5 1 2; do run TASI_DEFER=$m; done
cat $out
