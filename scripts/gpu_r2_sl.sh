#!/bin/bash
# round 2 (1 GPU): sliding key window (TASI_SLIDE=1): parity, then timings over window widths
mkdir -p gpurun_out
export TASI_TIMEOUT_S=5
TASI_SLIDE=1 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_slide.log 2>&1; echo "pytest slide rc=$?"; tail -n 5 gpurun_out/pytest_slide.log
out=gpurun_out/slide_sweep.txt; : > $out
run() { echo "== $* sizes $SIZES" >> $out; env "$@" timeout 300 python scripts/quick_bench.py $SIZES 2>&1 | python -c "
import json,sys
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: print(l.strip()[:300]); continue
    tot=d['cyc_pop']+d['cyc_stage']+d['cyc_relax']+d['cyc_post']
    print('kind %d N %5d ms %7.2f v/t %5.2f ev/el %5.1f cyc/visit %6.0f pop %.2f phases %d parked %d' % (d['kind'], d['N'], d['kernel_ms_solve'], d['visits_per_tile'], d['evals_per_elem'], tot/d['tile_visits'], d['cyc_pop']/tot, d['phases'], d['far_parked']))
" >> $out; }
SIZES="1024 4096"
run TASI_SLIDE=0
for f in 4 8 12 16; do run TASI_SLIDE=1 TASI_DELTA_FACTOR=$f; done
cat $out
