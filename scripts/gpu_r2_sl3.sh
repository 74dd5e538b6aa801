#!/bin/bash
# round 2 (1 GPU): sliding key window, window width sweep
export TASI_TIMEOUT_S=10
run() { echo "== $* sizes $SIZES"; env "$@" timeout 300 python scripts/quick_bench.py $SIZES 2>&1 | python -c "
import json,sys
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: print(l.strip()[:300]); continue
    tot=d['cyc_pop']+d['cyc_stage']+d['cyc_relax']+d['cyc_post']
    print('kind %d N %5d ms %7.2f v/t %5.2f ev/el %5.1f pop %.2f jumps %d parked %d' % (d['kind'], d['N'], d['kernel_ms_solve'], d['visits_per_tile'], d['evals_per_elem'], d['cyc_pop']/tot, d['phases'], d['far_parked']))
"; }
SIZES="4096"
for f in 24 32 48; do run TASI_SLIDE=1 TASI_DELTA_FACTOR=$f QUICK_KINDS=1; done
SIZES="16384"
run TASI_SLIDE=0
for f in 6 12 24; do run TASI_SLIDE=1 TASI_DELTA_FACTOR=$f; done
SIZES="1024"
run TASI_SLIDE=1 TASI_DELTA_FACTOR=32
