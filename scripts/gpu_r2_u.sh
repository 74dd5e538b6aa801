#!/bin/bash
# round 2, visit U (1 GPU): bucketed near queue, band factor at 8192^2 (all planners) and 32768^2 / 65536^2 (FD*)
mkdir -p gpurun_out
out=gpurun_out/buckets_sweep4.txt; : > $out
run() { echo "== $* sizes $SIZES" >> $out; env "$@" timeout 900 python scripts/quick_bench.py $SIZES 2>&1 | python -c "
import json,sys
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    tot=d['cyc_pop']+d['cyc_stage']+d['cyc_relax']+d['cyc_post']
    print('kind %d N %5d ms %7.2f v/t %5.2f r/v %4.1f ev/el %5.1f cyc/visit %6.0f pop %.2f phases %d' % (d['kind'], d['N'], d['kernel_ms_solve'], d['visits_per_tile'], d['rounds_per_visit'], d['evals_per_elem'], tot/d['tile_visits'], d['cyc_pop']/tot, d['phases']))
" >> $out; }
SIZES="8192"
run TASI_BUCKETS=0
for f in 8 16 24 32; do run TASI_BUCKETS=30 TASI_DELTA_FACTOR=$f; done
SIZES="32768"
run TASI_BUCKETS=0 QUICK_KINDS=0
for f in 4 8 12; do run TASI_BUCKETS=30 TASI_DELTA_FACTOR=$f QUICK_KINDS=0; done
SIZES="65536"
run TASI_BUCKETS=0 QUICK_KINDS=0
for f in 4 8; do run TASI_BUCKETS=30 TASI_DELTA_FACTOR=$f QUICK_KINDS=0; done
cat $out
