#!/bin/bash
# round 2, visit S (1 GPU): bucketed near queue with counted rings: parity, timings
mkdir -p gpurun_out
TASI_BUCKETS=8 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_focus.py -m gpu -x -q > gpurun_out/pytest_buckets.log 2>&1; echo "pytest buckets rc=$?"; tail -n 3 gpurun_out/pytest_buckets.log
out=gpurun_out/buckets_sweep2.txt; : > $out
run() { echo "== $*" >> $out; env "$@" timeout 600 python scripts/quick_bench.py $SIZES 2>&1 | python -c "
import json,sys
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    tot=d['cyc_pop']+d['cyc_stage']+d['cyc_relax']+d['cyc_post']
    print('kind %d N %5d ms %7.2f v/t %5.2f r/v %4.1f ev/el %5.1f cyc/visit %6.0f pop %.2f phases %d' % (d['kind'], d['N'], d['kernel_ms_solve'], d['visits_per_tile'], d['rounds_per_visit'], d['evals_per_elem'], tot/d['tile_visits'], d['cyc_pop']/tot, d['phases']))
" >> $out; }
SIZES="1024 4096 16384"
run TASI_BUCKETS=0
run TASI_BUCKETS=8
run TASI_BUCKETS=16
run TASI_BUCKETS=4
SIZES="4096"
for f in 6 8 12; do run TASI_BUCKETS=16 TASI_DELTA_FACTOR=$f; done
for f in 6 8; do run TASI_BUCKETS=8 TASI_DELTA_FACTOR=$f; done
cat $out
