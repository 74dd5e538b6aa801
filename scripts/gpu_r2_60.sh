#!/bin/bash
# round 2 (1 GPU): 60 rings against 30 (two rings per lane of the popping warp)
mkdir -p gpurun_out
export TASI_TIMEOUT_S=10
TASI_BUCKETS=60 timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -n 2
out=gpurun_out/buckets_60.txt; : > $out
run() { echo "== $* sizes $SIZES" >> $out; env "$@" timeout 600 python scripts/quick_bench.py $SIZES 2>&1 | python -c "
import json,sys
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    tot=d['cyc_pop']+d['cyc_stage']+d['cyc_relax']+d['cyc_post']
    print('kind %d N %5d ms %7.2f v/t %5.2f ev/el %5.1f cyc/visit %6.0f pop %.2f phases %d' % (d['kind'], d['N'], d['kernel_ms_solve'], d['visits_per_tile'], d['evals_per_elem'], tot/d['tile_visits'], d['cyc_pop']/tot, d['phases']))
" >> $out; }
SIZES="1024 4096 8192 16384"
run TASI_BUCKETS=30
run TASI_BUCKETS=60
run TASI_BUCKETS=45
SIZES="4096"
run TASI_BUCKETS=60 TASI_DELTA_FACTOR=48
run TASI_BUCKETS=60 TASI_DELTA_FACTOR=24
SIZES="16384"
run TASI_BUCKETS=60 TASI_DELTA_FACTOR=20
cat $out
