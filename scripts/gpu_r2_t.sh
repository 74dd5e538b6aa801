#!/bin/bash
# round 2, visit T (1 GPU): bucketed near queue: rings x band factor over the sizes
mkdir -p gpurun_out
out=gpurun_out/buckets_sweep3.txt; : > $out
run() { echo "== $*" >> $out; env "$@" timeout 600 python scripts/quick_bench.py $SIZES 2>&1 | python -c "
import json,sys
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    tot=d['cyc_pop']+d['cyc_stage']+d['cyc_relax']+d['cyc_post']
    print('kind %d N %5d ms %7.2f v/t %5.2f r/v %4.1f ev/el %5.1f cyc/visit %6.0f pop %.2f phases %d' % (d['kind'], d['N'], d['kernel_ms_solve'], d['visits_per_tile'], d['rounds_per_visit'], d['evals_per_elem'], tot/d['tile_visits'], d['cyc_pop']/tot, d['phases']))
" >> $out; }
SIZES="1024 4096 16384"
for nb in 16 30; do for f in 4 8 16 32; do run TASI_BUCKETS=$nb TASI_DELTA_FACTOR=$f; done; done
SIZES="4096"
run TASI_BUCKETS=30 TASI_DELTA_FACTOR=64
run TASI_BUCKETS=30 TASI_DELTA_FACTOR=12 TASI_STRIP=2
run TASI_BUCKETS=30 TASI_DELTA_FACTOR=12 TASI_BLOCKS_PER_SM=1
cat $out
