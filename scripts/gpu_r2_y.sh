#!/bin/bash
# round 2, visit Y (1 GPU): strip width x compaction under the bucketed queue (the solve is throughput-bound now)
mkdir -p gpurun_out
out=gpurun_out/buckets_strip_cmp.txt; : > $out
run() { echo "== $* sizes $SIZES" >> $out; env "$@" timeout 900 python scripts/quick_bench.py $SIZES 2>&1 | python -c "
import json,sys
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    tot=d['cyc_pop']+d['cyc_stage']+d['cyc_relax']+d['cyc_post']
    print('kind %d N %5d ms %7.2f blocks %d v/t %5.2f r/v %4.1f ev/el %5.1f cyc/round %5.0f cyc/visit %6.0f pop %.2f' % (d['kind'], d['N'], d['kernel_ms_solve'], d['solver_blocks'], d['visits_per_tile'], d['rounds_per_visit'], d['evals_per_elem'], d['cyc_relax']/d['tile_rounds'], tot/d['tile_visits'], d['cyc_pop']/tot))
" >> $out; }
SIZES="1024 4096 8192 16384"
for st in 2 4; do for c in 0 1; do run TASI_STRIP=$st TASI_COMPACT=$c; done; done
cat $out
