#!/bin/bash
# round 2, visit J (1 GPU): knob sweep for cfg2 (MS-DFM 4096^2): strip width, CTAs per SM, band factor
mkdir -p gpurun_out
out=gpurun_out/dfm4096_knobs.txt; : > $out
run() { echo "== $*" >> $out; env "$@" QUICK_KINDS=1 timeout 300 python scripts/quick_bench.py 4096 2>&1 | python -c "
import json,sys
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    tot=d['cyc_pop']+d['cyc_stage']+d['cyc_relax']+d['cyc_post']
    print('ms %.2f blocks %d v/t %.2f r/v %.1f ev/el %.1f cyc/round %.0f cyc/visit %.0f pop %.2f phases %d' % (d['kernel_ms_solve'], d['solver_blocks'], d['visits_per_tile'], d['rounds_per_visit'], d['evals_per_elem'], d['cyc_relax']/d['tile_rounds'], tot/d['tile_visits'], d['cyc_pop']/tot, d['phases']))
" >> $out; }
run A=0
run TASI_STRIP=2
run TASI_BLOCKS_PER_SM=1
for f in 2 3 5 6; do run TASI_DELTA_FACTOR=$f; done
for f in 3 4 6; do run TASI_DELTA_FACTOR=$f TASI_STRIP=2; done
for f in 3 4 6; do run TASI_DELTA_FACTOR=$f TASI_BLOCKS_PER_SM=1; done
cat $out
