#!/bin/bash
# round 2 (1 GPU): number of persistent CTAs (MS-DFM 4096^2 and the node planners at 16384^2) under the bucketed queue
run() { echo "== $*"; env "$@" python scripts/quick_bench.py $SIZES 2>&1 | python -c "
import json,sys
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    print('kind %d N %5d ms %7.2f v/t %5.2f ev/el %5.1f' % (d['kind'], d['N'], d['kernel_ms_solve'], d['visits_per_tile'], d['evals_per_elem']))"; }
SIZES=4096
for b in 148 185 222 259 296; do run QUICK_KINDS=1 QUICK_BLOCKS=$b; done
for b in 100 124 148; do run QUICK_KINDS=1 TASI_STRIP=2 QUICK_BLOCKS=$b; done
SIZES=16384
for b in 222 259 296; do run QUICK_KINDS=2,0 QUICK_BLOCKS=$b; done
