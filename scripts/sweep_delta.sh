#!/bin/bash
# development probe: sensitivity of the near/far band width
for f in ${FACTORS:-0 0.25 0.5 1 2 4}; do
  echo "== TASI_DELTA_FACTOR=$f"
  TASI_DELTA_FACTOR=$f timeout 300 python scripts/quick_bench.py "$@" | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    print(d['kind'], d['N'], 'ms=%.2f' % d['kernel_ms_solve'], 'Mc/s=%.0f' % d['mcells_s'], 'vis/tile=%.2f' % d['visits_per_tile'], 'rnd/vis=%.2f' % d['rounds_per_visit'], 'ev/el=%.1f' % d['evals_per_elem'], 'low/el=%.1f' % d['lowered_per_elem'], 'phases', d['phases'], 'busy%%=%.0f' % (100*(d['cyc_stage']+d['cyc_relax']+d['cyc_post'])/max(1,d['cyc_pop']+d['cyc_stage']+d['cyc_relax']+d['cyc_post'])), 'blk', d['solver_blocks'], 'rc', d['rc'])
"
done
